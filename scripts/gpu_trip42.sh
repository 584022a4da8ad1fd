#!/bin/bash
# round-1d evidence: full GPU test suite, both bench arms, ncu launch lists of the two workloads
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -4 > gpurun_out/r1d_pytest.txt
cat gpurun_out/r1d_pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 2> gpurun_out/r1d_bench.err | tail -1 > gpurun_out/r1d_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 2> gpurun_out/r1d_ref.err | tail -1 > gpurun_out/r1d_bench_reference.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1d_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/r1d_ncu_mc.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1d_train_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload train > gpurun_out/r1d_ncu_train.log 2>&1
python scripts/mc_launch_summary.py gpurun_out/r1d_mc_launches.csv > gpurun_out/r1d_mc_launch_summary.txt 2>&1
python scripts/summarize_launches.py gpurun_out/r1d_train_launches.csv > gpurun_out/r1d_train_launch_summary.txt 2>&1
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r1d_bench.json').read())
print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e'], 'roofline', d['roofline'], 'cpu', d['cpu_baseline'])
t = d['train']; print('train', t['value'], t['ms_per_step'], 'e2e', t['e2e'], t.get('roofline'), t.get('cpu_baseline'))
print(open('gpurun_out/r1d_bench_reference.json').read()[:600])
PY
