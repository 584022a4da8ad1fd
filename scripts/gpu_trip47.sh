#!/bin/bash
# round-1e evidence (CTA-pair conv, fused BN statistics): tests, both bench arms, launch lists, one full capture of the pair kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -4 > gpurun_out/r1e_pytest.txt
cat gpurun_out/r1e_pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 2> gpurun_out/r1e_bench.err | tail -1 > gpurun_out/r1e_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 2> gpurun_out/r1e_ref.err | tail -1 > gpurun_out/r1e_bench_reference.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1e_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/r1e_ncu_mc.log 2>&1
python scripts/mc_launch_summary.py gpurun_out/r1e_mc_launches.csv > gpurun_out/r1e_mc_launch_summary.txt 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_igemm_pair -c 2 -o gpurun_out/r1e_pair_800to512_32sq \
   python scripts/prof_conv.py 80 32 32 800 512 3 3 bf16 > gpurun_out/r1e_ncu_pair.log 2>&1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1e_train_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload train > gpurun_out/r1e_ncu_train.log 2>&1
python scripts/summarize_launches.py gpurun_out/r1e_train_launches.csv > gpurun_out/r1e_train_launch_summary.txt 2>&1
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r1e_bench.json').read())
print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], d['roofline']['whole_step'], 'cpu', d['cpu_baseline']['value'], d['clocks'])
t = d['train']; print('train', t['value'], t['ms_per_step'], 'e2e', t['e2e']['value'], t['roofline']['frac'], t['roofline']['whole_step'], t['cpu_baseline']['value'], t['gpu_launches'])
print(open('gpurun_out/r1e_bench_reference.json').read()[:300])
PY
