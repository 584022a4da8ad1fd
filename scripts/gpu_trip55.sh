#!/bin/bash
# round-1g: CTA pairs in the tap-reuse kernel too - full tests, both bench arms, MC launch list, one full capture
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -3 > gpurun_out/r1g_pytest.txt
cat gpurun_out/r1g_pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 2> gpurun_out/r1g_bench.err | tail -1 > gpurun_out/r1g_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 2> gpurun_out/r1g_ref.err | tail -1 > gpurun_out/r1g_bench_reference.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1g_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/r1g_ncu_mc.log 2>&1
python scripts/mc_launch_summary.py gpurun_out/r1g_mc_launches.csv > gpurun_out/r1g_mc_launch_summary.txt 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv3x3_halo -c 2 -o gpurun_out/r1g_halo_pair_192to64_256sq \
   python scripts/prof_conv.py 80 256 256 192 64 3 3 f16 > gpurun_out/r1g_ncu_halo.log 2>&1
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r1g_bench.json').read())
print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], d['roofline']['whole_step']['frac_of_sustained'], 'cpu', d['cpu_baseline']['value'], d['clocks'], d['gpu_launches'])
t = d['train']; print('train', t['value'], t['ms_per_step'], 'e2e', t['e2e']['value'], t['roofline']['frac'], t['roofline']['whole_step']['frac_of_sustained'], t['cpu_baseline']['value'], t['gpu_launches'])
PY
cat gpurun_out/r1g_mc_launch_summary.txt | head -12
