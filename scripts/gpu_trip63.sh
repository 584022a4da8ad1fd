#!/bin/bash
# round-1i: final state (fixed-order latent head on top of r1h): tests + both bench arms + smoke
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -3 > gpurun_out/r1i_pytest.txt
cat gpurun_out/r1i_pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 2> gpurun_out/r1i_bench.err | tail -1 > gpurun_out/r1i_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 2> gpurun_out/r1i_ref.err | tail -1 > gpurun_out/r1i_bench_reference.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r1i_bench.json').read())
print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], d['roofline']['whole_step']['frac_of_sustained'], 'cpu', d['cpu_baseline']['value'], d['clocks'], d['gpu_launches'])
t = d['train']; print('train', t['value'], t['ms_per_step'], 'e2e', t['e2e']['value'], t['roofline']['frac'], t['roofline']['whole_step']['frac_of_sustained'], t['cpu_baseline']['value'], t['gpu_launches'])
PY
