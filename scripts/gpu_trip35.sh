#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc --layers gpurun_out/trip35_layers.csv > /dev/null 2>&1
head -16 gpurun_out/trip35_layers.csv
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/trip35_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/trip35_ncu.log 2>&1
python - <<'PY'
import csv, re
from collections import defaultdict
lines = [l for l in open('gpurun_out/trip35_mc_launches.csv') if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
names = [re.sub(r'\(.*', '', r['Kernel Name']) for r in rows]
starts = [i for i, n in enumerate(names) if 'small_cin_conv_kernel<7' in n]
a, b = starts[2], starts[3]
for r, n in zip(rows[a:b], names[a:b]):
    us = float(r['Metric Value'].replace(',', '')) / 1e3
    if us > 100: print(f"{us:9.1f} us {n}")
PY
