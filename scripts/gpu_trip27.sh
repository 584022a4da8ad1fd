#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --layers gpurun_out/trip27_layers.csv > gpurun_out/trip27_bench.json 2> gpurun_out/trip27_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/trip27_mc_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/trip27_ncu.log 2>&1
cat gpurun_out/trip27_bench.json | cut -c1-300
python - <<'PY'
import csv, re
from collections import defaultdict
lines = [l for l in open('gpurun_out/trip27_mc_launches.csv') if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
# bench with warmup 3, steps 1: timed(False)=4 steps, timed(True)=4 steps, conv timing 2 steps => 10 steps
n = len(rows); per = n // 10
step = rows[3 * per:4 * per]
agg = defaultdict(lambda: [0, 0.0]); tot = 0
for r in step:
    us = float(r['Metric Value'].replace(',', '')) / 1e3
    name = re.sub(r'\(.*', '', r['Kernel Name']); agg[name][0] += 1; agg[name][1] += us; tot += us
print(f"launches {n} per step {per}; step total {tot/1e3:.3f} ms")
for name, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{us/1e3:8.3f} ms {100*us/tot:5.1f}% x{c:3d} {name}")
PY
head -30 gpurun_out/trip27_layers.csv
