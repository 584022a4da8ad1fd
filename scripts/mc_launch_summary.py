"""Per-step kernel summary of an MC bench run captured with `ncu --metrics gpu__time_duration.sum --csv` (one step = the
launches between two stem kernels). usage: python scripts/mc_launch_summary.py launches.csv [--list]"""
import csv, re, sys
from collections import defaultdict
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
names = [re.sub(r'\(.*', '', r['Kernel Name']) for r in rows]
starts = [i for i, n in enumerate(names) if 'small_cin_conv_kernel<7' in n or 'small_cin_mma_kernel<7' in n]
a, b = starts[2], starts[3]
agg = defaultdict(lambda: [0, 0.0]); tot = 0
for r, n in zip(rows[a:b], names[a:b]):
    us = float(r['Metric Value'].replace(',', '')) / 1e3
    agg[n][0] += 1; agg[n][1] += us; tot += us
    if '--list' in sys.argv: print(f"{us:9.1f} us {n}")
print(f"# one MC step from the ncu launch list (cold-cache, serialised times): {b-a} launches, {tot/1e3:.3f} ms")
for name, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{us/1e3:8.3f} ms {100*us/tot:5.1f}% x{c:3d} {name}")
