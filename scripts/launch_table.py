"""Per-launch table (in launch order) of an ncu CSV launch list taken with several --metrics:
time, tensor-pipe %, DRAM read/write bytes, L2 bytes, L2->SM bytes, grid, kernel. usage: launch_table.py x.csv [> table.txt]"""
import csv, re, sys
from collections import OrderedDict, defaultdict

with open(sys.argv[1]) as fh:
    lines = [l for l in fh if not l.startswith("==")]
rows = OrderedDict()
for r in csv.DictReader(lines):
    d = rows.setdefault(r["ID"], {"name": re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("vu::", ""),
                                  "grid": r["Grid Size"]})
    v = float(r["Metric Value"].replace(",", "")) if r["Metric Value"] not in ("", "n/a") else 0.0
    u = r.get("Metric Unit", "")
    scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3,
             "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
    d[r["Metric Name"]] = v * scale
T, P = "gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
print(f"{'#':>4} {'us':>8} {'tens%':>6} {'dramR MB':>9} {'dramW MB':>9} {'L2 MB':>8} {'L2>SM MB':>9} {'GB/s':>7}  grid  kernel")
agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
tot = 0.0
for i, d in enumerate(rows.values()):
    us = d.get(T, 0.0)
    dr, dw = d.get("dram__bytes_read.sum", 0.0), d.get("dram__bytes_write.sum", 0.0)
    gbs = (dr + dw) / us / 1e3 if us > 0 else 0.0
    print(f"{i:4d} {us:8.1f} {d.get(P, 0.0):6.1f} {dr / 1e6:9.2f} {dw / 1e6:9.2f} {d.get('lts__t_bytes.sum', 0.0) / 1e6:8.1f} "
          f"{d.get('l1tex__m_xbar2l1tex_read_bytes.sum', 0.0) / 1e6:9.1f} {gbs:7.0f}  {d['grid']:>14s} {d['name'][:70]}")
    a = agg[d["name"][:70]]
    a[0] += 1; a[1] += us; a[2] += d.get(P, 0.0) * us; a[3] += dr + dw
    tot += us
print(f"\ntotal {tot / 1e3:.3f} ms over {len(rows)} launches; per kernel: ms, share, launches, time-weighted tensor %, DRAM GB/s")
for name, (n, us, pw, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{us / 1e3:9.3f} ms {100 * us / tot:5.1f}%  x{n:4d}  tens {pw / us if us else 0:5.1f}%  {by / us / 1e3 if us else 0:6.0f} GB/s  {name}")
