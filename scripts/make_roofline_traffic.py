"""profiles/roofline_traffic.json from the committed ncu launch lists: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum)
of the tcgen05 conv launches of ONE step - one MC micro-batch of 16 patches / one training step of 16 images - i.e. the same
launches whose FLOPs and CUDA-event times make `roofline.achieved` in bench.py. Cold-cache figures (ncu flushes L2 between
launches), so they bound the live traffic from above.
usage: make_roofline_traffic.py <mc_launches.csv> <train_launches.csv> [top_kernel_full_summary.txt ...]"""
import csv, json, re, sys
from collections import OrderedDict
from pathlib import Path

CONV = ("conv_igemm_kernel", "conv_igemm_pair_kernel", "conv3x3_halo_kernel", "conv_wgrad_kernel", "conv_wgrad_row3_kernel")


def launches(path):
    with open(path) as fh:
        lines = [l for l in fh if not l.startswith("==")]
    rows = OrderedDict()
    for r in csv.DictReader(lines):
        d = rows.setdefault(r["ID"], {"name": re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("vu::", "")})
        v = float(r["Metric Value"].replace(",", "")) if r["Metric Value"] not in ("", "n/a") else 0.0
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0,
                 "ms": 1e3, "msecond": 1e3}.get(r.get("Metric Unit", ""), 1.0)
        d[r["Metric Name"]] = v * scale
    return list(rows.values())


def conv_sum(rows):
    sel = [d for d in rows if d["name"].startswith(CONV)]
    by = sum(d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0) for d in sel)
    us = sum(d.get("gpu__time_duration.sum", 0.0) for d in sel)
    top = max(sel, key=lambda d: d.get("gpu__time_duration.sum", 0.0))
    return {"dram_bytes_per_launch": by, "what": "sum over the tcgen05 conv launches of one step (cold-cache ncu launch list)",
            "conv_launches": len(sel), "conv_us_serialised": us,
            "top_launch": {"kernel": top["name"], "us": top.get("gpu__time_duration.sum"),
                           "dram_bytes": top.get("dram__bytes_read.sum", 0.0) + top.get("dram__bytes_write.sum", 0.0)}}


mc_rows = launches(sys.argv[1])
stems = [i for i, d in enumerate(mc_rows) if d["name"].startswith(("small_cin_conv_kernel<7", "small_cin_mma_kernel<7"))]
a, b = (stems[-2], stems[-1]) if len(stems) > 1 else (stems[0], len(mc_rows))   # one micro-batch = stem launch to stem launch
out = {"mc": conv_sum(mc_rows[a:b]), "train": conv_sum(launches(sys.argv[2]))}
out["mc"]["source"], out["train"]["source"] = Path(sys.argv[1]).name, Path(sys.argv[2]).name
Path(__file__).resolve().parent.parent.joinpath("profiles", "roofline_traffic.json").write_text(json.dumps(out, indent=1))
print(json.dumps(out, indent=1))
