"""Summarise an `ncu --set full` report (.ncu-rep) into the handful of metrics DESIGN.md / bench.py cite.
usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.summary.txt"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_uniform.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors.avg.per_cycle_elapsed", "lts__t_sectors.max.pct_of_peak_sustained_elapsed",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__warps_active.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "gpc__cycles_elapsed.avg.per_second",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {rep}: key metrics per captured launch (ncu --set full --clock-control none; times are cold-cache, serialised)")
    for vals in rows[2:]:
        name = vals[hdr.index("Kernel Name")]
        print(f"\nkernel: {name[:110]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:72s} {vals[i]:>18s} {units[i]}")
        if "dram__bytes_read.sum" in hdr:
            i, j = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            print(f"  traffic = dram read + write: {vals[i]} {units[i]} + {vals[j]} {units[j]}")


if __name__ == "__main__":
    main()
