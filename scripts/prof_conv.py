"""Run one conv layer shape of the bench through the C ABI a few times (target of the ncu captures).
usage: python scripts/prof_conv.py N H W Cin Cout k [reps] [dtype] [stats]   (stats: training epilogue, BN statistics on)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import ops

N, H, W, Cin, Cout, k = map(int, sys.argv[1:7])
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 5
dt = ops.F16 if (len(sys.argv) <= 8 or sys.argv[8] == "f16") else ops.BF16
tdt = torch.float16 if dt == ops.F16 else torch.bfloat16
stats = len(sys.argv) > 9 and sys.argv[9] == "stats"
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
x = ops.Act((torch.randn(N, H, W, Cin, device=dev, generator=g)).to(tdt))
w = torch.randn(Cout, Cin, k, k, device=dev, generator=g) * (2.0 / (Cin * k * k)) ** 0.5
wp = ops.pack_conv_weight(w, dt)
scale = torch.ones(Cout, device=dev); bias = torch.zeros(Cout, device=dev)
y = ops.new_act(N, H, W, Cout, dt, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for i in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if stats:
        ss = torch.zeros(2 * Cout, dtype=torch.float64, device=dev)
        e0.record()
        ops.conv2d(x, wp, k, 1, k // 2, None, None, False, out=y, stat_sums=ss)
    else:
        ops.conv2d(x, wp, k, 1, k // 2, scale, bias, True, out=y)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
fl = 2.0 * N * H * W * Cout * k * k * Cin
best = min(ts[1:]) if len(ts) > 1 else ts[0]
print(f"conv N{N} {H}x{W} Cin{Cin} Cout{Cout} k{k}: best {best:.4f} ms, {fl / best / 1e9:.1f} TFLOP/s; all {['%.4f' % t for t in ts]}")
