"""Run one wgrad shape through the C ABI a few times (target of the ncu captures).
usage: python scripts/prof_wgrad.py N H W Cin Cout k [reps]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import ops

N, H, W, Cin, Cout, k = map(int, sys.argv[1:7])
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 5
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
x = ops.Act(torch.randn(N, H, W, Cin, device=dev, generator=g).bfloat16())
dy = ops.Act(torch.randn(N, H, W, Cout, device=dev, generator=g).bfloat16())
dw = torch.zeros(Cout, k * k, Cin, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for i in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.conv2d_wgrad(x, dy, k, 1, k // 2, out=dw)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
fl = 2.0 * N * H * W * Cout * k * k * Cin
best = min(ts[1:]) if len(ts) > 1 else ts[0]
print(f"wgrad N{N} {H}x{W} Cin{Cin} Cout{Cout} k{k}: best {best:.4f} ms, {fl / best / 1e9:.1f} TFLOP/s; all {['%.4f' % t for t in ts]}")
