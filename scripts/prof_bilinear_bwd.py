"""Standalone timing of the bilinear (align_corners) backward / forward on the decoder's largest shapes (ncu target)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import ops, train_ops as T

dev = torch.device("cuda:0")
B = 16
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timed(fn, reps=5):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:])

for (H, ctot, coff, C) in [(256, 160, 0, 64), (256, 32, 0, 32), (128, 320, 0, 128), (256, 64, 0, 64)]:
    dcat = ops.Act(torch.randn(B, H, H, ctot, device=dev).bfloat16())
    dy = dcat.slice(coff, C)
    t = timed(lambda: T.bilinear_ac_bwd(dy, H // 2, H // 2))
    mb = (B * H * H * C * 2 + B * (H // 2) ** 2 * C * 2) / 1e6
    print(f"bilinear_bwd {H}->{H//2} C{C} of ctot {ctot}: {t*1e3:.1f} us, {mb:.0f} MB algorithmic, {mb / t / 1e3:.2f} TB/s")
    x = ops.Act(torch.randn(B, H // 2, H // 2, C, device=dev).bfloat16())
    out = ops.Act(torch.empty(B, H, H, ctot, device=dev, dtype=torch.bfloat16)).slice(coff, C)
    t = timed(lambda: ops.bilinear_ac(x, H, H, out=out))
    print(f"bilinear_fwd {H//2}->{H} C{C} into ctot {ctot}: {t*1e3:.1f} us, {mb / t / 1e3:.2f} TB/s")
