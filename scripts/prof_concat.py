"""Time vu_decoder_concat at a decoder-block shape. usage: prof_concat.py D rep hs H Cx Cs F [reps] [Cz]
(Cz = 0 with Cx == Cs takes the fp16 fast kernel; VU_NO_FAST_CONCAT=1 forces the generic one)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import ops

D, rep, hs, H, Cx, Cs, F = map(int, sys.argv[1:8])
reps = int(sys.argv[8]) if len(sys.argv) > 8 else 4
Cz = int(sys.argv[9]) if len(sys.argv) > 9 else 32
dev = torch.device("cuda:0")
dt = torch.float16
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s: ops.Act(torch.randn(*s, device=dev, generator=g).to(dt))
x, g1 = r(D, hs, hs, Cx), r(D, hs, hs, F)
x1, skip = r(D // rep, H, H, F), r(D // rep, H, H, Cs)
pw = torch.randn(F, device=dev, generator=g)
zv = torch.randn(D, Cz, device=dev, generator=g) if Cz else None
out = ops.new_act(D, H, H, Cx + Cs + Cz, ops.F16, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for i in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.decoder_concat(x, g1, x1, skip, pw, 0.5, 0.1, zv, out)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
wbytes = out.t.numel() * 2
rbytes = (x.t.numel() + g1.t.numel() + x1.t.numel() + skip.t.numel()) * 2
best = min(ts[1:])
print(f"decoder_concat D{D} {hs}->{H} Cx{Cx} Cs{Cs} F{F}: best {best:.4f} ms, {(wbytes + rbytes) / best / 1e6:.0f} GB/s algorithmic "
      f"({wbytes / 1e6:.0f} MB written, {rbytes / 1e6:.0f} MB read); all {['%.4f' % t for t in ts]}")
