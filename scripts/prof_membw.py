"""Write-only / read-only / copy bandwidth of plain torch kernels at a size far above L2: the rooflines of the
write-dominated producers (decoder concat: 2.7 GB written, 0.7 GB read per launch)."""
import torch
dev = torch.device("cuda:0")
n = 2684 << 20
a = torch.empty(n, dtype=torch.uint8, device=dev)
b = torch.empty(n, dtype=torch.uint8, device=dev)
def t(fn, reps=5):
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:])
ms = t(lambda: a.zero_()); print(f"memset {n/1e6:.0f} MB: {ms:.3f} ms = {n/ms/1e6:.0f} GB/s written")
ms = t(lambda: b.copy_(a)); print(f"copy   {n/1e6:.0f} MB: {ms:.3f} ms = {2*n/ms/1e6:.0f} GB/s read+written")
af = a.view(torch.float16)
ms = t(lambda: af.sum()); print(f"read (sum f16) {n/1e6:.0f} MB: {ms:.3f} ms = {n/ms/1e6:.0f} GB/s read")
bf = b.view(torch.float16)
ms = t(lambda: torch.mul(af, 0.5, out=bf)); print(f"scale  {n/1e6:.0f} MB: {ms:.3f} ms = {2*n/ms/1e6:.0f} GB/s read+written")
