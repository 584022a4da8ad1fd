import torch
dev = torch.device("cuda:0")
n = 1 << 30
a = torch.empty(n, dtype=torch.float16, device=dev); b = torch.empty(n, dtype=torch.float16, device=dev)
def t(f, reps=5):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: a.fill_(1.0)); print(f"fill 2 GiB: {ms:.3f} ms  {2 * n / ms / 1e6:.0f} GB/s write-only")
ms = t(lambda: b.copy_(a)); print(f"copy 2 GiB -> 2 GiB: {ms:.3f} ms  {4 * n / ms / 1e6:.0f} GB/s read+write")
ms = t(lambda: a.sum()); print(f"sum 2 GiB: {ms:.3f} ms  {2 * n / ms / 1e6:.0f} GB/s read-only")
