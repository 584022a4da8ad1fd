"""Per-tensor cosine between the gradients of the fused TrainStep and of the autograd path (same kernels, same 16-bit storage,
different evaluation orders / fusions): the wiring pin of tests/test_autograd_gpu.py::test_module_train_matches_trainstep."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import torch
import test_autograd_gpu as T
from vaeunet_b200 import ops, testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet
dev = torch.device("cuda:0")
sd, mkw = testing.resnet_fixture()
for B, S in ((8, 64), (8, 128)):
    x, m = synthetic_fundus(B, S, S, seed=41)
    eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(42))
    model = UNetResNet(**mkw); model.load_state_dict(sd); model = model.to(dev); model.compute_dtype = "f16"
    T._reference_step_body(model, x.to(dev), m.to(dev), eps.to(dev), 0.001, "EX")
    ga = {k: p.grad.detach().clone() / T.SCALE for k, p in model.named_parameters()}
    twin = UNetResNet(**mkw); twin.load_state_dict(sd); twin = twin.to(dev).train()
    ts = TrainStep(twin, lesion="EX", dtype=ops.F16, loss_scale=T.SCALE)
    ts.zero_grad(); ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev)); torch.cuda.synchronize()
    rows = []
    for k in ga:
        if k.endswith('.0.bias') and ('attention' in k or 'z_proj' in k or 'z_initial' in k):
            continue                       # a conv bias in front of a BatchNorm: the true gradient is exactly zero
        a, b = ga[k].flatten().float(), (ts.G[k] / T.SCALE).flatten().float()
        c = (a @ b / (a.norm() * b.norm() + 1e-30)).item()
        r = ((a - b).norm() / (a.norm() + 1e-30)).item()
        rows.append((c, r, a.numel(), k))
    rows.sort()
    print(f"B={B} S={S}: {len(rows)} tensors; worst cosines:")
    for c, r, n, k in rows[:12]:
        print(f"  cos {c:.4f} rel {r:.3f} n={n:7d} {k}")
    big = [x for x in rows if x[2] >= 64]
    print("  min cos over tensors with >= 64 elements:", min(big)[0], " median:", sorted(c for c, *_ in big)[len(big) // 2])
