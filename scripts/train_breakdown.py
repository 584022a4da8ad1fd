"""Where the training step's time goes (B=16, 512x512, one B200), by CUDA events on the current stream:
forward only, forward + loss + backward, whole step; run with VU_DEBUG_SKIP_WGRAD=1 / VU_TRAIN_STREAMS=0 for the variants.
usage: python scripts/train_breakdown.py [B] [S]"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
S = int(sys.argv[2]) if len(sys.argv) > 2 else 512
dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
ts = TrainStep(model)
x, m = synthetic_fundus(2, S, S, seed=5)
x = x.repeat(B // 2, 1, 1, 1).to(dev); m = m.repeat(B // 2, 1, 1, 1).to(dev)
eps = torch.randn(B, 32, device=dev)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def fwd():
    ts.repack()
    return ts.forward(x, eps)


def fwd_bwd():
    ts.zero_grad()
    ts.forward_backward(x, m, 0.001, eps)
    ts._join_side()


import time
t_f = timed(fwd)
t_fb = timed(fwd_bwd)
t_s = timed(lambda: ts.step(x, m, 0.001, eps))
t0 = time.perf_counter()
for _ in range(5):
    ts.step(x, m, 0.001, eps)
t_host = (time.perf_counter() - t0) / 5
torch.cuda.synchronize()
print(f"env: VU_TRAIN_STREAMS={os.environ.get('VU_TRAIN_STREAMS', '1')} VU_DEBUG_SKIP_WGRAD={os.environ.get('VU_DEBUG_SKIP_WGRAD', '0')}")
print(f"B={B} S={S}: repack+forward {t_f:.3f} ms | zero_grad+forward+loss+backward {t_fb:.3f} ms | whole step {t_s:.3f} ms | "
      f"host enqueue per step {t_host * 1e3:.3f} ms")
