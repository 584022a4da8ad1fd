"""One training step (B=16, 512x512) between cudaProfilerStart/Stop, for `ncu --profile-from-start off`."""
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
for _ in range(2):
    ts.step(x, m, 0.001, eps)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
ts.step(x, m, 0.001, eps)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("done")
