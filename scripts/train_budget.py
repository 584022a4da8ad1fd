"""How much of the gradient error is the 16-bit storage format itself? Runs the CPU oracle's training_grads with
activation/filter storage rounding (fp16 / bf16) and compares with the fp32 oracle, tensor by tensor. CPU only."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from oracle import vaeunet_oracle as O
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus

B, S = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2, 64)
sd, kw = testing.resnet_fixture()
x, m = synthetic_fundus(B, S, S, seed=21, lesion="EX")
eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(22))
loss0, _, kl0, g0, _ = O.training_grads(sd, O.ResNetCfg(), x, m, eps, 0.001, 1e-3, "EX")
for dt in (torch.float16, torch.bfloat16):
    with O.storage_rounding(dt):
        loss1, _, kl1, g1, _ = O.training_grads(sd, O.ResNetCfg(), x, m, eps, 0.001, 1e-3, "EX")
    f0 = torch.cat([g.flatten() for g in g0.values()]); f1 = torch.cat([g1[k].flatten() for k in g0])
    print(f"== {dt} B={B} S={S}: loss {loss1.item():.6f} vs {loss0.item():.6f}; cosine {(f0 @ f1 / f0.norm() / f1.norm()).item():.5f} "
          f"total_rel {((f1 - f0).norm() / f0.norm()).item():.4f}")
    for k in g0:
        if "decoder_blocks.3" in k or k.startswith("final") or "layer1.0" in k or k.startswith("encoder.conv1") or "z_initial" in k:
            print(f"   {((g1[k] - g0[k]).norm() / (g0[k].norm() + 1e-20)).item():9.4f} |g|={g0[k].norm().item():.3e} {k}")
