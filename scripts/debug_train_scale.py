"""Where does the TrainStep forward diverge from the oracle as the shape grows? (debug aid)
usage: python scripts/debug_train_scale.py B S dtype"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from oracle import vaeunet_oracle as O
from vaeunet_b200 import ops, testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet

B, S = int(sys.argv[1]), int(sys.argv[2])
dtn = sys.argv[3] if len(sys.argv) > 3 else "f16"
dt = ops.F16 if dtn == "f16" else ops.BF16
dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
x, m = synthetic_fundus(B, S, S, seed=161)
eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(162))
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
ts = TrainStep(model, lesion="EX", dtype=dt, loss_scale=1024.0 if dtn == "f16" else 1.0)
ts.repack()
logits, mu, lv, St = ts.forward(x.to(dev), eps.to(dev))
torch.cuda.synchronize()
st = O.BNState(training=True)
with torch.no_grad():
    feats = O.resnet34_features(sd, x, "encoder.", st)
    mu_r, lv_r = O.latent_heads(sd, feats[-1])
    z = mu_r + eps * torch.exp(0.5 * lv_r)
    out = O.decode_from_features(sd, O.ResNetCfg(), feats, z, True, st)
    out = O._up_ac(out, x.shape[2:])
rel = lambda a, b: ((a.float().cpu() - b).norm() / (b.norm() + 1e-12)).item()
print(f"B={B} S={S} {dtn}")
for i, (fa, fr) in enumerate(zip(St["feats"], feats)):
    g = fa.t[..., fa.coff:fa.coff + fa.C].float().permute(0, 3, 1, 2)
    print(f"  f{i}: rel {rel(g, fr):.3e}  max|gpu| {g.abs().max().item():.1f} max|ref| {fr.abs().max().item():.1f}")
print(f"  mu rel {rel(mu, mu_r):.3e} logvar rel {rel(lv, lv_r):.3e}")
for i, r in enumerate(St["dec"]):
    y2 = r["y2"]
    print(f"  dec{i}.y2 max {y2.t.float().abs().max().item():.1f} finite {bool(torch.isfinite(y2.t.float()).all())}")
print(f"  logits rel {rel(logits, out):.3e}  max|gpu| {logits.abs().max().item():.2f} max|ref| {out.abs().max().item():.2f}")
loss_r = O.combined_loss(out, m).item()
sums = ops.seg_loss_sums(logits.reshape(-1).contiguous(), m.to(dev).reshape(-1).contiguous(), False, 0.8, 2.0)
print(f"  oracle recon {loss_r:.5f}")
out2 = ts.loss_and_grads(St, m.to(dev), 0.001)[0]
print(f"  gpu loss {out2['loss'].item():.5f} recon {out2['recon'].item():.5f} kl {out2['kl'].item():.5f}")
