"""Generate the committed golden fixtures by running the UNMODIFIED reference (/root/reference).

Run in the build container only:   python tests/golden/make_golden.py
Writes (small files, committed):
  fixture_resnet.pt / fixture_resnet_<mode>.pt / fixture_unet*.pt
        calibrated BatchNorm running stats + rescaled heads/output conv that overlay the seeded
        init of vaeunet_b200.testing.seeded_state_dict (the 30 M conv weights are regenerated from
        the seed, never stored)
  golden_resnet.pt, golden_unet.pt, golden_misc.pt
        inputs (seeds), injected eps and the reference's outputs for every hot-path entry point.
"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from unet.unet_resnet import UNetResNet as RefUNetResNet  # noqa: E402  (reference)
from unet.unet_model import UNet as RefUNet  # noqa: E402
from utils import vae_utils as ref_vae  # noqa: E402
from utils import loss as ref_loss  # noqa: E402
from utils import uncertainty_metrics as ref_um  # noqa: E402
from utils import metrics as ref_metrics  # noqa: E402
import visualize_vae as ref_vis  # noqa: E402

from vaeunet_b200.synthetic import synthetic_fundus  # noqa: E402
from vaeunet_b200.testing import seeded_state_dict  # noqa: E402

OUT = Path(__file__).resolve().parent
torch.set_num_threads(8)
L = 32
TRAIN_STEPS = 400


def calibrate(model, run_train_forward):
    """Exact BN calibration: cumulative-average mode, one train-mode pass => running stats = batch stats."""
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.reset_running_stats()
            m.momentum = None
    model.train()
    with torch.no_grad():
        run_train_forward()
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.momentum = 0.1
            m.num_batches_tracked.zero_()
    model.eval()


def small_overlay(sd, extra_keys):
    keep = {}
    for k, v in sd.items():
        if k.endswith("running_mean") or k.endswith("running_var") or k in extra_keys:
            keep[k] = v.detach().clone()
    return keep


def train_last_block(model, steps, log_every=50):
    torch.manual_seed(4242)
    for p in model.parameters():
        p.requires_grad_(False)
    params = list(model.decoder_blocks[3].parameters()) + list(model.final_conv.parameters())
    for p in params:
        p.requires_grad_(True)
    opt = torch.optim.Adam(params, lr=2e-3)
    crit = ref_loss.CombinedLoss()
    model.eval()                                      # BN uses the calibrated running stats throughout
    sizes = [(6, 64), (4, 96), (4, 128), (2, 192)]
    for it in range(steps):
        B, S = sizes[it % len(sizes)]
        x, m = synthetic_fundus(B, S, seed=5000 + it)
        out, _, _ = model(x)                          # draws its own eps (torch.manual_seed above)
        loss = crit(out, m)
        opt.zero_grad(); loss.backward(); opt.step()
        if it % log_every == 0 or it == steps - 1:
            with torch.no_grad():
                pos = (out > 0).float().mean().item()
            print(f"  fixture train step {it}: loss {loss.item():.4f} pos {pos:.4f} target {m.mean().item():.4f}", flush=True)
    for p in model.parameters():
        p.requires_grad_(True)
    model.zero_grad()


def make_resnet(mode="all", seed=0, train_steps=0):
    model = RefUNetResNet(3, 1, pretrained=False, latent_injection=mode)
    model.load_state_dict(seeded_state_dict(model.state_dict(), seed))
    imgs, _ = synthetic_fundus(8, 128, seed=77)
    g = torch.Generator().manual_seed(78)
    # heads first: mu ~ O(1), logvar ~ -1 +- 0.5 so that samples differ visibly
    model.eval()
    with torch.no_grad():
        for m in model.encoder.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.reset_running_stats(); m.momentum = None
        model.encoder.train(); model.encoder(imgs); model.encoder.eval()
        for m in model.encoder.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.momentum = 0.1; m.num_batches_tracked.zero_()
        f4 = model.encoder(imgs)[-1]
        mu = model.mu_head(f4).flatten(1); lv = model.logvar_head(f4).flatten(1)
        model.mu_head[0].weight.mul_(1.0 / mu.std()); model.mu_head[0].bias.copy_(-(mu.mean(0) / mu.std()))
        model.logvar_head[0].weight.mul_(0.5 / lv.std()); model.logvar_head[0].bias.copy_(-1.0 - 0.5 * lv.mean(0) / lv.std())
    eps = torch.randn(8, L, generator=g)
    with ref_shim.inject_eps([eps]):
        calibrate(model, lambda: model(imgs))
    extra = {"mu_head.0.weight", "mu_head.0.bias", "logvar_head.0.weight", "logvar_head.0.bias",
             "final_conv.weight", "final_conv.bias"}
    if train_steps > 0:
        # "trained-like" output stage: fit decoder_blocks.3 + final_conv (175 k parameters, the
        # only trained tensors stored) to the synthetic lesion masks with the reference's own CombinedLoss,
        # everything upstream frozen at its seeded init. Logits become bimodal with sharp lesion borders,
        # as a trained network's are, so the >= 99.9 % mask-agreement bar measures the kernels and not the
        # number of pixels a random network parks on the threshold.
        train_last_block(model, train_steps)
        extra |= {k for k in model.state_dict() if k.startswith("decoder_blocks.3.") and not k.endswith("num_batches_tracked")}
    else:
        with torch.no_grad():
            with ref_shim.inject_eps([eps]):
                out, _, _ = model(imgs)
            s = out.std()
            model.final_conv.weight.mul_(2.5 / s)
            model.final_conv.bias.copy_((model.final_conv.bias - out.mean()) * (2.5 / s) - 4.0)   # lesion-like prior: few positive pixels
    name = "fixture_resnet" + ("" if mode == "all" else "_" + mode)
    torch.save(small_overlay(model.state_dict(), extra), OUT / f"{name}.pt")
    return model


def golden_resnet():
    G = {}
    model = make_resnet("all", train_steps=TRAIN_STEPS)
    g = torch.Generator().manual_seed(101)
    # --- forward (eval), 2 x 3 x 96 x 96 and an odd size
    for tag, (B, H, W) in {"fwd96": (2, 96, 96), "fwd_odd": (1, 100, 76)}.items():
        x, _ = synthetic_fundus(B, H, W, seed=200 + H)
        eps = torch.randn(B, L, generator=g)
        with torch.no_grad(), ref_shim.inject_eps([eps]):
            out, mu, lv = model(x)
        G[tag] = dict(seed=200 + H, shape=(B, H, W), eps=eps, logits=out, mu=mu, logvar=lv)
    # --- per-stage intermediates on one input (encoder feats, decoder block 0 and 3 outputs)
    x, _ = synthetic_fundus(1, 96, 96, seed=300)
    with torch.no_grad():
        feats = model.encoder(x)
        mu, lv = model.encode(x)
        z = mu.unsqueeze(-1).unsqueeze(-1)
        zs = torch.nn.functional.interpolate(z, size=feats[-1].shape[2:], mode="bilinear", align_corners=True)
        x0 = model.z_initial(zs)
        d0 = model.decoder_blocks[0](x0, feats[-2], zs)
        att = model.decoder_blocks[1].attention(torch.nn.functional.interpolate(d0, size=feats[-3].shape[2:], mode="bilinear", align_corners=True), feats[-3])
    G["stages"] = dict(seed=300, feats=[f[:, ::4].clone() for f in feats], mu=mu, logvar=lv, x0=x0, d0=d0, att1=att)
    # --- generate_predictions: T=1.0 N=3 on 2 x 96 (config 1 shape family) and T=2.0 N=2
    for tag, (T, N) in {"gen_T1_N3": (1.0, 3), "gen_T2_N2": (2.0, 2)}.items():
        x, _ = synthetic_fundus(2, 96, 96, seed=400)
        eps = torch.randn(N, 2, L, generator=g)
        with ref_shim.inject_eps(list(eps)):
            pred = ref_vae.generate_predictions(model, x, T, N)
        G[tag] = dict(seed=400, T=T, N=N, eps=eps, pred=pred.detach())
    # --- predict_full_image + calculate_uncertainty_metrics (N=4, T=1)
    x, _ = synthetic_fundus(1, 96, 96, seed=500)
    eps = torch.randn(4, 1, L, generator=g)
    with torch.no_grad():
        mu, lv = ref_vae.encode_images(model, x)
        segs = []
        for n in range(4):
            z = (mu + eps[n] * torch.exp(0.5 * lv) * 1.0).unsqueeze(-1).unsqueeze(-1)
            segs.append(ref_vis.predict_full_image(model, x, z))
        segs = torch.stack(segs)[:, 0]           # [N,1,H,W]
        um = ref_vis.calculate_uncertainty_metrics(segs)
    G["full_image_mc"] = dict(seed=500, eps=eps, segs=segs, **{"um_" + k: v for k, v in um.items()})
    # --- predict_with_patches on an odd-sized image (last-tile snapping), one z
    x, _ = synthetic_fundus(1, 150, 170, seed=600)
    with torch.no_grad():
        mu, lv = ref_vae.encode_images(model, x)
        z = mu.unsqueeze(-1).unsqueeze(-1)
        pw = ref_vis.predict_with_patches(model, x, z, patch_size=64, overlap=16, batch_size=4)
    G["patches"] = dict(seed=600, shape=(150, 170), patch=64, overlap=16, mu=mu, out=pw)
    # --- encode / decode
    with torch.no_grad():
        zz = torch.randn(2, L, generator=g)
        dec = model.decode(zz)
        dec_r = model.decode(zz[:1], input_size=(40, 48))
    G["decode"] = dict(z=zz, out=dec[:, :, ::8, ::8].clone(), out_sum=dec.sum(), out_resized=dec_r)
    # --- training step: loss + grads (train-mode BN), B=2 x 64 x 64, EX and MA criteria
    for lesion in ("EX", "MA"):
        x, m = synthetic_fundus(2, 64, 64, seed=700, lesion=lesion)
        eps = torch.randn(2, L, generator=g)
        model.train(); model.zero_grad()
        sd_before = {k: v.clone() for k, v in model.state_dict().items()}
        with ref_shim.inject_eps([eps]):
            seg, mu, lv = model(x)
        crit = ref_loss.MASegmentationLoss(class_weight=0.9) if lesion == "MA" else ref_loss.CombinedLoss()
        recon = crit(seg, m)
        kl = ref_loss.kl_with_free_bits(mu, lv, free_bits=1e-3)
        loss = recon + 0.001 * kl
        loss.backward()
        names = ["encoder.conv1.weight", "encoder.layer1.0.bn1.weight", "encoder.layer2.0.downsample.0.weight",
                 "encoder.layer4.2.conv2.weight", "mu_head.0.weight", "logvar_head.0.bias", "z_initial.0.weight",
                 "decoder_blocks.0.attention.psi.0.weight", "decoder_blocks.0.conv1.0.weight",
                 "decoder_blocks.1.z_proj.0.weight", "decoder_blocks.3.conv2.1.bias", "final_conv.weight"]
        grads = {}
        for n_, p in model.named_parameters():
            gr = p.grad if p.grad is not None else torch.zeros_like(p)
            grads[n_] = dict(norm=gr.norm(), sum=gr.sum())
            if n_ in names:
                grads[n_]["full"] = gr.clone() if gr.numel() <= 300000 else gr.flatten()[:4096].clone()
        new_stats = {k: v.clone() for k, v in model.state_dict().items() if k.endswith("running_mean") and ("decoder_blocks.3" in k or "encoder.bn1" in k)}
        G["train_" + lesion] = dict(seed=700, eps=eps, loss=loss.detach(), recon=recon.detach(), kl=kl.detach(),
                                     seg=seg.detach(), grads=grads, new_running_mean=new_stats)
        model.load_state_dict(sd_before); model.eval()
    torch.save(G, OUT / "golden_resnet.pt")
    # --- other injection modes: forward only
    G2 = {}
    for mode in ("none", "bottleneck", "inject_no_bottleneck", "first"):
        m2 = make_resnet(mode)
        x, _ = synthetic_fundus(1, 64, 64, seed=800)
        eps = torch.randn(1, L, generator=g)
        with torch.no_grad(), ref_shim.inject_eps([eps]):
            out, mu, lv = m2(x)
        G2[mode] = dict(seed=800, eps=eps, logits=out, mu=mu, logvar=lv)
    torch.save(G2, OUT / "golden_resnet_modes.pt")


def golden_unet():
    G = {}
    for bilinear in (False, True):
        model = RefUNet(3, 1, bilinear)
        model.load_state_dict(seeded_state_dict(model.state_dict(), 1))
        imgs, _ = synthetic_fundus(4, 64, seed=79)
        calibrate(model, lambda: model(imgs))
        with torch.no_grad():
            out = model(imgs)
            s = out.std()
            model.outc.conv.weight.mul_(2.5 / s)
            model.outc.conv.bias.copy_((model.outc.conv.bias - out.mean()) * (2.5 / s) - 4.0)   # lesion-like prior: few positive pixels
        name = "fixture_unet" + ("_bilinear" if bilinear else "")
        torch.save(small_overlay(model.state_dict(), {"outc.conv.weight", "outc.conv.bias"}), OUT / f"{name}.pt")
        x, _ = synthetic_fundus(2, 64, 80, seed=900)
        with torch.no_grad():
            out = model(x)
            x1 = model.inc(x); x2 = model.down1(x1)
            u = model.up4(torch.nn.functional.relu(torch.randn(2, 64 if bilinear else 128, 32, 40, generator=torch.Generator().manual_seed(5))), x1)
        G["bilinear" if bilinear else "convT"] = dict(seed=900, shape=(2, 64, 80), logits=out, x2=x2[:, ::8].clone(), up4=u[:, ::8].clone())
    torch.save(G, OUT / "golden_unet.pt")


def golden_misc():
    G = {}
    g = torch.Generator().manual_seed(1000)
    x = torch.randn(3, 1, 40, 44, generator=g) * 3
    t = (torch.rand(3, 1, 40, 44, generator=g) > 0.85).float()
    G["loss"] = dict(x=x, t=t, dice=ref_loss.dice_loss(x, t), combined=ref_loss.CombinedLoss()(x, t),
                     focal=ref_loss.MAFocalLoss(alpha=0.9)(x, t), ma=ref_loss.MASegmentationLoss(class_weight=0.9)(x, t))
    mu = torch.randn(5, L, generator=g) * 0.5
    lv = torch.randn(5, L, generator=g) * 0.4
    G["kl"] = dict(mu=mu, lv=lv, kl_1e3=ref_loss.kl_with_free_bits(mu, lv, 1e-3), kl_0=ref_loss.kl_with_free_bits(mu, lv, 0.0),
                   kl_big=ref_loss.kl_with_free_bits(mu, lv, 0.05), stats=ref_vae.calculate_latent_stats(mu, lv))
    ann = {}
    for strat in ("linear", "cyclical", "constant"):
        a = ref_loss.KLAnnealer(0.0, 0.001, 10, strat)
        ann[strat] = [a.get_weight(e) for e in range(0, 14)] + [a.get_weight(3, 5, 20)]
    G["anneal"] = ann
    p = torch.rand(2, 50, 60, generator=g); p[0, 0, :4] = 1.0; p[0, 1, :4] = 0.0
    gt = (torch.rand(2, 50, 60, generator=g) > 0.7).float()
    ece, accs, confs, counts = ref_um.calculate_expected_calibration_error(p, gt)
    unc = torch.rand(2, 50, 60, generator=g) * 0.5
    G["metrics"] = dict(p=p, gt=gt, unc=unc, ece=ece, accs=torch.tensor(accs), confs=torch.tensor(confs),
                        counts=torch.tensor(counts), brier=ref_um.brier_score(p, gt),
                        nll=ref_um.calculate_negative_log_likelihood(p, gt),
                        ue_dice=ref_um.calculate_uncertainty_error_dice(unc, (p > 0.5).float(), gt, 0.2),
                        all_metrics=ref_metrics.get_all_metrics(p, gt))
    segs = torch.rand(6, 2, 1, 20, 24, generator=g)
    um = ref_vis.calculate_uncertainty_metrics(segs)
    G["uncertainty"] = dict(segs=segs, **um)
    s = ref_vae.sample_from_latent
    e = torch.randn(5, L, generator=g)
    with ref_shim.inject_eps([e]):
        G["sample"] = dict(eps=e, z=s(mu, lv, 1.7))
    torch.save(G, OUT / "golden_misc.pt")


if __name__ == "__main__":
    import logging
    logging.disable(logging.CRITICAL)
    golden_misc()
    golden_unet()
    golden_resnet()
    for f in sorted(OUT.glob("*.pt")):
        print(f.name, f.stat().st_size)
