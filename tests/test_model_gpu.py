"""GPU parity of the model-level paths (through the drop-in module surface and the C ABI) against
the FP32 oracle on the same seeded inputs and injected eps, for both 16-bit storage types of the kernels.

f16 (default; the reference's own AMP dtype) is held to BASELINE.json's bar on EVERY output: relative L2
error <= 1e-2 on logits, mu/logvar and uncertainty maps, thresholded masks agree on >= 99.9 % of pixels.
bf16 (selectable) has 3 fewer mantissa bits; through ~40 chained seeded-random layers its rounding noise
alone (oracle with bf16 storage rounding vs fp32 oracle) is 0.3-7 %, so it gets the bounds in TOLS["bf16"]
and its measured errors are reported next to that rounding budget (DESIGN.md "Numerics").
Every measured error is appended to gpurun_out/parity_report.txt."""
import os
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import vaeunet_oracle as O  # noqa: E402
from vaeunet_b200 import testing  # noqa: E402
from vaeunet_b200.synthetic import synthetic_fundus  # noqa: E402

TOL = 1e-2
#            logits/maps of fixture_resnet | mu, logvar, encoder features | fully random nets (modes, UNet)
TOLS = {"f16": dict(out=1e-2, mu=1e-2, random=1e-2, mask=0.999),
        "bf16": dict(out=3e-2, mu=5e-2, random=0.12, mask=0.999)}
MU_TOL = {k: v["mu"] for k, v in TOLS.items()}
_REPORT = Path(__file__).resolve().parent.parent / "gpurun_out" / "parity_report.txt"


def report(line: str):
    print(line)
    try:
        _REPORT.parent.mkdir(exist_ok=True)
        with open(_REPORT, "a") as f:
            f.write(line + "\n")
    except OSError:
        pass


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def mask_agree(a, b, thr=0.0):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a > thr) == (b > thr)).float().mean().item()


@pytest.fixture(scope="module", params=["bf16", "f16"])
def model(cuda_dev, request):
    from vaeunet_b200.unet import UNetResNet
    sd, kw = testing.resnet_fixture()
    m = UNetResNet(**kw)
    m.load_state_dict(sd)
    m = m.to(cuda_dev).eval()
    m.compute_dtype = request.param
    m.dt_name = request.param
    return m, sd


CFG = O.ResNetCfg()


@pytest.mark.parametrize("B,H,W", [(2, 96, 96), (1, 100, 76), (2, 256, 256)])
def test_forward_parity(model, cuda_dev, B, H, W):
    m, sd = model
    x, _ = synthetic_fundus(B, H, W, seed=200 + H)
    eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(1))
    ref, mu_r, lv_r = O.unetresnet_forward(sd, CFG, x, eps)
    with torch.no_grad():
        out, mu, lv = m(x.to(cuda_dev), eps=eps.to(cuda_dev))
    assert out.shape == ref.shape and mu.shape == (B, 32)
    r, agree = rel(out, ref), mask_agree(out, ref)
    near = ((ref.abs() < 0.05).float().mean().item())
    line = (f"forward {m.dt_name} {B}x{H}x{W}: logits rel={r:.5f} mask_agree={agree:.5f} mu rel={rel(mu, mu_r):.5f} "
            f"logvar rel={rel(lv, lv_r):.5f} frac|logit|<0.05={near:.5f} pos={(ref > 0).float().mean().item():.4f}")
    if m.dt_name == "bf16":
        with O.storage_rounding(torch.bfloat16):
            emu, mu_e, _ = O.unetresnet_forward(sd, CFG, x, eps)
        line += f" | vs bf16-rounded oracle: logits {rel(out, emu):.5f} mu {rel(mu, mu_e):.5f} (rounded vs fp32 oracle {rel(emu, ref):.5f})"
    report(line)
    assert rel(mu, mu_r) < MU_TOL[m.dt_name] and rel(lv, lv_r) < MU_TOL[m.dt_name]
    assert r < TOLS[m.dt_name]["out"]
    assert agree >= TOLS[m.dt_name]["mask"]


def test_module_surface_stages(model, cuda_dev):
    """The reference's scripts call sub-modules one by one (utils/vae_utils.py:18-67)."""
    m, sd = model
    x, _ = synthetic_fundus(1, 96, 96, seed=300)
    feats_r = O.resnet34_features(sd, x)
    with torch.no_grad():
        feats = m.encoder(x.to(cuda_dev))
        assert [tuple(f.shape) for f in feats] == [tuple(f.shape) for f in feats_r]
        for i, (a, b) in enumerate(zip(feats, feats_r)):
            report(f"stages {m.dt_name}: encoder feature f{i} rel={rel(a, b):.5f}")
            assert rel(a, b) < MU_TOL[m.dt_name]
        x_enc = feats_r[-1].to(cuda_dev)
        mu = m.mu_head(x_enc).squeeze(-1).squeeze(-1)
        lv = m.logvar_head(x_enc).squeeze(-1).squeeze(-1)
        mu_r, lv_r = O.latent_heads(sd, feats_r[-1])
        assert rel(mu, mu_r) < TOL and rel(lv, lv_r) < TOL
        z = mu_r.unsqueeze(-1).unsqueeze(-1)
        zs = torch.nn.functional.interpolate(z, size=feats_r[-1].shape[2:], mode="bilinear", align_corners=True)
        x0 = m.z_initial(zs.to(cuda_dev))
        x0_r = torch.relu(O._bn(sd, "z_initial.1.", O._conv(sd, "z_initial.0.", zs), None))
        assert rel(x0, x0_r) < TOL
        hooked = []
        h = m.decoder_blocks[0].attention.psi.register_forward_hook(lambda mod, i, o: hooked.append(o))
        d0 = m.decoder_blocks[0](x0_r.to(cuda_dev), feats_r[-2].to(cuda_dev), zs.to(cuda_dev))
        h.remove()
        d0_r = O.decoder_block(sd, "decoder_blocks.0.", x0_r, feats_r[-2], zs, True, True, True)
        assert rel(d0, d0_r) < TOLS[m.dt_name]['out']
        assert len(hooked) == 1 and hooked[0].shape == (1, 1, 6, 6)      # analyze_model.py:733-738
        fin = m.final_conv(torch.relu(torch.randn(1, 64, 48, 48, generator=torch.Generator().manual_seed(2))).to(cuda_dev))
        fin_r = O._conv(sd, "final_conv.", torch.relu(torch.randn(1, 64, 48, 48, generator=torch.Generator().manual_seed(2))))
        assert rel(fin, fin_r) < TOL
        mu2, lv2 = m.encode(x.to(cuda_dev))
        assert rel(mu2, mu_r) < MU_TOL[m.dt_name]


def test_generate_predictions(model, cuda_dev):
    from vaeunet_b200.utils.vae_utils import generate_predictions, sample_from_latent
    m, sd = model
    x, _ = synthetic_fundus(2, 96, 96, seed=400)
    eps = torch.randn(3, 2, 32, generator=torch.Generator().manual_seed(3))
    ref = O.generate_predictions(sd, CFG, x, 2.0, 3, eps)
    got = generate_predictions(m, x.to(cuda_dev), 2.0, 3, eps=eps.to(cuda_dev))
    assert got.shape == ref.shape == (2, 1, 48, 48)
    report(f"generate_predictions {m.dt_name}: rel={rel(got, ref):.5f}")
    assert rel(got, ref) < TOLS[m.dt_name]["out"]
    mu = torch.randn(4, 32); lv = torch.randn(4, 32) * 0.3; e = torch.randn(4, 32)
    z = sample_from_latent(mu.to(cuda_dev), lv.to(cuda_dev), 1.7, eps=e.to(cuda_dev))
    assert (z.cpu() - O.sample_from_latent(mu, lv, 1.7, e)).abs().max() < 1e-5


@pytest.mark.parametrize("B,S,N,T", [(2, 128, 3, 1.0), (1, 512, 10, 1.0)])
def test_mc_predict(model, cuda_dev, B, S, N, T):
    """BASELINE config 1/2 shapes: mean / std / entropy maps over N latent samples."""
    m, sd = model
    x, _ = synthetic_fundus(B, S, seed=1234)
    eps = torch.randn(N, B, 32, generator=torch.Generator().manual_seed(4))
    ref = O.mc_patch_uncertainty(sd, CFG, x, T, N, eps)
    with torch.no_grad():
        res = m.mc_predict(x.to(cuda_dev), T, N, eps=eps.to(cuda_dev), return_samples=True)
    tol = TOLS[m.dt_name]["out"]
    report(f"mc_predict {m.dt_name} B={B} {S}x{S} N={N}: " + " ".join(
        f"{k} rel={rel(res[k], ref[k]):.5f} maxabs={(res[k].cpu() - ref[k]).abs().max().item():.5f}"
        for k in ("mean", "std", "entropy", "mutual_info")) + f" mask_agree={mask_agree(res['mean'], ref['mean'], 0.5):.5f}")
    for k in ("mean", "entropy"):
        assert rel(res[k], ref[k]) < tol, k
    # std/MI are differences of nearly equal samples: compare on an absolute scale tied to p in [0,1]
    for k in ("std", "mutual_info"):
        err = (res[k].cpu() - ref[k]).abs().max().item()
        print(f"mc {S} N={N} {k}: max abs err {err:.5f} (ref max {ref[k].abs().max().item():.4f}) rel {rel(res[k], ref[k]):.4f}")
        assert err < (2e-2 if m.dt_name == "f16" else 5e-2), k
        if m.dt_name == "f16":
            assert rel(res[k], ref[k]) < tol, k
    assert rel(res["samples"].permute(1, 0, 2, 3), ref["samples"]) < tol
    assert mask_agree(res["mean"], ref["mean"], 0.5) >= 0.999
    assert ref["std"].max() > 1e-3            # the fixture really has sample-to-sample variation


@pytest.mark.parametrize("S", [512, 200])
def test_mc_decode_batch_invariance(model, cuda_dev, S):
    """One decode's logits do not depend on how many other decodes share the launch, bit for bit, and are reproducible run
    to run - the property sample-sharded MC (distributed.py, BASELINE config 4) relies on for sharded == unsharded. The
    latent head sums its global-average-pool partials in a fixed order for this reason (vu_gate_latent.cu)."""
    m, _ = model
    x, _ = synthetic_fundus(1, S, seed=21)
    xf = x.to(cuda_dev).float().contiguous()
    eps = torch.randn(15, 1, 32, generator=torch.Generator().manual_seed(3)).to(cuda_dev)
    eng = m.engine()
    full = eng.mc_logits(xf, 2.0, 15, eps, True)[0].clone()
    again = eng.mc_logits(xf, 2.0, 15, eps, True)[0].clone()
    a = eng.mc_logits(xf, 2.0, 8, eps[:8], True)[0].clone()
    b = eng.mc_logits(xf, 2.0, 7, eps[8:], True)[0].clone()
    assert torch.equal(full, again)
    assert torch.equal(full.reshape(15, -1), torch.cat([a, b], dim=1).reshape(15, -1))
    assert full.reshape(15, -1).std(dim=0).max() > 0      # the samples really differ from each other


def test_predict_paths(model, cuda_dev):
    from vaeunet_b200 import predict as P
    m, sd = model
    x, _ = synthetic_fundus(1, 150, 170, seed=600)
    mu_r, lv_r = O.unetresnet_encode(sd, x)
    z = mu_r.unsqueeze(-1).unsqueeze(-1)
    ref = O.predict_with_patches(sd, CFG, x, mu_r, 64, 16)
    got = P.predict_with_patches(m, x.to(cuda_dev), z.to(cuda_dev), 64, 16, batch_size=4)
    assert got.shape == ref.shape
    tol = TOLS[m.dt_name]["out"]
    ref_f = O.predict_full_image(sd, CFG, x, mu_r)
    got_f = P.predict_full_image(m, x.to(cuda_dev), z.to(cuda_dev))
    report(f"predict_with_patches {m.dt_name}: rel={rel(got, ref):.5f} maxabs={(got.cpu() - ref).abs().max().item():.5f}; "
           f"predict_full_image rel={rel(got_f, ref_f):.5f}")
    assert (got.cpu() - ref).abs().max() < 2 * tol and rel(got, ref) < tol
    assert rel(got_f, ref_f) < tol
    eps = torch.randn(3, 1, 32, generator=torch.Generator().manual_seed(5))
    segs_r, _, _ = O.segmentation_distribution(sd, CFG, x, 3, 1.5, eps, 64, 16)
    segs, mu, lv = P.segmentation_distribution(m, x.to(cuda_dev), 3, 64, 16, 1.5, 4, eps=eps.to(cuda_dev))
    assert rel(segs, segs_r) < tol
    um = P.calculate_uncertainty_metrics(segs)
    um_r = O.calculate_uncertainty_metrics(segs_r.unsqueeze(1))
    assert rel(um["mean"], um_r["mean"]) < tol


def test_evaluate_validation_pass(model, cuda_dev):
    """evaluate.evaluate (evaluate.py:19-101) - same averaged metrics as the oracle restatement (pinned to
    the unmodified reference by tests/test_oracle_golden.py::test_evaluate), incl. max_samples and dict batches."""
    from vaeunet_b200.evaluate import evaluate
    m, sd = model
    seeds = [900, 901, 902]
    batches = [synthetic_fundus(2, 64, 80, seed=s) for s in seeds]
    g = torch.Generator().manual_seed(77)
    eps = [torch.randn(2, 32, generator=g) for _ in seeds]
    ref, kept = O.evaluate_metrics(sd, CFG, batches, eps)
    got, samples = evaluate(m, batches, cuda_dev, amp=False, eps=eps)
    report(f"evaluate {m.dt_name}: " + " ".join(f"{k} {got[k]:.5f}/{ref[k]:.5f}" for k in ref))
    for k, v in ref.items():
        assert abs(got[k] - v) <= 2e-2 * max(abs(v), 1e-3), (k, got[k], v)       # thresholded counts: pixels near logit 0.5 flip
    assert len(samples) == 4 and samples[0][1].shape == (1, 64, 80)
    assert abs(float(samples[0][1].mean()) - kept[0][0].mean().item()) < 5e-3
    ref2, _ = O.evaluate_metrics(sd, CFG, batches, eps, max_samples=2)
    got2, _ = evaluate(m, [dict(image=a, mask=b) for a, b in batches], cuda_dev, amp=False, max_samples=2, eps=eps)
    for k, v in ref2.items():
        assert abs(got2[k] - v) <= 2e-2 * max(abs(v), 1e-3), (k, got2[k], v)


def test_temperature_ensemble(model, cuda_dev):
    """visualize_vae.generate_ensemble_prediction (:1083-1163) - all (T, sample) draws in one tiled sweep."""
    from vaeunet_b200 import predict as P
    m, sd = model
    x, _ = synthetic_fundus(1, 100, 120, seed=910)
    temps = [0.5, 1.0, 2.5]
    eps = torch.randn(6, 1, 32, generator=torch.Generator().manual_seed(12))
    tol = TOLS[m.dt_name]["out"]
    for weighted in (True, False):
        ref = O.generate_ensemble_prediction(sd, CFG, x, temps, 2, eps, weighted, 64, 16)
        got = P.generate_ensemble_prediction(m, x.to(cuda_dev), temps, 2, weighted, 64, 16, 4, eps=eps.to(cuda_dev))
        assert got.shape == ref.shape == (1, 1, 100, 120)
        report(f"ensemble {m.dt_name} weighted={weighted}: rel={rel(got, ref):.5f} maxabs={(got.cpu() - ref).abs().max().item():.5f}")
        assert rel(got, ref) < tol and (got.cpu() - ref).abs().max() < 2 * tol
    w = P.ensemble_weights(temps, True)
    assert abs(w.sum().item() - 1.0) < 1e-6 and w[1] > w[0] > w[2]


# The injection-mode and basic-UNet fixtures are fully seeded-random networks (no fitted output stage):
# they amplify rounding noise ~10x more than fixture_resnet -> TOLS[dt]["random"].
_TDT = {"bf16": torch.bfloat16, "f16": torch.float16}


@pytest.mark.parametrize("dt", ["bf16", "f16"])
def test_injection_modes(cuda_dev, dt):
    from vaeunet_b200.unet import UNetResNet
    for mode in ("none", "bottleneck", "inject_no_bottleneck", "first"):
        sd, kw = testing.resnet_fixture(latent_injection=mode)
        m = UNetResNet(**kw); m.load_state_dict(sd); m = m.to(cuda_dev).eval()
        m.compute_dtype = dt
        x, _ = synthetic_fundus(1, 64, 64, seed=800)
        eps = torch.randn(1, 32, generator=torch.Generator().manual_seed(6))
        cfg = O.ResNetCfg(latent_injection=mode)
        ref, _, _ = O.unetresnet_forward(sd, cfg, x, eps)
        with O.storage_rounding(_TDT[dt]):
            emu, _, _ = O.unetresnet_forward(sd, cfg, x, eps)
        with torch.no_grad():
            out, _, _ = m(x.to(cuda_dev), eps=eps.to(cuda_dev))
        report(f"injection {mode} {dt}: rel vs fp32 oracle={rel(out, ref):.5f} vs {dt}-rounded oracle={rel(out, emu):.5f} "
               f"(rounded vs fp32 oracle {rel(emu, ref):.5f})")
        assert rel(out, emu) < TOLS[dt]["random"], mode
        assert rel(out, ref) < TOLS[dt]["random"], mode


@pytest.mark.parametrize("dt", ["bf16", "f16"])
@pytest.mark.parametrize("bilinear", [False, True])
def test_basic_unet(cuda_dev, bilinear, dt):
    from vaeunet_b200.unet import UNet
    sd = testing.unet_fixture(1, bilinear)
    m = UNet(3, 1, bilinear); m.load_state_dict(sd); m = m.to(cuda_dev).eval()
    m.compute_dtype = dt
    x, _ = synthetic_fundus(2, 64, 80, seed=900)
    ref = O.unet_forward(sd, x, bilinear)
    with O.storage_rounding(_TDT[dt]):
        emu = O.unet_forward(sd, x, bilinear)
    xr = torch.relu(torch.randn(2, 64 if bilinear else 128, 32, 40, generator=torch.Generator().manual_seed(5)))
    with torch.no_grad():
        out = m(x.to(cuda_dev))
        x1 = m.inc(x.to(cuda_dev))
        u = m.up4(xr.to(cuda_dev), x1)
    report(f"unet bilinear={bilinear} {dt}: rel vs fp32 oracle={rel(out, ref):.5f} vs {dt}-rounded oracle={rel(out, emu):.5f} "
           f"mask_agree={mask_agree(out, ref):.5f}")
    assert out.shape == ref.shape and rel(out, emu) < TOLS[dt]["random"]
    x1_r = O.double_conv(sd, "inc.", x)
    assert rel(x1, x1_r) < TOL
    u_r = O.unet_up(sd, "up4.", xr, x1_r, bilinear)
    assert rel(u, u_r) < TOL
    assert rel(out, ref) < TOLS[dt]["random"]
    if dt == "f16":
        assert mask_agree(out, ref) >= 0.999


def test_losses_autograd(cuda_dev):
    from vaeunet_b200.utils import loss as Lz
    g = torch.Generator().manual_seed(7)
    x = (torch.randn(2, 1, 64, 64, generator=g) * 3)
    t = (torch.rand(2, 1, 64, 64, generator=g) > 0.9).float()
    for mine, ref_fn in ((Lz.CombinedLoss(), O.combined_loss), (Lz.MASegmentationLoss(class_weight=0.9), O.ma_segmentation_loss),
                         (Lz.DiceLoss(), O.dice_loss)):
        xr = x.clone().requires_grad_(True)
        lr = ref_fn(xr, t); gr, = torch.autograd.grad(lr, xr)
        xd = x.to(cuda_dev).requires_grad_(True)
        ld = mine(xd, t.to(cuda_dev)); ld.backward()
        assert abs(ld.item() - lr.item()) < 1e-5
        assert (xd.grad.cpu() - gr).abs().max() <= 1e-3 * gr.abs().max() + 1e-9
    mu = (torch.randn(4, 32, generator=g) * 0.5); lv = torch.randn(4, 32, generator=g) * 0.3
    mud = mu.to(cuda_dev).requires_grad_(True); lvd = lv.to(cuda_dev).requires_grad_(True)
    kl = Lz.kl_with_free_bits(mud, lvd, 1e-3); kl.backward()
    assert abs(kl.item() - O.kl_with_free_bits(mu, lv, 1e-3).item()) < 1e-5
    a = Lz.KLAnnealer(0.0, 0.001, 10)
    assert abs(a.get_weight(5) - O.kl_anneal_weight(5, 0.0, 0.001, 10)) < 1e-12


def test_uncertainty_metric_functions(cuda_dev):
    from vaeunet_b200.utils import uncertainty_metrics as U, metrics as Mx
    g = torch.Generator().manual_seed(8)
    p = torch.rand(2, 50, 60, generator=g); gt = (torch.rand(2, 50, 60, generator=g) > 0.7).float()
    unc = torch.rand(2, 50, 60, generator=g) * 0.5
    d = cuda_dev
    ece, accs, confs, counts = U.calculate_expected_calibration_error(p.to(d), gt.to(d))
    ece_r, accs_r, confs_r, counts_r = O.expected_calibration_error(p, gt)
    assert abs(ece - ece_r) < 1e-6 and (counts == counts_r).all()
    assert abs(U.brier_score(p.to(d), gt.to(d)) - O.brier_score(p, gt)) < 1e-6
    assert abs(U.calculate_negative_log_likelihood(p.to(d), gt.to(d)) - O.negative_log_likelihood(p, gt)) < 1e-5
    pb = (p > 0.5).float()
    assert abs(U.calculate_uncertainty_error_dice(unc.to(d), pb.to(d), gt.to(d), 0.2) - O.uncertainty_error_dice(unc, pb, gt, 0.2)) < 1e-6
    am, ar = Mx.get_all_metrics(p.to(d), gt.to(d)), O.threshold_metrics(p, gt)
    for k in ar:
        assert abs(am[k] - ar[k]) < 1e-6, k


def test_no_cpu_fallback():
    """The product path must fail loudly off-device (BASELINE: no CPU fallback)."""
    from vaeunet_b200.unet import UNetResNet
    from vaeunet_b200._lib import VaeUnetError
    sd, kw = testing.resnet_fixture()
    m = UNetResNet(**kw); m.load_state_dict(sd); m.eval()
    with pytest.raises(VaeUnetError):
        m(torch.zeros(1, 3, 64, 64))
