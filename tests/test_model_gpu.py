"""GPU parity of the model-level paths (through the drop-in module surface and the C ABI) against
the oracle on the same seeded inputs and injected eps. Tolerances are BASELINE.json's:
bf16 path: relative L2 error <= 1e-2 on logits / uncertainty maps, masks agree on >= 99.9 % of pixels."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import vaeunet_oracle as O  # noqa: E402
from vaeunet_b200 import testing  # noqa: E402
from vaeunet_b200.synthetic import synthetic_fundus  # noqa: E402

TOL = 1e-2


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def mask_agree(a, b, thr=0.0):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a > thr) == (b > thr)).float().mean().item()


@pytest.fixture(scope="module")
def model(cuda_dev):
    from vaeunet_b200.unet import UNetResNet
    sd, kw = testing.resnet_fixture()
    m = UNetResNet(**kw)
    m.load_state_dict(sd)
    return m.to(cuda_dev).eval(), sd


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
    assert rel(mu, mu_r) < TOL and rel(lv, lv_r) < TOL
    r = rel(out, ref)
    near = ((ref.abs() < 0.05).float().mean().item())
    agree = mask_agree(out, ref)
    print(f"forward {B}x{H}x{W}: rel={r:.4f} mask_agree={agree:.5f} frac|logit|<0.05={near:.4f}")
    assert r < TOL
    assert agree >= 0.999


def test_module_surface_stages(model, cuda_dev):
    """The reference's scripts call sub-modules one by one (utils/vae_utils.py:18-67)."""
    m, sd = model
    x, _ = synthetic_fundus(1, 96, 96, seed=300)
    feats_r = O.resnet34_features(sd, x)
    with torch.no_grad():
        feats = m.encoder(x.to(cuda_dev))
        assert [tuple(f.shape) for f in feats] == [tuple(f.shape) for f in feats_r]
        for a, b in zip(feats, feats_r):
            assert rel(a, b) < TOL
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
        assert rel(d0, d0_r) < TOL
        assert len(hooked) == 1 and hooked[0].shape == (1, 1, 6, 6)      # analyze_model.py:733-738
        fin = m.final_conv(torch.relu(torch.randn(1, 64, 48, 48, generator=torch.Generator().manual_seed(2))).to(cuda_dev))
        fin_r = O._conv(sd, "final_conv.", torch.relu(torch.randn(1, 64, 48, 48, generator=torch.Generator().manual_seed(2))))
        assert rel(fin, fin_r) < TOL
        mu2, lv2 = m.encode(x.to(cuda_dev))
        assert rel(mu2, mu_r) < TOL


def test_generate_predictions(model, cuda_dev):
    from vaeunet_b200.utils.vae_utils import generate_predictions, sample_from_latent
    m, sd = model
    x, _ = synthetic_fundus(2, 96, 96, seed=400)
    eps = torch.randn(3, 2, 32, generator=torch.Generator().manual_seed(3))
    ref = O.generate_predictions(sd, CFG, x, 2.0, 3, eps)
    got = generate_predictions(m, x.to(cuda_dev), 2.0, 3, eps=eps.to(cuda_dev))
    assert got.shape == ref.shape == (2, 1, 48, 48)
    assert rel(got, ref) < TOL
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
    for k in ("mean", "entropy"):
        assert rel(res[k], ref[k]) < TOL, k
    # std/MI are differences of nearly equal samples: compare on an absolute scale tied to p in [0,1]
    for k in ("std", "mutual_info"):
        err = (res[k].cpu() - ref[k]).abs().max().item()
        print(f"mc {S} N={N} {k}: max abs err {err:.5f} (ref max {ref[k].abs().max().item():.4f}) rel {rel(res[k], ref[k]):.4f}")
        assert err < 2e-2, k
    assert rel(res["samples"].permute(1, 0, 2, 3), ref["samples"]) < TOL
    assert mask_agree(res["mean"], ref["mean"], 0.5) >= 0.999
    assert ref["std"].max() > 1e-3            # the fixture really has sample-to-sample variation


def test_predict_paths(model, cuda_dev):
    from vaeunet_b200 import predict as P
    m, sd = model
    x, _ = synthetic_fundus(1, 150, 170, seed=600)
    mu_r, lv_r = O.unetresnet_encode(sd, x)
    z = mu_r.unsqueeze(-1).unsqueeze(-1)
    ref = O.predict_with_patches(sd, CFG, x, mu_r, 64, 16)
    got = P.predict_with_patches(m, x.to(cuda_dev), z.to(cuda_dev), 64, 16, batch_size=4)
    assert got.shape == ref.shape
    assert (got.cpu() - ref).abs().max() < 2e-2 and rel(got, ref) < TOL
    ref_f = O.predict_full_image(sd, CFG, x, mu_r)
    got_f = P.predict_full_image(m, x.to(cuda_dev), z.to(cuda_dev))
    assert rel(got_f, ref_f) < TOL
    eps = torch.randn(3, 1, 32, generator=torch.Generator().manual_seed(5))
    segs_r, _, _ = O.segmentation_distribution(sd, CFG, x, 3, 1.5, eps, 64, 16)
    segs, mu, lv = P.segmentation_distribution(m, x.to(cuda_dev), 3, 64, 16, 1.5, 4, eps=eps.to(cuda_dev))
    assert rel(segs, segs_r) < TOL
    um = P.calculate_uncertainty_metrics(segs)
    um_r = O.calculate_uncertainty_metrics(segs_r.unsqueeze(1))
    assert rel(um["mean"], um_r["mean"]) < TOL


def test_injection_modes(cuda_dev):
    from vaeunet_b200.unet import UNetResNet
    for mode in ("none", "bottleneck", "inject_no_bottleneck", "first"):
        sd, kw = testing.resnet_fixture(latent_injection=mode)
        m = UNetResNet(**kw); m.load_state_dict(sd); m = m.to(cuda_dev).eval()
        x, _ = synthetic_fundus(1, 64, 64, seed=800)
        eps = torch.randn(1, 32, generator=torch.Generator().manual_seed(6))
        ref, _, _ = O.unetresnet_forward(sd, O.ResNetCfg(latent_injection=mode), x, eps)
        with torch.no_grad():
            out, _, _ = m(x.to(cuda_dev), eps=eps.to(cuda_dev))
        assert rel(out, ref) < TOL, mode


@pytest.mark.parametrize("bilinear", [False, True])
def test_basic_unet(cuda_dev, bilinear):
    from vaeunet_b200.unet import UNet
    sd = testing.unet_fixture(1, bilinear)
    m = UNet(3, 1, bilinear); m.load_state_dict(sd); m = m.to(cuda_dev).eval()
    x, _ = synthetic_fundus(2, 64, 80, seed=900)
    ref = O.unet_forward(sd, x, bilinear)
    with torch.no_grad():
        out = m(x.to(cuda_dev))
        x1 = m.inc(x.to(cuda_dev))
        u = m.up4(torch.relu(torch.randn(2, 64 if bilinear else 128, 32, 40, generator=torch.Generator().manual_seed(5))).to(cuda_dev), x1)
    assert out.shape == ref.shape and rel(out, ref) < TOL and mask_agree(out, ref) >= 0.999
    x1_r = O.double_conv(sd, "inc.", x)
    assert rel(x1, x1_r) < TOL
    u_r = O.unet_up(sd, "up4.", torch.relu(torch.randn(2, 64 if bilinear else 128, 32, 40, generator=torch.Generator().manual_seed(5))), x1_r, bilinear)
    assert rel(u, u_r) < TOL


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
