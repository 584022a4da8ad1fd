"""GPU parity at the shapes BASELINE.json names (configs[2..4]) and for the module-surface paths round 1 left untested:
  * mc_predict at 700 x 700, T = 2.0, N = 15 (configs[3]: odd feature pyramid 350 / 175 / 88 / 44 / 22);
  * one TrainStep at B = 16 x 512 x 512 (configs[2]) against the fp32 oracle's loss and gradients;
  * a 3 x 3 tiled image (patch 512, overlap 100) through segmentation_distribution + calculate_uncertainty_metrics
    (configs[4]'s pipeline, visualize_vae.py:243-415,578-652);
  * UNetResNet.decode(z, input_size) (unet_resnet.py:250-279), predict_single_patch (visualize_vae.py:418-476);
  * use_skip=False / use_attention=False models (the reference's --no-skip): forward, generate_predictions, mc_predict.
Bars (BASELINE.json): relative L2 <= 1e-2 at 16 bit (f16, the product dtype), thresholded masks >= 99.9 %."""
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


def _model(dev, dtype="f16", **kw):
    from vaeunet_b200.unet import UNetResNet
    if "use_skip" in kw or "use_attention" in kw:
        sd, mkw = testing.resnet_variant_state(**kw)
    else:
        sd, mkw = testing.resnet_fixture(**kw)
    m = UNetResNet(**mkw)
    m.load_state_dict(sd)
    m = m.to(dev).eval()
    m.compute_dtype = dtype
    return m, sd, O.ResNetCfg(use_attention=mkw.get("use_attention", True), use_skip=mkw.get("use_skip", True),
                              latent_injection=mkw.get("latent_injection", "all"))


def test_mc_predict_700_T2_N15(cuda_dev):
    """configs[3] on one GPU: MA-like 700 x 700 patch, T = 2.0, N = 15."""
    m, sd, cfg = _model(cuda_dev)
    x, _ = synthetic_fundus(1, 700, 700, seed=703, lesion="MA")
    eps = torch.randn(15, 1, 32, generator=torch.Generator().manual_seed(704))
    ref = O.mc_patch_uncertainty(sd, cfg, x, 2.0, 15, eps)
    with torch.no_grad():
        res = m.mc_predict(x.to(cuda_dev), 2.0, 15, eps=eps.to(cuda_dev))
    torch.cuda.synchronize()
    assert res["mean"].shape == (1, 700, 700)
    for k in ("mean", "std", "entropy", "mutual_info"):
        assert rel(res[k], ref[k]) <= TOL, (k, rel(res[k], ref[k]))
    agree = ((res["mean"].cpu() > 0.5) == (ref["mean"] > 0.5)).float().mean().item()
    assert agree >= 0.999, agree
    assert rel(res["mu"], ref["mu"]) <= TOL and rel(res["logvar"], ref["logvar"]) <= TOL


def test_train_step_b16_512(cuda_dev):
    """configs[2]: ONE training step at batch 16 of 512 x 512 against the oracle's autograd (the shape bench.py times).
    f16 + loss scale (the reference trains under fp16 autocast + GradScaler) and bf16 (what the bench ran in round 1)."""
    import psutil
    if psutil.virtual_memory().available < 96 * 2 ** 30:
        pytest.skip("the fp32 CPU oracle of a B=16 x 512^2 step needs ~60 GB of host memory")
    from vaeunet_b200 import ops
    from vaeunet_b200.training import TrainStep
    from vaeunet_b200.unet import UNetResNet
    dev = cuda_dev
    sd, kw = testing.resnet_fixture()
    x, msk = synthetic_fundus(16, 512, 512, seed=161)
    eps = torch.randn(16, 32, generator=torch.Generator().manual_seed(162))
    beta = 0.001
    loss_r, recon_r, kl_r, grads_r, _ = O.training_grads(sd, O.ResNetCfg(), x, msk, eps, beta, 1e-3, "EX")
    fr = torch.cat([g.flatten() for g in grads_r.values()])
    for dtype, scale, loss_tol, cos_floor in ((ops.F16, 1024.0, 2e-3, 0.985), (ops.BF16, 1.0, 1e-2, 0.90)):
        model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
        ts = TrainStep(model, lesion="EX", dtype=dtype, loss_scale=scale)
        ts.zero_grad()
        out = ts.forward_backward(x.to(dev), msk.to(dev), beta, eps.to(dev))
        torch.cuda.synchronize()
        assert abs(out["loss"].item() - loss_r.item()) <= loss_tol * abs(loss_r.item()), (dtype, out["loss"].item(), loss_r.item())
        assert abs(out["kl"].item() - kl_r.item()) <= loss_tol * abs(kl_r.item()) + 1e-5
        fg = torch.cat([(ts.G[k] / scale).cpu().flatten() for k in grads_r])
        cos = (fg.double() @ fr.double() / (fg.double().norm() * fr.double().norm())).item()
        assert cos >= cos_floor, (dtype, cos)
        del ts, model
        torch.cuda.empty_cache()


def test_tiled_image_3x3_uncertainty(cuda_dev):
    """configs[4]'s pipeline on a 1336 x 1336 image = exactly 3 x 3 tiles of 512 with overlap 100: whole-image mu / logvar,
    N = 2 samples decoded tile by tile, ramp blending, per-pixel uncertainty maps - against the oracle."""
    from vaeunet_b200 import predict
    m, sd, cfg = _model(cuda_dev)
    x, _ = synthetic_fundus(1, 1336, 1336, seed=1336)
    eps = torch.randn(2, 1, 32, generator=torch.Generator().manual_seed(1337))
    assert predict.patch_grid(1336, 1336, 512, 100)[:2] == (3, 3)
    segs_r, mu_r, lv_r = O.segmentation_distribution(sd, cfg, x, 2, 1.0, eps, patch_size=512, overlap=100)
    um_r = O.calculate_uncertainty_metrics(segs_r.unsqueeze(1))
    segs, mu, lv = predict.segmentation_distribution(m, x.to(cuda_dev), 2, 512, 100, 1.0, batch_size=4, eps=eps.to(cuda_dev))
    um = predict.calculate_uncertainty_metrics(segs)
    torch.cuda.synchronize()
    assert segs.shape == (2, 1, 1336, 1336)
    assert rel(segs, segs_r) <= TOL, rel(segs, segs_r)
    assert rel(mu, mu_r) <= TOL
    for k in ("mean", "std", "entropy"):
        assert rel(um[k][0], um_r[k][0]) <= 2e-2 if k == "std" else rel(um[k][0], um_r[k][0]) <= TOL, (k, rel(um[k][0], um_r[k][0]))
    agree = ((segs.cpu() > 0.5) == (segs_r > 0.5)).float().mean().item()
    assert agree >= 0.999, agree


def test_decode_and_single_patch(cuda_dev):
    from vaeunet_b200 import predict
    m, sd, cfg = _model(cuda_dev)
    g = torch.Generator().manual_seed(250)
    z = torch.randn(2, 32, generator=g)
    ref = O.unetresnet_decode(sd, cfg, z)
    got = m.decode(z.to(cuda_dev))
    assert got.shape == ref.shape == (2, 1, 256, 256)
    assert rel(got, ref) <= TOL, rel(got, ref)
    ref_r = O.unetresnet_decode(sd, cfg, z[:1], input_size=(40, 48))
    got_r = m.decode(z[:1].to(cuda_dev), input_size=(40, 48))
    assert got_r.shape == (1, 1, 40, 48) and rel(got_r, ref_r) <= TOL
    # predict_single_patch (visualize_vae.py:418-476): tile (1, 0) of a 2 x 3 grid -> ramps on top, bottom?, right
    patch, _ = synthetic_fundus(1, 256, 320, seed=251)
    zz = torch.randn(1, 32, generator=g)
    pred, wt = predict.predict_single_patch(m, patch.to(cuda_dev), zz.to(cuda_dev).view(1, 32, 1, 1), cuda_dev, overlap=50,
                                            patch_grid=(2, 3), patch_idx=3)
    feats = O.resnet34_features(sd, patch)
    pr = torch.sigmoid(O.decode_from_features(sd, cfg, feats, zz, honour_use_skip=False))
    pr = O._up_ac(pr, (256, 320))
    assert rel(pred, pr) <= TOL
    wr = O.blend_weight(256, 320, 1, 0, 2, 3, 50)
    assert (wt[0, 0].cpu() - wr).abs().max() < 1e-6


@pytest.mark.parametrize("kw", [dict(use_skip=False), dict(use_attention=False)], ids=["no_skip", "no_attention"])
def test_no_skip_and_no_attention_models(cuda_dev, kw):
    """The reference's --no-skip / attention-less models (use_skip=False used to crash in decode)."""
    from vaeunet_b200.utils.vae_utils import generate_predictions
    m, sd, cfg = _model(cuda_dev, **kw)
    x, _ = synthetic_fundus(2, 96, 96, seed=960)
    g = torch.Generator().manual_seed(961)
    eps = torch.randn(2, 32, generator=g)
    ref, mu_r, lv_r = O.unetresnet_forward(sd, cfg, x, eps)
    with torch.no_grad():
        out, mu, lv = m(x.to(cuda_dev), eps=eps.to(cuda_dev))
    assert out.shape == ref.shape and rel(out, ref) <= TOL, rel(out, ref)
    assert rel(mu, mu_r) <= TOL
    e3 = torch.randn(3, 2, 32, generator=g)
    ref_g = O.generate_predictions(sd, cfg, x, 1.0, 3, e3)
    got_g = generate_predictions(m, x.to(cuda_dev), 1.0, 3, eps=e3.to(cuda_dev))
    assert rel(got_g, ref_g) <= TOL, rel(got_g, ref_g)
    ref_mc = O.mc_patch_uncertainty(sd, cfg, x, 1.0, 3, e3)
    with torch.no_grad():
        res = m.mc_predict(x.to(cuda_dev), 1.0, 3, eps=e3.to(cuda_dev))
    assert rel(res["mean"], ref_mc["mean"]) <= TOL and rel(res["entropy"], ref_mc["entropy"]) <= TOL
