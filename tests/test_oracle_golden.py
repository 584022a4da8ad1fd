"""Pin the oracle (oracle/vaeunet_oracle.py) against golden vectors produced by the UNMODIFIED
reference (tests/golden/make_golden.py). CPU only. Tolerance: fp32 restatement of the same torch
ops -> 1e-4 relative (BASELINE.json's fp32 bar); most entries agree to ~1e-6."""
import math
from pathlib import Path

import pytest
import torch

from oracle import vaeunet_oracle as O
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200 import testing

G = Path(__file__).resolve().parent / "golden"


def rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.fixture(scope="module")
def gold():
    return torch.load(G / "golden_resnet.pt", map_location="cpu", weights_only=False)


@pytest.fixture(scope="module")
def sd():
    return testing.resnet_fixture()[0]


CFG = O.ResNetCfg()


def test_state_dict_layout(sd):
    assert len(sd) == 389
    assert sd["decoder_blocks.0.conv1.0.weight"].shape == (512, 800, 3, 3)
    assert sd["final_conv.weight"].shape == (1, 64, 1, 1)
    n = sum(v.numel() for k, v in sd.items() if not (k.endswith("running_mean") or k.endswith("running_var") or k.endswith("num_batches_tracked")))
    assert n == 30455789


@pytest.mark.parametrize("tag", ["fwd96", "fwd_odd"])
def test_forward(gold, sd, tag):
    g = gold[tag]
    B, H, W = g["shape"]
    x, _ = synthetic_fundus(B, H, W, seed=g["seed"])
    out, mu, lv = O.unetresnet_forward(sd, CFG, x, g["eps"])
    assert rel(out, g["logits"]) < 1e-4 and rel(mu, g["mu"]) < 1e-4 and rel(lv, g["logvar"]) < 1e-4
    assert out.std() > 1.0                                  # fixture logits span a real range (H1)
    assert ((out > 0) == (g["logits"] > 0)).float().mean() >= 0.999


def test_stages(gold, sd):
    g = gold["stages"]
    x, _ = synthetic_fundus(1, 96, 96, seed=g["seed"])
    feats = O.resnet34_features(sd, x)
    for f, r in zip(feats, g["feats"]):
        assert rel(f[:, ::4], r) < 1e-4
    mu, lv = O.unetresnet_encode(sd, x)
    assert rel(mu, g["mu"]) < 1e-4 and rel(lv, g["logvar"]) < 1e-4
    zs = mu.view(1, -1, 1, 1).expand(-1, -1, *feats[-1].shape[2:])
    x0 = torch.relu(O._bn(sd, "z_initial.1.", O._conv(sd, "z_initial.0.", zs), None))
    assert rel(x0, g["x0"]) < 1e-4
    d0 = O.decoder_block(sd, "decoder_blocks.0.", x0, feats[-2], zs, True, True, True)
    assert rel(d0, g["d0"]) < 1e-4
    att = O.attention_gate(sd, "decoder_blocks.1.attention.", O._up_ac(d0, feats[-3].shape[2:]), feats[-3])
    assert rel(att, g["att1"]) < 1e-4


@pytest.mark.parametrize("tag", ["gen_T1_N3", "gen_T2_N2"])
def test_generate_predictions(gold, sd, tag):
    g = gold[tag]
    x, _ = synthetic_fundus(2, 96, 96, seed=g["seed"])
    pred = O.generate_predictions(sd, CFG, x, g["T"], g["N"], g["eps"])
    assert pred.shape == g["pred"].shape == (2, 1, 48, 48)   # mean of logits at H/2 (vae_utils.py:76)
    assert rel(pred, g["pred"]) < 1e-4


def test_full_image_mc(gold, sd):
    g = gold["full_image_mc"]
    x, _ = synthetic_fundus(1, 96, 96, seed=g["seed"])
    segs, mu, lv = O.segmentation_distribution(sd, CFG, x, 4, 1.0, g["eps"])
    assert rel(segs, g["segs"]) < 1e-4
    um = O.calculate_uncertainty_metrics(segs.unsqueeze(1))
    for k in ("mean", "std", "entropy", "mutual_info"):
        assert (um[k] - g["um_" + k]).abs().max() < 1e-4, k
    res = O.mc_patch_uncertainty(sd, CFG, x, 1.0, 4, g["eps"])
    assert rel(res["mean"], g["um_mean"]) < 1e-4 and (res["std"] - g["um_std"]).abs().max() < 1e-4


def test_predict_with_patches(gold, sd):
    g = gold["patches"]
    H, W = g["shape"]
    x, _ = synthetic_fundus(1, H, W, seed=g["seed"])
    mu, _ = O.unetresnet_encode(sd, x)
    assert rel(mu, g["mu"]) < 1e-4
    out = O.predict_with_patches(sd, CFG, x, mu, g["patch"], g["overlap"])
    assert (out - g["out"]).abs().max() < 1e-4
    nh, nw, info = O.patch_grid(2848, 4288, 512, 100)
    assert (nh, nw) == (7, 11) and info[-1] == (2848 - 512, 2848, 4288 - 512, 4288)


def test_decode(gold, sd):
    g = gold["decode"]
    dec = O.unetresnet_decode(sd, CFG, g["z"])
    assert rel(dec[:, :, ::8, ::8], g["out"]) < 1e-4
    assert abs(dec.sum().item() - g["out_sum"].item()) < 1e-3 * abs(g["out_sum"].item()) + 1e-2
    dec_r = O.unetresnet_decode(sd, CFG, g["z"][:1], (40, 48))
    assert rel(dec_r, g["out_resized"]) < 1e-4


@pytest.mark.parametrize("lesion", ["EX", "MA"])
def test_training_step(gold, sd, lesion):
    g = gold["train_" + lesion]
    x, m = synthetic_fundus(2, 64, 64, seed=g["seed"], lesion=lesion)
    loss, recon, kl, grads, updates = O.training_grads(sd, CFG, x, m, g["eps"], 0.001, 1e-3, lesion)
    assert abs(loss.item() - g["loss"].item()) < 1e-5 and abs(kl.item() - g["kl"].item()) < 1e-4
    assert len(grads) == len(g["grads"])
    worst = 0.0
    for k, ref in g["grads"].items():
        gn = grads[k].norm().item()
        assert abs(gn - ref["norm"].item()) <= 2e-3 * ref["norm"].item() + 1e-7, k
        if "full" in ref:
            mine = grads[k] if grads[k].numel() <= 300000 else grads[k].flatten()[:4096]
            worst = max(worst, rel(mine, ref["full"]))
    assert worst < 2e-3
    for k, v in g["new_running_mean"].items():
        assert rel(updates[k], v) < 1e-4, k


def test_injection_modes():
    g = torch.load(G / "golden_resnet_modes.pt", map_location="cpu", weights_only=False)
    for mode, r in g.items():
        sd, kw = testing.resnet_fixture(latent_injection=mode)
        x, _ = synthetic_fundus(1, 64, 64, seed=r["seed"])
        out, mu, lv = O.unetresnet_forward(sd, O.ResNetCfg(latent_injection=mode), x, r["eps"])
        assert rel(out, r["logits"]) < 1e-4, mode


def test_basic_unet():
    g = torch.load(G / "golden_unet.pt", map_location="cpu", weights_only=False)
    for tag, bilinear in (("convT", False), ("bilinear", True)):
        sd = testing.unet_fixture(1, bilinear)
        assert len(sd) == (194 if bilinear else 202)
        r = g[tag]
        x, _ = synthetic_fundus(*r["shape"], seed=r["seed"])
        out = O.unet_forward(sd, x, bilinear)
        assert rel(out, r["logits"]) < 1e-4
        x1 = O.double_conv(sd, "inc.", x)
        x2 = O.double_conv(sd, "down1.maxpool_conv.1.", torch.nn.functional.max_pool2d(x1, 2))
        assert rel(x2[:, ::8], r["x2"]) < 1e-4
        xin = torch.relu(torch.randn(2, 64 if bilinear else 128, 32, 40, generator=torch.Generator().manual_seed(5)))
        assert rel(O.unet_up(sd, "up4.", xin, x1, bilinear)[:, ::8], r["up4"]) < 1e-4


def test_losses_metrics_misc():
    g = torch.load(G / "golden_misc.pt", map_location="cpu", weights_only=False)
    L = g["loss"]
    assert abs(O.dice_loss(L["x"], L["t"]).item() - L["dice"].item()) < 1e-6
    assert abs(O.combined_loss(L["x"], L["t"]).item() - L["combined"].item()) < 1e-6
    assert abs(O.ma_focal_loss(L["x"], L["t"], 0.9).item() - L["focal"].item()) < 1e-6
    assert abs(O.ma_segmentation_loss(L["x"], L["t"]).item() - L["ma"].item()) < 1e-6
    K = g["kl"]
    assert abs(O.kl_with_free_bits(K["mu"], K["lv"], 1e-3).item() - K["kl_1e3"].item()) < 1e-5
    assert abs(O.kl_with_free_bits(K["mu"], K["lv"], 0.0).item() - K["kl_0"].item()) < 1e-5
    assert abs(O.kl_with_free_bits(K["mu"], K["lv"], 0.05).item() - K["kl_big"].item()) < 1e-5
    st = O.latent_stats(K["mu"], K["lv"])
    for k, v in K["stats"].items():
        assert abs(st[k] - v) < 1e-5, k
    for strat, vals in g["anneal"].items():
        mine = [O.kl_anneal_weight(e, 0.0, 0.001, 10, strat) for e in range(14)] + \
               [O.kl_anneal_weight(3, 0.0, 0.001, 10, strat, 5, 20)]
        assert all(math.isclose(a, b, rel_tol=1e-9, abs_tol=1e-12) for a, b in zip(mine, vals)), strat
    M = g["metrics"]
    ece, accs, confs, counts = O.expected_calibration_error(M["p"], M["gt"])
    assert abs(ece - M["ece"]) < 1e-7 and torch.equal(torch.tensor(counts), M["counts"])
    assert abs(O.brier_score(M["p"], M["gt"]) - M["brier"]) < 1e-7
    nll = O.negative_log_likelihood(M["p"], M["gt"])
    # reference quirk: clamp(p, 1e-9, 1 - 1e-9) is a no-op at 1.0 in fp32 -> inf when p == 1, t == 0
    assert (math.isinf(nll) and math.isinf(M["nll"])) or abs(nll - M["nll"]) < 1e-6
    pc = M["p"].clamp(1e-3, 1 - 1e-3)
    assert abs(O.negative_log_likelihood(pc, M["gt"]) -
               torch.nn.functional.binary_cross_entropy(pc, M["gt"]).item()) < 1e-6
    assert abs(O.uncertainty_error_dice(M["unc"], (M["p"] > 0.5).float(), M["gt"], 0.2) - M["ue_dice"]) < 1e-6
    am = O.threshold_metrics(M["p"], M["gt"])
    for k, v in M["all_metrics"].items():
        assert abs(am[k] - v) < 1e-6, k
    U = g["uncertainty"]
    um = O.calculate_uncertainty_metrics(U["segs"])
    for k in ("mean", "std", "entropy", "mutual_info", "coeff_var"):
        assert torch.allclose(um[k], U[k], rtol=1e-5, atol=1e-6), k
    S = g["sample"]
    assert torch.allclose(O.sample_from_latent(K["mu"], K["lv"], 1.7, S["eps"]), S["z"], rtol=1e-6, atol=1e-6)


# ---------------------------------------------------------------------------------------------------------
# evaluate() and the temperature ensemble, pinned by tests/golden/make_golden_next.py
# ---------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gold_next():
    return torch.load(G / "golden_next.pt", map_location="cpu", weights_only=False)


def test_evaluate(gold_next, sd):
    g = gold_next["evaluate"]
    B, H, W = g["shape"]
    batches = [synthetic_fundus(B, H, W, seed=s) for s in g["seeds"]]
    avg, kept = O.evaluate_metrics(sd, CFG, batches, g["eps"])
    for k, v in g["avg"].items():
        assert abs(avg[k] - v) <= 1e-6 * max(1.0, abs(v)), (k, avg[k], v)
    avg2, _ = O.evaluate_metrics(sd, CFG, batches, g["eps"], max_samples=2)
    for k, v in g["avg_max2"].items():
        assert abs(avg2[k] - v) <= 1e-6 * max(1.0, abs(v)), (k, avg2[k], v)
    assert (kept[0][0] - g["sample0_pred"]).abs().max() < 1e-5


def test_temperature_ensemble(gold_next, sd):
    g = gold_next["ensemble"]
    H, W = g["shape"]
    x, _ = synthetic_fundus(1, H, W, seed=g["seed"])
    out = O.generate_ensemble_prediction(sd, CFG, x, g["temps"], g["samples_per_temp"], g["eps"], True, g["patch"], g["overlap"])
    assert (out - g["out"]).abs().max() < 1e-5
    out_u = O.generate_ensemble_prediction(sd, CFG, x, g["temps"], g["samples_per_temp"], g["eps"], False, g["patch"], g["overlap"])
    assert (out_u - g["out_unweighted"]).abs().max() < 1e-5


def test_sort_based_metrics_f3():
    """the oracle's numpy restatements of calculate_sparsification_metrics / calculate_uncertainty_error_auc
    (incl. its own restatement of sklearn's ROC-AUC / average precision) against the reference's outputs."""
    import math
    from vaeunet_b200.testing import f3_inputs
    F3 = torch.load(G / "golden_f3.pt", weights_only=False)
    for tag in ("sp_small", "sp_one"):
        g = F3[tag]
        B, H, W = g["shape"]
        p, gt, unc = f3_inputs(g["seed"], B, H, W, False)
        fr, er, eu = O.sparsification_metrics(p, unc, gt, 20)
        assert abs(fr - g["frac"].numpy()).max() == 0
        assert abs(er - g["rand"].numpy()).max() < 1e-12 and abs(eu - g["unc"].numpy()).max() < 1e-12
    g = F3["sp_degenerate"]
    p, gt, unc = f3_inputs(13, 2, 20, 24, False)
    p[1] = 0.5; gt[1] = 1.0
    _, er, eu = O.sparsification_metrics(p, unc, gt, 10)
    assert abs(er - g["rand"].numpy()).max() < 1e-12 and abs(eu - g["unc"].numpy()).max() < 1e-12
    for tag in ("auc_cont", "auc_ties"):
        g = F3[tag]
        H, W = g["shape"]
        p, gt, unc = f3_inputs(g["seed"], 1, H, W, g["ties"])
        a, b = O.uncertainty_error_auc(p[0], gt[0], unc[0])
        assert abs(a - g["auroc"]) < 1e-12 and abs(b - g["auprc"]) < 1e-12
    p, gt, unc = f3_inputs(23, 1, 16, 16, False)
    a, b = O.uncertainty_error_auc(p[0], (p[0] > 0.5).float(), unc[0])
    assert math.isnan(a) and math.isnan(b) and math.isnan(F3["auc_one_class"]["auroc"])


def test_patch_pipeline_f4():
    """the oracle's restatement of Pillow's 8-bit resampling + the reference's patch loop against the golden
    hashes from the unmodified reference (IDRIDDataset.preprocess / is_valid_patch), and directly against PIL when present."""
    import hashlib
    import numpy as np
    from vaeunet_b200.testing import f4_inputs
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    F4 = torch.load(G / "golden_f4.pt", weights_only=False)
    for tag, g in F4.items():
        img, mask = f4_inputs(g["seed"], *g["shape"])
        imgs, masks, coords, lesion = O.patch_pipeline(img, mask, g["patch"], g["scale"], g["split"])
        assert coords == g["coords"] and lesion == g["lesion"], tag
        hs = [sha(imgs[i].contiguous().numpy())[:16] + sha(masks[i].contiguous().numpy())[:16] for i in range(len(coords))]
        assert hs == g["hashes"], tag                       # bit-exact patches
    try:
        from PIL import Image
    except ImportError:
        return
    rng = np.random.default_rng(5)
    for (H, W, nh, nw) in [(97, 131, 24, 32), (50, 70, 75, 105), (233, 177, 58, 44)]:
        a = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        assert np.array_equal(O.pil_resize_u8(a, nw, nh, False), np.array(Image.fromarray(a).resize((nw, nh), resample=Image.BICUBIC)))
        m = (rng.integers(0, 2, (H, W), dtype=np.uint8) * 255)
        assert np.array_equal(O.pil_resize_u8(m, nw, nh, True), np.array(Image.fromarray(m).resize((nw, nh), resample=Image.NEAREST)))
