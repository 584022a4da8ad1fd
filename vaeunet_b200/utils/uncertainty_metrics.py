"""Drop-in for the numeric functions of the reference's utils/uncertainty_metrics.py.
ECE / Brier / NLL / uncertainty-error Dice are single-pass device reductions; the sort-based sparsification curve and
the uncertainty-vs-error AUROC / AUPRC run a device radix sort + prefix sums (csrc/vu_sort_metrics.cu)
instead of numpy argsort / sklearn on the host. Plotting helpers are out of scope."""
from __future__ import annotations

import numpy as np
import torch

from .. import ops


def calculate_expected_calibration_error(pred_probs, ground_truth, num_bins=10):
    """utils/uncertainty_metrics.py:108-135 -> (ece, bin_accs, bin_confs, bin_counts)."""
    h = ops.ece_hist(pred_probs.float(), ground_truth.float(), num_bins).cpu()
    counts, sgt, sp = h[0], h[1], h[2]
    nz = counts > 0
    accs = torch.where(nz, sgt / counts.clamp(min=1), torch.zeros_like(sgt))
    confs = torch.where(nz, sp / counts.clamp(min=1), torch.zeros_like(sp))
    ece = ((counts * (accs - confs).abs()).sum() / counts.sum()).item()
    return ece, accs.float().numpy(), confs.float().numpy(), counts.float().numpy()


def brier_score(pred_probs, ground_truth):
    """utils/uncertainty_metrics.py:137-139."""
    s = ops.seg_loss_sums(pred_probs.float(), ground_truth.float(), True).cpu()
    return (s[6] / s[5]).item()


def calculate_negative_log_likelihood(predictions, targets, epsilon=1e-9):
    """utils/uncertainty_metrics.py:334-338 (epsilon fixed at the reference default 1e-9)."""
    if epsilon != 1e-9:
        raise ValueError("the fused kernel implements the reference default epsilon=1e-9")
    s = ops.seg_loss_sums(predictions.float(), targets.float(), True).cpu()
    return (s[7] / s[5]).item()


def calculate_uncertainty_error_dice(uncertainty_map, predictions_binary, targets, uncertainty_threshold=0.2):
    """utils/uncertainty_metrics.py:340-353: Dice between (uncertainty > thr) and (pred != target)."""
    err = (predictions_binary != targets).float()     # mask construction: host-side glue on a bool map
    c = ops.confusion_counts(uncertainty_map.float(), err, float(uncertainty_threshold)).cpu()
    tp, fp, tn, fn = [v.item() for v in c]
    inter, den = tp, (tp + fp) + (tp + fn)
    if den == 0:
        return 1.0 if inter == 0 else 0.0
    return float((2.0 * inter) / (den + 1e-8))


def calculate_uncertainty_error_auc(mean_pred, gt_mask, uncertainty_map):
    """utils/uncertainty_metrics.py:270-283 -> (auroc, auprc) of "uncertainty predicts the segmentation error".
    errors = ((mean_pred > 0.5) != gt); both scores follow sklearn's definitions (ties grouped: the curve has one point per
    distinct uncertainty value; AUROC by trapezoids, average precision as the step-function sum). One device sort of the
    pixels by uncertainty, one prefix scan, one host read of four numbers. NaN when only one class is present (the
    reference catches sklearn's ValueError)."""
    labels = ops.error_labels(mean_pred.reshape(-1), gt_mask.reshape(-1), 0.5)
    keys, lab_sorted, ws = ops.sort_pairs_desc(uncertainty_map.reshape(-1).float(), labels)
    a, ap, pos, neg = [float(v) for v in ops.binary_curve_auc(keys, lab_sorted, ws).cpu()]
    if pos == 0.0 or neg == 0.0:
        return float('nan'), float('nan')
    return a / (pos * neg), ap / pos


def calculate_sparsification_metrics(pred_probs, uncertainties, ground_truth, num_points=20):
    """utils/uncertainty_metrics.py:175-240 -> (fraction_removed, errors_random, errors_uncertainty) numpy arrays.
    Per image: pixel NLL errors, pixels sorted by uncertainty (descending) on the device, the error of the kept pixels at
    every removal fraction from sums of the sorted errors between the cut points. The random-removal baseline is DEFINED by
    numpy's RNG stream (np.random.seed(i + 1000 b); np.random.choice without replacement): the subsets are drawn on the host
    exactly as the reference draws them and cross PCIe as packed bit masks (N / 8 bytes each); the sums run on the device."""
    pred_probs = pred_probs.float()
    B = pred_probs.shape[0]
    fraction_removed = np.linspace(0, 0.99, num_points)
    errors_random = np.zeros(num_points)
    errors_uncertainty = np.zeros(num_points)
    p2 = pred_probs.reshape(B, -1)
    u2 = uncertainties.float().reshape(B, -1)
    g2 = ground_truth.float().reshape(B, -1)
    N = p2.shape[1]
    rem = [int(N * f) for f in fraction_removed]
    cuts = torch.tensor(rem + [N], dtype=torch.int64, device=p2.device)
    ws = None
    for b in range(B):
        err, stats = ops.pixel_nll_errors(p2[b], g2[b], 1e-7)
        total, nans, lo, hi = ops.decode_nll_stats(stats)
        if nans > 0 or lo == hi:                       # np.isnan(...).any() / all errors equal (uncertainty_metrics.py:194)
            continue
        initial_error = total / N
        if initial_error <= 0 or initial_error != initial_error:
            continue
        _, err_sorted, ws = ops.sort_pairs_desc(u2[b], err.view(torch.int32), ws)
        bins = ops.sorted_bin_sums(err_sorted, cuts).cpu().numpy()          # sums over [rem_k, rem_{k+1})
        removed = np.concatenate([[0.0], np.cumsum(bins)])[:num_points]     # sum of the rem_k most uncertain pixels
        masks = np.zeros((num_points, (N + 7) // 8), dtype=np.uint8)
        live = []
        for i, frac in enumerate(fraction_removed):
            keep = N - rem[i]
            if frac >= 1.0 or keep <= 0:
                errors_random[i] += 1.0
                errors_uncertainty[i] += 1.0
                continue
            np.random.seed(i + b * 1000)
            idx = np.random.choice(np.arange(N), keep, replace=False)
            m = np.zeros(N, dtype=bool)
            m[idx] = True
            masks[i] = np.packbits(m, bitorder="little")
            live.append(i)
            errors_uncertainty[i] += ((total - removed[i]) / keep) / initial_error
        dm = torch.from_numpy(masks).to(p2.device)
        sums = torch.cat([ops.masked_sum(err, dm[i]) for i in live]).cpu().numpy() if live else []
        for j, i in enumerate(live):
            errors_random[i] += (sums[j] / (N - rem[i])) / initial_error
    errors_random /= B
    errors_uncertainty /= B
    if errors_random[0] > 0:
        errors_random = errors_random / errors_random[0]
    if errors_uncertainty[0] > 0:
        errors_uncertainty = errors_uncertainty / errors_uncertainty[0]
    for i in range(1, num_points):
        if errors_uncertainty[i] > errors_uncertainty[i - 1]:
            errors_uncertainty[i] = errors_uncertainty[i - 1]
    errors_random = np.nan_to_num(errors_random, nan=1.0, posinf=1.0, neginf=0.0)
    errors_uncertainty = np.nan_to_num(errors_uncertainty, nan=1.0, posinf=1.0, neginf=0.0)
    return fraction_removed, errors_random, errors_uncertainty
