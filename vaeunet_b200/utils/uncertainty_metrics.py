"""Drop-in for the numeric functions of the reference's utils/uncertainty_metrics.py.
ECE / Brier / NLL / uncertainty-error Dice are single-pass device reductions. The sort-based
sparsification curve and AUROC/AUPRC stay on the host through numpy/sklearn exactly as the
reference does (SURVEY.md 8f row 3 — a "next" row); plotting helpers are out of scope."""
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
    """utils/uncertainty_metrics.py:270-283 — host/sklearn, as in the reference."""
    from sklearn.metrics import average_precision_score, roc_auc_score
    pred_binary = (mean_pred > 0.5).float().view(-1).cpu().numpy()
    gt_flat = gt_mask.view(-1).cpu().numpy()
    uncertainty = uncertainty_map.view(-1).cpu().numpy()
    errors = (pred_binary != gt_flat).astype(np.int32)
    try:
        return roc_auc_score(errors, uncertainty), average_precision_score(errors, uncertainty)
    except ValueError:
        return float('nan'), float('nan')
