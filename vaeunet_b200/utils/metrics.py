"""Drop-in for utils/metrics.get_all_metrics (utils/metrics.py:98-117): one fused confusion-count
reduction instead of six thresholded passes. Inputs are thresholded at 0.5 as given (the reference
calls this on raw logits, evaluate.py:76 — the quirk is the caller's and is preserved)."""
from __future__ import annotations

from typing import Dict

from .. import ops


def get_all_metrics(pred, target, epsilon: float = 1e-6) -> Dict[str, float]:
    tp, fp, tn, fn = [v.item() for v in ops.confusion_counts(pred.float(), target.float(), 0.5).cpu()]
    den = (tp + fp) + (tp + fn)
    dice = 1.0 if den == 0 else (2.0 * tp + epsilon) / (den + epsilon)
    return {'dice': dice, 'iou': (tp + epsilon) / (den - tp + epsilon),
            'precision': (tp + epsilon) / (tp + fp + epsilon), 'recall': (tp + epsilon) / (tp + fn + epsilon),
            'specificity': (tn + epsilon) / (tn + fp + epsilon), 'accuracy': (tp + tn) / (tp + fp + tn + fn)}
