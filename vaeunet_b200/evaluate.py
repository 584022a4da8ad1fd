"""Drop-in for evaluate.evaluate (evaluate.py:19-101): the validation pass that runs between training epochs.

Same signature, return value and quirks as the reference (metrics of `get_all_metrics` on the RAW logits thresholded at
0.5, :76; sum over batches divided by the number of batches, :97; `max_samples` caps the batch count, :27-34; up to four
(image, sigmoid(prediction), mask, batch metrics) samples kept for plotting, :79-88). What changes is where the work runs:
the forward is the tcgen05 inference engine, the six thresholded passes of `get_all_metrics` are ONE fused confusion-count
reduction per batch (`vu_confusion_counts`), and the per-batch metric arithmetic stays on the device - the host reads the
averaged metrics once at the end instead of synchronising six times per batch.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from . import ops

_KEYS = ("dice", "iou", "precision", "recall", "specificity", "accuracy")


def batch_metrics_device(counts: torch.Tensor, epsilon: float = 1e-6) -> torch.Tensor:
    """(tp, fp, tn, fn) fp32/fp64 [4] -> the six metrics of utils/metrics.py:98-117 as a device tensor [6] (order _KEYS)."""
    c = counts.double()
    tp, fp, tn, fn = c[0], c[1], c[2], c[3]
    den = (tp + fp) + (tp + fn)
    dice = torch.where(den == 0, torch.ones_like(den), (2.0 * tp + epsilon) / (den + epsilon))     # metrics.py:13-14
    return torch.stack([dice, (tp + epsilon) / (den - tp + epsilon), (tp + epsilon) / (tp + fp + epsilon),
                        (tp + epsilon) / (tp + fn + epsilon), (tn + epsilon) / (tn + fp + epsilon),
                        (tp + tn) / (tp + fp + tn + fn)])


@torch.inference_mode()
def evaluate(model, dataloader, device, amp=False, max_samples: Optional[int] = None,
             eps: Optional[Sequence[torch.Tensor]] = None):
    """-> (metrics_avg: dict of floats, samples: list of (img_np, pred_np, mask_np, batch_metrics)).
    `amp` is accepted for signature compatibility (the engine computes in its own 16-bit storage type);
    `eps` (additive): per-batch reparameterisation noise for reproducible runs."""
    model.eval()
    device = torch.device(device)
    num_val_batches = len(dataloader)
    if max_samples is not None and max_samples < num_val_batches:
        num_val_batches = max_samples
    total = torch.zeros(6, dtype=torch.float64, device=device)
    samples: List = []
    for i, batch in enumerate(dataloader):
        if i >= num_val_batches:
            break
        if isinstance(batch, dict):
            images, masks = batch["image"], batch["mask"]
        else:
            images, masks = batch
        if isinstance(images, list):                       # ragged batch: the reference skips it (:41-43) but still divides
            continue                                       # by the full batch count
        images = images.to(device=device, non_blocking=True)
        masks = masks.to(device=device, non_blocking=True)
        # normalize() of the reference is Normalize(mean 0, std 1): the identity (:15-16,52-53)
        outputs = model(images) if eps is None else model(images, eps=eps[i].to(device))
        mask_pred = outputs[0] if isinstance(outputs, tuple) else outputs
        if mask_pred.shape != masks.shape:
            mask_pred = F.interpolate(mask_pred, size=masks.shape[2:], mode="bilinear", align_corners=False)
        m = batch_metrics_device(ops.confusion_counts(mask_pred.float().contiguous(), masks.float().contiguous(), 0.5))
        total += m
        if len(samples) < 4:
            bm = dict(zip(_KEYS, m.tolist()))
            pred = torch.sigmoid(mask_pred)
            for j in range(min(2, images.shape[0])):
                samples.append((images[j].cpu().numpy(), pred[j].cpu().numpy(), masks[j].cpu().numpy(), bm))
    avg = (total / num_val_batches).tolist()
    return dict(zip(_KEYS, avg)), samples
