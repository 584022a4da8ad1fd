"""Drop-in for the reference's utils/loss.py on fused sm_100a reduction kernels.
Each loss is ONE pass over logits+targets (sigmoid, BCE, dice sums, focal term) and, in backward,
one pass that emits d(loss)/d(logits)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops


class _SegLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inputs, targets, w_bce, w_dice, w_focal, alpha, gamma, smooth):
        x = inputs.detach().float().contiguous()
        t = targets.detach().float().contiguous()
        sums = ops.seg_loss_sums(x, t, False, alpha, gamma)
        n = float(x.numel())
        half = smooth / 2.0
        P = torch.clamp(sums[0], min=half)
        T = torch.clamp(sums[1], min=half)
        dice = 1.0 - (2.0 * sums[2] + smooth) / (P + T + smooth)
        loss = w_bce * sums[3] / n + w_dice * dice + w_focal * sums[4] / n
        ctx.save_for_backward(x, t, sums)
        ctx.cfg = (w_bce, w_dice, w_focal, alpha, gamma, smooth, inputs.dtype, inputs.shape)
        return loss.float()

    @staticmethod
    def backward(ctx, gout):
        x, t, sums = ctx.saved_tensors
        w_bce, w_dice, w_focal, alpha, gamma, smooth, dt, shp = ctx.cfg
        dx = ops.seg_loss_bwd(x, t, sums, w_bce, w_dice, w_focal, alpha, gamma, smooth, 1.0)
        return (dx * gout).reshape(shp).to(dt), None, None, None, None, None, None, None


def dice_loss(inputs, targets, smooth=1.0):
    """utils/loss.py:6-28."""
    return _SegLossFn.apply(inputs, targets, 0.0, 1.0, 0.0, 0.8, 2.0, float(smooth))


class DiceLoss(nn.Module):
    def __init__(self, smooth=1.0):
        super().__init__()
        self.smooth = smooth

    def forward(self, inputs, targets):
        return dice_loss(inputs, targets, self.smooth)


class CombinedLoss(nn.Module):
    """utils/loss.py:44-63: bce_weight * BCEWithLogits(mean) + dice_weight * dice."""

    def __init__(self, bce_weight=0.5, dice_weight=0.5):
        super().__init__()
        self.bce_weight = bce_weight
        self.dice_weight = dice_weight

    def forward(self, inputs, targets):
        return _SegLossFn.apply(inputs, targets, self.bce_weight, self.dice_weight, 0.0, 0.8, 2.0, 1.0)


class MAFocalLoss(nn.Module):
    """utils/loss.py:66-92."""

    def __init__(self, alpha=0.8, gamma=2.0):
        super().__init__()
        self.alpha = alpha
        self.gamma = gamma
        self.eps = 1e-6

    def forward(self, inputs, targets):
        return _SegLossFn.apply(inputs, targets, 0.0, 0.0, 1.0, self.alpha, self.gamma, 1.0)


class MASegmentationLoss(nn.Module):
    """utils/loss.py:95-111."""

    def __init__(self, dice_weight=0.5, focal_weight=0.5, focal_gamma=2.0, class_weight=0.9):
        super().__init__()
        self.dice_weight = dice_weight
        self.focal_weight = focal_weight
        self.focal_loss = MAFocalLoss(alpha=class_weight, gamma=focal_gamma)
        self.dice_loss = DiceLoss()

    def forward(self, inputs, targets):
        return _SegLossFn.apply(inputs, targets, 0.0, self.dice_weight, self.focal_weight, self.focal_loss.alpha,
                                self.focal_loss.gamma, 1.0)


class KLAnnealer:
    """utils/loss.py:114-145 (host-side schedule)."""

    def __init__(self, kl_start=0.0, kl_end=1.0, warmup_epochs=10, strategy='linear'):
        self.kl_start = kl_start
        self.kl_end = kl_end
        self.warmup_epochs = warmup_epochs
        self.strategy = strategy

    def get_weight(self, epoch, batch=None, num_batches=None):
        if self.strategy == 'constant':
            return self.kl_end
        if batch is not None and num_batches is not None:
            progress = (epoch + batch / num_batches) / self.warmup_epochs
        else:
            progress = epoch / self.warmup_epochs
        progress = min(progress, 1.0)
        if self.strategy == 'linear':
            return self.kl_start + progress * (self.kl_end - self.kl_start)
        elif self.strategy == 'cyclical':
            return self.kl_start + (progress % 1.0) * (self.kl_end - self.kl_start)
        return self.kl_end


class _KLFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mu, logvar, free_bits):
        kl, dmu, dlv = ops.kl_free_bits(mu.detach().float(), logvar.detach().float(), free_bits, want_grad=True)
        ctx.save_for_backward(dmu, dlv)
        return kl

    @staticmethod
    def backward(ctx, gout):
        dmu, dlv = ctx.saved_tensors
        return dmu * gout, dlv * gout, None


def kl_with_free_bits(mu, logvar, free_bits=1e-4):
    """utils/loss.py:148-170: per-element free bits, sum over latent dims, mean over batch."""
    return _KLFn.apply(mu, logvar, float(free_bits))
