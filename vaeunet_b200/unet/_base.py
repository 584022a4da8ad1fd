"""Shared plumbing of the drop-in nn.Modules: parameters live in ordinary nn.Conv2d/nn.BatchNorm2d
containers (so `state_dict()` keys match the reference exactly) while `forward` runs this repo's
CUDA kernels through a lazily (re)built packed-weight unit."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops
from .._lib import VaeUnetError


class KernelModule(nn.Module):
    """Mixin: caches a packed-weight unit, rebuilt when any parameter/buffer changes."""

    # 16-bit storage / tensor-core operand type of the kernels (accumulation is always fp32).
    # Default: ops.DEFAULT_DTYPE = f16, the reference's own AMP dtype (train.py autocast + GradScaler):
    # 3 more mantissa bits than bf16 at the same tcgen05 kind::f16 rate, and the only 16-bit type that
    # meets BASELINE's 1e-2 bar on every output (DESIGN.md "Numerics"). "bf16" is selectable; setting
    # compute_dtype on a model propagates to every kernel sub-module.
    @property
    def compute_dtype(self):
        return self.__dict__.get("_compute_dtype", ops.DEFAULT_DTYPE)

    @compute_dtype.setter
    def compute_dtype(self, dt):
        if isinstance(dt, str):
            dt = {"bf16": ops.BF16, "f16": ops.F16, "fp16": ops.F16}[dt]
        for m in self.modules():
            if isinstance(m, KernelModule):
                m.__dict__["_compute_dtype"] = dt

    def _param_version(self):
        dev = None
        ver = 0
        for t in list(self.parameters()) + list(self.buffers()):
            ver += t._version
            dev = t.device
        return (ver, str(dev), self.compute_dtype)

    def _build_unit(self, sd, dev):  # pragma: no cover - abstract
        raise NotImplementedError

    def _unit(self):
        key = self._param_version()
        cached = self.__dict__.get("_unit_cache")
        if cached is not None and cached[0] == key:
            return cached[1]
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise VaeUnetError(
                f"{type(self).__name__}: parameters are on {dev}; vaeunet_b200 runs on a B200 only "
                "(no CPU fallback). Move the model with .cuda().")
        unit = self._build_unit({k: v for k, v in self.state_dict().items()}, dev)
        self.__dict__["_unit_cache"] = (key, unit)
        return unit

    def _require_eval(self):
        if self.training:
            raise VaeUnetError(
                f"{type(self).__name__}.forward in train() mode: sub-modules run the eval-mode (folded BatchNorm) "
                "kernels. Train through the top-level model (UNetResNet / UNet .forward in train() mode uses the "
                "torch.ops.vaeunet.* autograd path) or vaeunet_b200.training.TrainStep.")


def to_act(x: torch.Tensor, dtype=None) -> ops.Act:
    """NCHW float tensor on the device -> NHWC 16-bit activation."""
    if not x.is_cuda:
        raise VaeUnetError("input tensor is not on a CUDA device (no CPU fallback)")
    return ops.nchw_to_nhwc(x.float(), dtype=ops.DEFAULT_DTYPE if dtype is None else dtype)


def from_act(a: ops.Act) -> torch.Tensor:
    return ops.nhwc_to_nchw(a)
