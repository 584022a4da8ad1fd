"""NVTX ranges around the phases of the hot path (encoder / latent heads / decoder / reduction; forward / loss / backward /
optimizer), for Nsight timelines. Off by default (no host cost); `VU_NVTX=1` turns them on."""
from __future__ import annotations

import contextlib
import os

ENABLED = os.environ.get("VU_NVTX", "0") == "1"

_NULL = contextlib.nullcontext()


class _Range:
    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        import torch
        torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        import torch
        torch.cuda.nvtx.range_pop()
        return False


def rng(name: str):
    """`with rng("vaeunet.encode"):` - an NVTX range when VU_NVTX=1, a no-op context otherwise."""
    return _Range(name) if ENABLED else _NULL
