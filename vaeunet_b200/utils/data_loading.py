"""Device side of the reference's patch pipeline (utils/data_loading.py:302-446 precompute_all_patches, :580-601 preprocess;
SURVEY 8f-4): once the JPEG / TIFF is decoded (host, PIL - out of scope), the 8-bit arrays go to the B200 and stay there:

    resize (Pillow-exact BICUBIC for images, NEAREST for masks) -> /255 or (> 0) -> 50 %-overlap tiling into NCHW fp32 patch
    batches -> per-patch black-border ratio and lesion flag

The reference does all of this with PIL / numpy / torch on the host and writes one `torch.save` file per patch; here the
result is a device-resident batch that feeds the network (training) or `predict.blended_samples` (inference) without host
staging. Resampling reproduces Pillow's ImagingResample bit for bit: the coefficient tables are built on the host with
Pillow's own formulas in double precision (O(W + H) numbers) and the two 8-bit passes run as kernels
(csrc/vu_tiles.cu).

Pillow (`pillow~=10.3.0`, requirements.txt:8) is a third-party dependency of the reference that is not vendored; the
algorithm restated here is src/libImaging/Resample.c (precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc
/ Vertical_8bpc) and Geometry.c (ImagingScaleAffine, nearest) of that release, and tests pin it against PIL.Image.resize."""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import ops

PRECISION_BITS = 32 - 8 - 2


def _bicubic(x: float) -> float:
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def resample_tables(in_size: int, out_size: int, support: float = 2.0, filt=_bicubic) -> Tuple[np.ndarray, np.ndarray]:
    """Pillow's precompute_coeffs + normalize_coeffs_8bpc for one axis (box = the whole axis).
    -> (bounds int32 [out, 2] = (first input index, taps), kk int32 [out, ksize] in 22-bit fixed point)."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    sup = support * filterscale
    ksize = int(math.ceil(sup)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - sup + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + sup + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [filt((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        if ww != 0.0:
            w = [v / ww for v in w]
        for x, v in enumerate(w):
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def nearest_table(in_size: int, out_size: int) -> np.ndarray:
    """Source index per output index of Image.resize(..., NEAREST) (Geometry.c ImagingScaleAffine: the coordinate
    advances by repeated addition in double precision, starting half a step in)."""
    a = in_size / out_size
    xo = a * 0.5
    tab = np.zeros(out_size, dtype=np.int32)
    for x in range(out_size):
        xin = -1 if xo < 0.0 else int(xo)
        tab[x] = xin if 0 <= xin < in_size else -1
        xo += a
    return tab


def resize_u8(img_u8_hwc: torch.Tensor, new_w: int, new_h: int, is_mask: bool) -> torch.Tensor:
    """PIL.Image.resize((new_w, new_h), NEAREST if is_mask else BICUBIC) of a uint8 HWC device tensor."""
    assert img_u8_hwc.dtype == torch.uint8 and img_u8_hwc.dim() == 3 and img_u8_hwc.is_cuda
    x = img_u8_hwc.contiguous()
    H, W, _ = x.shape
    dev = x.device
    if is_mask:
        yi = torch.from_numpy(nearest_table(H, new_h)).to(dev)
        xi = torch.from_numpy(nearest_table(W, new_w)).to(dev)
        return ops.nearest_u8(x, yi, xi)
    if new_w != W:                                   # horizontal pass first (Resample.c ImagingResampleInner)
        b, k = resample_tables(W, new_w)
        x = ops.resample_u8(x, new_w, 0, torch.from_numpy(b).to(dev), torch.from_numpy(k).to(dev))
    if new_h != H:
        b, k = resample_tables(H, new_h)
        x = ops.resample_u8(x, new_h, 1, torch.from_numpy(b).to(dev), torch.from_numpy(k).to(dev))
    return x


def patch_coords(h: int, w: int, patch_size: int) -> List[Tuple[int, int]]:
    """(y, x) of the 50 %-overlap grid of precompute_all_patches (data_loading.py:373-376)."""
    stride = patch_size // 2
    return [(y, x) for y in range(0, h - patch_size + 1, stride) for x in range(0, w - patch_size + 1, stride)]


class DevicePatchPipeline:
    """One image (+ mask) -> its training / validation patches, on the device.

        pipe = DevicePatchPipeline(patch_size=512, scale=1.0, split="train")
        out = pipe(img_u8_hwc, mask_u8_hw)          # uint8 device tensors (the decoded JPEG / TIFF)
        out["images"] [T,3,p,p] fp32 in [0,1], out["masks"] [T,1,p,p] {0,1}, out["coords"] [(y,x)], out["has_lesion"] [T] bool

    Mirrors IDRIDDataset.preprocess + the patch loop of precompute_all_patches: scale both with PIL's resampling
    (BICUBIC / NEAREST), /255 and (> 0), cut the 50 %-overlap grid, drop patches with too much black border
    (is_valid_patch: ratio of pixels whose channel mean < 0.1 above 0.1, or 0.5 for the test split), flag lesions
    (any mask pixel > 0.5). Only the validity / lesion flags are read back (T x 2 floats)."""

    def __init__(self, patch_size: int = 512, scale: float = 1.0, split: str = "train", skip_border_check: bool = False):
        self.patch_size, self.scale, self.split, self.skip_border_check = patch_size, scale, split, skip_border_check

    def preprocess(self, arr_u8: torch.Tensor, is_mask: bool) -> torch.Tensor:
        """-> the SCALED uint8 image (HWC) still in 8 bits (the /255 or > 0 conversion is fused into the tile gather)."""
        if arr_u8.dim() == 2:
            arr_u8 = arr_u8.unsqueeze(-1)
        H, W, _ = arr_u8.shape
        new_w, new_h = int(self.scale * W), int(self.scale * H)
        if new_w < 1 or new_h < 1:
            raise ValueError(f"image scaled too small => {new_w}x{new_h}")
        if (new_w, new_h) == (W, H):
            return arr_u8.contiguous()
        return resize_u8(arr_u8, new_w, new_h, is_mask)

    def __call__(self, img_u8_hwc: torch.Tensor, mask_u8_hw: Optional[torch.Tensor] = None) -> Dict[str, object]:
        img = self.preprocess(img_u8_hwc, False)
        if img.shape[2] == 1:
            img = img.expand(-1, -1, 3).contiguous()             # grey image -> 3 channels (data_loading.py:597-598)
        msk = self.preprocess(mask_u8_hw, True) if mask_u8_hw is not None else None
        if msk is not None and msk.shape[2] > 1:
            msk = msk[..., :1].contiguous()                       # :593-594
        h, w, _ = img.shape
        p = self.patch_size
        if h < p or w < p:
            return {"images": None, "masks": None, "coords": [], "has_lesion": None, "scaled_shape": (h, w)}
        coords = patch_coords(h, w, p)
        cd = torch.tensor(coords, dtype=torch.int32).to(img.device)
        tiles = ops.tile_gather(img, cd, p, p, 255.0)
        masks = ops.tile_gather(msk, cd, p, p, 1.0, binarise=True) if msk is not None else None
        stats = ops.tile_stats(tiles, masks.reshape(len(coords), p, p) if masks is not None else None, 0.1).cpu()
        thr = 0.5 if self.split == "test" else 0.1
        keep = [i for i in range(len(coords)) if self.skip_border_check or float(stats[i, 0]) <= thr]
        idx = torch.tensor(keep, dtype=torch.long, device=img.device)
        return {"images": tiles.index_select(0, idx), "masks": masks.index_select(0, idx) if masks is not None else None,
                "coords": [coords[i] for i in keep], "has_lesion": stats[keep, 1] > 0.5, "scaled_shape": (h, w),
                "black_ratio": stats[keep, 0]}
