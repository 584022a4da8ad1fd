"""Golden vectors for SURVEY 8(f) row 4 (patch pipeline) from the UNMODIFIED reference's IDRIDDataset.preprocess
(utils/data_loading.py:580-601, i.e. PIL.Image.resize BICUBIC / NEAREST) and its patch loop (:368-398).

Run in the build container only:   python tests/golden/make_golden_f4.py      (writes golden_f4.pt, a few KB)
`albumentations` (augmentation library, not on the hot path) is absent here: an empty module stands in for the import.
Inputs are regenerated from seeds; stored: SHA-256 of the resized arrays, patch coordinates, lesion flags, patch hashes."""
import hashlib
import sys
import types
from pathlib import Path

import numpy as np
import torch
from PIL import Image

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")
sys.modules.setdefault("albumentations", types.ModuleType("albumentations"))
from utils.data_loading import IDRIDDataset  # noqa: E402  (reference)

from vaeunet_b200.testing import f4_inputs  # noqa: E402

OUT = Path(__file__).resolve().parent
sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


class _Stub:
    """Just enough of IDRIDDataset for the two methods called here."""
    skip_border_check = False
    is_full_image = False

    def __init__(self, split):
        self.split = split


def main():
    G = {}
    for tag, (seed, H, W, scale, patch, split) in {"half": (41, 320, 448, 0.5, 64, "train"),
                                                    "third_test": (42, 300, 420, 1.0 / 3.0, 32, "test"),
                                                    "unit": (43, 96, 160, 1.0, 48, "train")}.items():
        img, mask = f4_inputs(seed, H, W)
        stub = _Stub(split)
        ia = IDRIDDataset.preprocess(stub, Image.fromarray(img), scale, is_mask=False)      # [3,h,w] fp32 in [0,1]
        ma = IDRIDDataset.preprocess(stub, Image.fromarray(mask), scale, is_mask=True)      # [h,w] fp32 {0,1}
        it, mt = torch.as_tensor(ia), torch.as_tensor(ma).unsqueeze(0)
        _, h, w = it.shape
        coords, lesion, hashes = [], [], []
        stride = patch // 2
        for y in range(0, h - patch + 1, stride):                                          # data_loading.py:373-398
            for x in range(0, w - patch + 1, stride):
                ip, mp = it[:, y:y + patch, x:x + patch], mt[:, y:y + patch, x:x + patch]
                if not IDRIDDataset.is_valid_patch(stub, ip):
                    continue
                coords.append((y, x)); lesion.append(bool(torch.any(mp > 0.5)))
                hashes.append(sha(ip.contiguous().numpy())[:16] + sha(mp.contiguous().numpy())[:16])
        G[tag] = dict(seed=seed, shape=(H, W), scale=scale, patch=patch, split=split, scaled=(h, w),
                      img_sha=sha(ia), mask_sha=sha(ma), coords=coords, lesion=lesion, hashes=hashes)
        print(tag, (h, w), len(coords), sum(lesion), sha(ia)[:12])
    torch.save(G, OUT / "golden_f4.pt")


if __name__ == "__main__":
    main()
