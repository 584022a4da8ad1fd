"""GPU parity of the patch-pipeline kernels and the batched tile blending (csrc/vu_tiles.cu) through the C ABI:
bit-exact against the oracle's restatement of PIL.Image.resize + the reference's patch loop, and against the golden hashes
of the unmodified reference."""
import hashlib
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_device_patch_pipeline_bit_exact(cuda_dev):
    from oracle import vaeunet_oracle as O
    from vaeunet_b200.testing import f4_inputs
    from vaeunet_b200.utils.data_loading import DevicePatchPipeline, resize_u8
    G = torch.load(GOLD / "golden_f4.pt", weights_only=False)
    for tag, g in G.items():
        img, mask = f4_inputs(g["seed"], *g["shape"])
        pipe = DevicePatchPipeline(g["patch"], g["scale"], g["split"])
        out = pipe(torch.from_numpy(img).to(cuda_dev), torch.from_numpy(mask).to(cuda_dev))
        torch.cuda.synchronize()
        assert out["coords"] == g["coords"], tag
        assert [bool(v) for v in out["has_lesion"]] == g["lesion"], tag
        hs = [sha(out["images"][i].cpu().numpy())[:16] + sha(out["masks"][i].cpu().numpy())[:16] for i in range(len(g["coords"]))]
        assert hs == g["hashes"], tag                       # the reference's patches, bit for bit
    # resize alone, up- and down-scaling, against the oracle (== PIL)
    rng = np.random.default_rng(7)
    for (H, W, nh, nw) in [(97, 131, 24, 32), (50, 70, 75, 105), (233, 177, 58, 44), (64, 80, 64, 40)]:
        a = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        got = resize_u8(torch.from_numpy(a).to(cuda_dev), nw, nh, False).cpu().numpy()
        assert np.array_equal(got, O.pil_resize_u8(a, nw, nh, False))
        m = (rng.integers(0, 2, (H, W, 1), dtype=np.uint8) * 255)
        gm = resize_u8(torch.from_numpy(m).to(cuda_dev), nw, nh, True).cpu().numpy()
        assert np.array_equal(gm[..., 0], O.pil_resize_u8(m[..., 0], nw, nh, True))


def test_full_scale_image_tiling(cuda_dev):
    """configs[4] geometry: a 2848 x 4288 8-bit image -> the 50 %-overlap 512 grid (10 x 15 = 150 patches before the border
    check) entirely on the device; spot-checked against slices of the host array, and the prediction grid (overlap 100:
    7 x 11 = 77 tiles) through tile_gather against torch slicing."""
    from vaeunet_b200 import ops, predict
    from vaeunet_b200.testing import f4_inputs
    from vaeunet_b200.utils.data_loading import DevicePatchPipeline, patch_coords
    img, mask = f4_inputs(44, 2848, 4288)
    d = cuda_dev
    out = DevicePatchPipeline(512, 1.0, "train", skip_border_check=True)(torch.from_numpy(img).to(d), torch.from_numpy(mask).to(d))
    coords = patch_coords(2848, 4288, 512)
    assert len(coords) == 10 * 15 and out["coords"] == coords and out["images"].shape == (150, 3, 512, 512)
    for i in (0, 17, 77, 149):
        y, x = coords[i]
        ref = torch.from_numpy(img[y:y + 512, x:x + 512].astype(np.float32) / 255.0).permute(2, 0, 1)
        assert torch.equal(out["images"][i].cpu(), ref)
        assert torch.equal(out["masks"][i, 0].cpu(), torch.from_numpy((mask[y:y + 512, x:x + 512] > 0).astype(np.float32)))
    nh, nw, info = predict.patch_grid(2848, 4288, 512, 100)
    assert (nh, nw) == (7, 11)
    f = torch.from_numpy(img.astype(np.float32) / 255.0).permute(2, 0, 1).unsqueeze(0).to(d)
    idx = [0, 10, 38, 76]
    cd = torch.tensor([[info[k][0], info[k][2]] for k in idx], dtype=torch.int32, device=d)
    tiles = ops.tile_gather(f, cd, 512, 512)
    for t, k in enumerate(idx):
        assert torch.equal(tiles[t], f[0, :, info[k][0]:info[k][1], info[k][2]:info[k][3]])


def test_blend_accumulate_batch_equals_per_tile(cuda_dev):
    """One launch per tile batch == one launch per (tile, sample): same ramp weights, same fp32 summation order (bit-exact),
    with ONE weight map; finalize divides every sample by it."""
    from vaeunet_b200 import ops, predict
    d = cuda_dev
    H, W, P, ov, Ns = 150, 170, 64, 16, 3
    nh, nw, info = predict.patch_grid(H, W, P, ov)
    g = torch.Generator().manual_seed(8)
    probs = torch.rand(len(info), Ns, P, P, generator=g).to(d)
    acc_a = torch.zeros(Ns, H, W, device=d); ws_a = torch.zeros(Ns, H, W, device=d)
    for k, (sh, eh, sw, ew) in enumerate(info):
        for n in range(Ns):
            ops.blend_accumulate(probs[k, n].contiguous(), sh, sw, k // nw, k % nw, nh, nw, ov, acc_a[n], ws_a[n])
    acc_b = torch.zeros(Ns, H, W, device=d); ws_b = torch.zeros(H, W, device=d)
    for lo in range(0, len(info), 4):
        grp = list(range(lo, min(lo + 4, len(info))))
        tinfo = torch.tensor([[info[k][0], info[k][2], k // nw, k % nw] for k in grp], dtype=torch.int32, device=d)
        bbox = (min(info[k][0] for k in grp), max(info[k][1] for k in grp), min(info[k][2] for k in grp), max(info[k][3] for k in grp))
        ops.blend_accumulate_batch(probs[grp].contiguous(), tinfo, bbox, nh, nw, ov, acc_b, ws_b)
    torch.cuda.synchronize()
    assert torch.equal(acc_a, acc_b) and torch.equal(ws_a[0], ws_b)
    assert torch.equal(ops.blend_finalize(acc_a, ws_a), ops.blend_finalize_shared(acc_b, ws_b))
