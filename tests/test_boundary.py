"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/vaeunet_b200.h declares, the ctypes table covers the header, and the product path fails
loudly (no CPU fallback) when no B200 is visible. No compute call is made here."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

from vaeunet_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "vaeunet_b200.h"


def _declared():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(vu_[a-zA-Z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 25
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_ctypes_table_matches_header():
    names = set(_declared()) - {"vu_last_error", "vu_sort_workspace_bytes"}   # the two non-int-returning entry points
    assert names == set(_lib.SIGNATURES), (names ^ set(_lib.SIGNATURES))
    _lib.load()  # resolves restype/argtypes for each of them


def test_version_and_error_text():
    lib = _lib.load()
    assert lib.vu_version() >= 1
    assert isinstance(lib.vu_last_error(), bytes)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback_without_device():
    from vaeunet_b200 import ops
    with pytest.raises(_lib.VaeUnetError):
        ops.reparameterize(torch.zeros(4), torch.zeros(4), torch.zeros(2, 4))
    from vaeunet_b200.unet import UNetResNet
    m = UNetResNet(3, 1, pretrained=False).eval()
    with pytest.raises(_lib.VaeUnetError):
        m(torch.zeros(1, 3, 64, 64))


def test_module_surface_names():
    """importable names and constructor signatures of the reference."""
    import inspect
    from vaeunet_b200.unet import UNet, UNetResNet
    from vaeunet_b200.unet.unet_parts import AttentionGate, DoubleConv, Down, OutConv, Up  # noqa: F401
    from vaeunet_b200.unet.unet_resnet import DecoderBlock
    from vaeunet_b200.utils import loss, uncertainty_metrics, vae_utils
    assert list(inspect.signature(UNet.__init__).parameters)[1:] == ["n_channels", "n_classes", "bilinear"]
    assert list(inspect.signature(UNetResNet.__init__).parameters)[1:] == [
        "n_channels", "n_classes", "backbone", "pretrained", "latent_dim", "use_attention", "use_skip",
        "latent_injection"]
    assert list(inspect.signature(DecoderBlock.__init__).parameters)[1:8] == [
        "in_channels", "skip_channels", "out_channels", "latent_dim", "use_attention", "use_skip", "use_latent"]
    for n in ("sample_from_latent", "encode_images", "generate_predictions", "calculate_latent_stats"):
        assert hasattr(vae_utils, n)
    for n in ("dice_loss", "DiceLoss", "CombinedLoss", "MAFocalLoss", "MASegmentationLoss", "KLAnnealer",
              "kl_with_free_bits"):
        assert hasattr(loss, n)
    assert hasattr(uncertainty_metrics, "calculate_expected_calibration_error")


def test_torch_custom_ops_registered_and_fake():
    """the ops are torch.library custom ops (torch.ops.vaeunet.*) with fake (meta) implementations - shape
    propagation works without a device - and autograd registrations on the differentiable ones."""
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    from vaeunet_b200 import torch_ops
    for n in torch_ops.OP_NAMES:
        assert hasattr(torch.ops.vaeunet, n), n
    with FakeTensorMode():
        x = torch.empty(2, 32, 32, 64, dtype=torch.float16)
        assert torch.ops.vaeunet.conv2d(x, torch.empty(128, 64, 3, 3), None, 2, 1).shape == (2, 16, 16, 128)
        assert torch.ops.vaeunet.conv_transpose2x2(x, torch.empty(64, 32, 2, 2), torch.empty(32)).shape == (2, 64, 64, 32)
        y, mean, invstd = torch.ops.vaeunet.batch_norm_train(x, torch.empty(64), torch.empty(64), None, True)
        assert y.shape == x.shape and mean.shape == (64,)
        p, code = torch.ops.vaeunet.maxpool(x, 2, 2, 0)
        assert p.shape == (2, 16, 16, 64) and code.dtype == torch.uint8
        assert torch.ops.vaeunet.bilinear_ac(x, 64, 64).shape == (2, 64, 64, 64)
        img = torch.empty(2, 3, 64, 64)
        assert torch.ops.vaeunet.conv_small_cin(img, torch.empty(64, 3, 7, 7), 2, 3, 1).shape == (2, 32, 32, 64)
        m5 = torch.ops.vaeunet.uncertainty_reduce(torch.empty(2, 10, 32, 32), 64, 64)
        assert len(m5) == 5 and m5[0].shape == (2, 64, 64)


def test_module_train_mode_needs_a_device():
    """model.train(); model(x) goes to the autograd path (torch.ops.vaeunet.*), which - like everything else - has no CPU
    fallback: it raises VaeUnetError on a CPU tensor instead of silently running torch ops."""
    import pytest
    import torch
    from vaeunet_b200._lib import VaeUnetError
    from vaeunet_b200.unet import UNet, UNetResNet
    m = UNetResNet(3, 1, pretrained=False).train()
    with pytest.raises(VaeUnetError):
        m(torch.zeros(2, 3, 64, 64))
    u = UNet(3, 1).train()
    with pytest.raises(VaeUnetError):
        u(torch.zeros(2, 3, 64, 64))
