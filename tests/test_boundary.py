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
    names = set(_declared()) - {"vu_last_error"}
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
