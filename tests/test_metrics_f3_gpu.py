"""GPU parity of the sort-based uncertainty metrics (csrc/vu_sort_metrics.cu) through the C ABI:
the device radix sort against torch.sort, the sparsification curve and the uncertainty-vs-error AUROC / AUPRC against
the golden vectors of the unmodified reference (tests/golden/golden_f3.pt) and the numpy oracle."""
import math
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


@pytest.mark.parametrize("n,quant", [(1, 0), (37, 0), (2048, 0), (2049, 16), (100003, 0), (262144, 200), (1 << 20, 7)])
def test_radix_sort_pairs_desc_is_torch_stable_sort(cuda_dev, n, quant):
    """bit-exact: keys descending, payload order == a stable descending sort (ties keep input order); negatives, zeros and
    heavy ties included; sizes that are not multiples of the 2048-element block tile."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(n)
    s = torch.randn(n, generator=g)
    if quant:
        s = torch.round(s * quant) / quant
    if n > 4:
        s[1] = 0.0; s[2] = -0.0; s[3] = s[0]
    payload = torch.arange(n, dtype=torch.int32)
    keys, vals, _ = ops.sort_pairs_desc(s.to(cuda_dev), payload.to(cuda_dev))
    torch.cuda.synchronize()
    # reference: stable sort of the ORDER-PRESERVING keys (so that -0.0 < +0.0 like the device's total order)
    u = s.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    asc = torch.where(u >> 31 == 1, u ^ 0xFFFFFFFF, u ^ 0x80000000)
    order = torch.sort(-asc, stable=True).indices
    assert torch.equal(vals.cpu().to(torch.int64), order)
    got = s[vals.cpu().long()]
    assert bool((got[:-1] >= got[1:]).all())


def _run_sparsification(p, unc, gt, num_points, dev):
    from vaeunet_b200.utils.uncertainty_metrics import calculate_sparsification_metrics
    return calculate_sparsification_metrics(p.to(dev), unc.to(dev), gt.to(dev), num_points)


def test_sparsification_curve_matches_reference_golden(cuda_dev):
    from oracle import vaeunet_oracle as O
    from vaeunet_b200.testing import f3_inputs
    G = torch.load(GOLD / "golden_f3.pt", weights_only=False)
    for tag in ("sp_small", "sp_one"):
        g = G[tag]
        B, H, W = g["shape"]
        p, gt, unc = f3_inputs(g["seed"], B, H, W, False)
        fr, er, eu = _run_sparsification(p, unc, gt, 20, cuda_dev)
        assert abs(fr - g["frac"].numpy()).max() == 0
        # fp32 pixel errors (device logf vs torch CPU log) summed in fp64: 1e-5 on the normalised curve
        assert abs(er - g["rand"].numpy()).max() < 1e-5, tag
        assert abs(eu - g["unc"].numpy()).max() < 1e-5, tag
    g = G["sp_degenerate"]
    p, gt, unc = f3_inputs(13, 2, 20, 24, False)
    p[1] = 0.5; gt[1] = 1.0                          # all errors equal: the image is skipped but still divides
    _, er, eu = _run_sparsification(p, unc, gt, 10, cuda_dev)
    assert abs(er - g["rand"].numpy()).max() < 1e-5 and abs(eu - g["unc"].numpy()).max() < 1e-5
    # a full 512 x 512 patch against the oracle (BASELINE configs[1] map size)
    p, gt, unc = f3_inputs(99, 1, 512, 512, False)
    _, er_o, eu_o = O.sparsification_metrics(p, unc, gt, 20)
    _, er, eu = _run_sparsification(p, unc, gt, 20, cuda_dev)
    assert abs(er - er_o).max() < 1e-5 and abs(eu - eu_o).max() < 1e-5
    assert eu[-1] < eu[0] and all(eu[i + 1] <= eu[i] + 1e-12 for i in range(len(eu) - 1))     # monotone envelope


def test_uncertainty_error_auc_matches_reference_golden(cuda_dev):
    from oracle import vaeunet_oracle as O
    from vaeunet_b200.testing import f3_inputs
    from vaeunet_b200.utils.uncertainty_metrics import calculate_uncertainty_error_auc
    G = torch.load(GOLD / "golden_f3.pt", weights_only=False)
    d = cuda_dev
    for tag in ("auc_cont", "auc_ties"):
        g = G[tag]
        H, W = g["shape"]
        p, gt, unc = f3_inputs(g["seed"], 1, H, W, g["ties"])
        a, b = calculate_uncertainty_error_auc(p[0].to(d), gt[0].to(d), unc[0].to(d))
        assert abs(a - g["auroc"]) < 1e-10 and abs(b - g["auprc"]) < 1e-10, (tag, a, b)       # integer counts, fp64 sums
    p, gt, unc = f3_inputs(23, 1, 16, 16, False)
    a, b = calculate_uncertainty_error_auc(p[0].to(d), (p[0] > 0.5).float().to(d), unc[0].to(d))
    assert math.isnan(a) and math.isnan(b)           # one class only: the reference's except branch
    # full patch, heavy ties, against the oracle's restatement of sklearn
    for ties in (False, True):
        p, gt, unc = f3_inputs(77, 1, 512, 512, ties)
        a_o, b_o = O.uncertainty_error_auc(p[0], gt[0], unc[0])
        a, b = calculate_uncertainty_error_auc(p[0].to(d), gt[0].to(d), unc[0].to(d))
        assert abs(a - a_o) < 1e-10 and abs(b - b_o) < 1e-10, (ties, a, a_o, b, b_o)
