"""Golden vectors for SURVEY 8(f) row 3 from the UNMODIFIED reference (/root/reference/utils/uncertainty_metrics.py):
  calculate_sparsification_metrics (:175-240)   - numpy argsort + np.random baselines
  calculate_uncertainty_error_auc (:270-283)    - sklearn roc_auc_score / average_precision_score

Run in the build container only:   python tests/golden/make_golden_f3.py      (writes golden_f3.pt, < 200 KB)
Inputs are regenerated from seeds by the tests; only seeds + the reference's outputs are stored."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from utils import uncertainty_metrics as ref_um  # noqa: E402  (reference)

from vaeunet_b200.testing import f3_inputs  # noqa: E402

OUT = Path(__file__).resolve().parent


def main():
    G = {}
    for tag, (seed, B, H, W) in {"sp_small": (11, 2, 37, 53), "sp_one": (12, 1, 64, 64)}.items():
        p, gt, unc = f3_inputs(seed, B, H, W, ties=False)
        fr, er, eu = ref_um.calculate_sparsification_metrics(p, unc, gt, num_points=20)
        G[tag] = dict(seed=seed, shape=(B, H, W), frac=torch.from_numpy(fr), rand=torch.from_numpy(er), unc=torch.from_numpy(eu))
    # degenerate image in the batch (all errors equal) is skipped but still counted in the divisor
    p, gt, unc = f3_inputs(13, 2, 20, 24, ties=False)
    p[1] = 0.5; gt[1] = 1.0
    fr, er, eu = ref_um.calculate_sparsification_metrics(p, unc, gt, num_points=10)
    G["sp_degenerate"] = dict(seed=13, shape=(2, 20, 24), rand=torch.from_numpy(er), unc=torch.from_numpy(eu))
    for tag, (seed, H, W, ties) in {"auc_cont": (21, 61, 47, False), "auc_ties": (22, 80, 96, True)}.items():
        p, gt, unc = f3_inputs(seed, 1, H, W, ties)
        auroc, auprc = ref_um.calculate_uncertainty_error_auc(p[0], gt[0], unc[0])
        G[tag] = dict(seed=seed, shape=(H, W), ties=ties, auroc=float(auroc), auprc=float(auprc))
    # one class only -> (nan, nan)
    p, gt, unc = f3_inputs(23, 1, 16, 16, False)
    gt0 = (p[0] > 0.5).float()
    a, b = ref_um.calculate_uncertainty_error_auc(p[0], gt0, unc[0])
    # scikit-learn ~= 1.6.1 (the reference's pin, requirements.txt:10) raises ValueError for a single class and the reference's
    # except branch returns (nan, nan); newer sklearn (1.9 in this container) returns nan + a defined AP with a warning.
    # The golden stores the PINNED behaviour.
    assert a != a
    G["auc_one_class"] = dict(seed=23, shape=(16, 16), auroc=float("nan"), auprc=float("nan"))
    torch.save(G, OUT / "golden_f3.pt")
    for k, v in G.items():
        print(k, {kk: (vv if not torch.is_tensor(vv) else vv[:4].tolist()) for kk, vv in v.items()})


if __name__ == "__main__":
    import logging
    logging.disable(logging.CRITICAL)
    main()
