"""Golden vectors for the SURVEY 8(f) "next" rows, from the UNMODIFIED reference (/root/reference):
  evaluate.evaluate (evaluate.py:19-101)                          -> validation metrics of a small synthetic loader
  visualize_vae.generate_ensemble_prediction (visualize_vae.py:1083-1163) -> temperature-ensemble probability map

Run in the build container only:   python tests/golden/make_golden_next.py      (writes golden_next.pt, ~100 KB)
Uses the committed fixture state_dict (tests/golden/fixture_resnet.pt over the seeded init); does not touch the other
golden files.
"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from unet.unet_resnet import UNetResNet as RefUNetResNet  # noqa: E402  (reference)
import evaluate as ref_eval  # noqa: E402
import visualize_vae as ref_vis  # noqa: E402

from vaeunet_b200 import testing  # noqa: E402
from vaeunet_b200.synthetic import synthetic_fundus  # noqa: E402

OUT = Path(__file__).resolve().parent
torch.set_num_threads(8)
L = 32


def main():
    sd, kw = testing.resnet_fixture()
    model = RefUNetResNet(3, 1, pretrained=False)
    model.load_state_dict(sd)
    model.eval()
    G = {}

    # --- evaluate(): 3 batches of (images, masks) tuples, then max_samples = 2; the model's forward draws eps per batch
    seeds = [900, 901, 902]
    loader = [synthetic_fundus(2, 64, 80, seed=s) for s in seeds]
    g = torch.Generator().manual_seed(77)
    eps = [torch.randn(2, L, generator=g) for _ in seeds]
    with ref_shim.inject_eps(list(eps)):
        avg, samples = ref_eval.evaluate(model, loader, torch.device("cpu"), amp=False)
    with ref_shim.inject_eps(list(eps)):
        avg2, _ = ref_eval.evaluate(model, loader, torch.device("cpu"), amp=False, max_samples=2)
    G["evaluate"] = dict(seeds=seeds, shape=(2, 64, 80), eps=eps, avg=avg, avg_max2=avg2, n_samples=len(samples),
                         sample0_pred=torch.as_tensor(samples[0][1]), sample0_metrics=samples[0][3])

    # --- generate_ensemble_prediction: odd-sized image, 3 temperatures x 2 samples, patches of 64 with overlap 16
    x, _ = synthetic_fundus(1, 100, 120, seed=910)
    temps = [0.5, 1.0, 2.5]
    eps_e = [torch.randn(1, L, generator=g) for _ in range(len(temps) * 2)]
    with ref_shim.inject_eps(list(eps_e)):
        ens = ref_vis.generate_ensemble_prediction(model, x, temps=temps, samples_per_temp=2, weighted=True, patch_size=64,
                                                   overlap=16, batch_size=4)
    with ref_shim.inject_eps(list(eps_e)):
        ens_u = ref_vis.generate_ensemble_prediction(model, x, temps=temps, samples_per_temp=2, weighted=False,
                                                     patch_size=64, overlap=16, batch_size=4)
    G["ensemble"] = dict(seed=910, shape=(100, 120), temps=temps, samples_per_temp=2, patch=64, overlap=16,
                         eps=torch.stack(eps_e), out=ens, out_unweighted=ens_u)
    torch.save(G, OUT / "golden_next.pt")
    print("evaluate:", avg, "| max_samples=2:", avg2)
    print("ensemble:", tuple(ens.shape), float(ens.mean()), float(ens_u.mean()))


if __name__ == "__main__":
    main()
