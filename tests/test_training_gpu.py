"""GPU parity of the whole training step (reference train.py:385-411) against the CPU oracle
(oracle/vaeunet_oracle.py training_grads / clip_grad_norm_ / adamw_step, pinned to the unmodified reference's
autograd by tests/test_oracle_golden.py::test_training_step).

Tolerances. BASELINE's bar (1e-2 relative, 16-bit) is stated for logits; the loss is held to it. Gradients of this
seeded random-weight network are far more sensitive to storage rounding than its outputs: the CPU oracle itself,
evaluated with 16-bit rounding of activations / activation gradients / filter banks (oracle.storage_rounding) and
fp32 arithmetic, moves the full gradient by ~10 % (fp16) / ~30 % (bf16) relative L2 at this fixture size, and is
chaotic at batch 2 (BatchNorm over two constant latent maps). The test therefore measures that "format budget" in the
same run and requires the GPU gradients to stay within a small factor of it, tensor by tensor (a wiring mistake
shows up as an O(1) error on specific tensors), plus absolute floors on the cosine of the full gradient.
The bf16 margins carry the spread between equally valid evaluation orders: the same step with the BatchNorm statistics
taken in the conv epilogue or in a separate pass (identical to ~1e-8) lands at cosine 0.897 / 0.911 on this fixture
(budget 0.953), 0.925 / 0.918 at B=8 S=128 (budget 0.963) - scripts/train_report.py prints the table.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(lesion, dev, dtype, loss_scale, B=2, S=64, seed=21):
    from oracle import vaeunet_oracle as O
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.training import TrainStep
    from vaeunet_b200.unet import UNetResNet
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw)
    model.load_state_dict(sd)
    model = model.to(dev).train()
    ts = TrainStep(model, lesion=lesion, dtype=dtype, loss_scale=loss_scale)
    x, m = synthetic_fundus(B, S, S, seed=seed, lesion=lesion)
    eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(seed + 1))
    return O, sd, model, ts, x, m, eps


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def training_report(lesion, dev, dtype, loss_scale, B=8, S=64):
    """-> dict with loss / kl errors, per-tensor gradient errors, running-stat errors, post-step parameter errors.
    "budget" = the same errors for the CPU oracle evaluated with 16-bit storage rounding of activations, activation
    gradients and filter banks (oracle.storage_rounding): what the number format alone costs on this fixture."""
    from vaeunet_b200 import ops
    O, sd, model, ts, x, m, eps = _setup(lesion, dev, dtype, loss_scale, B, S)
    beta = 0.001
    loss_r, recon_r, kl_r, grads_r, upd_r = O.training_grads(sd, O.ResNetCfg(), x, m, eps, beta, 1e-3, lesion)
    with O.storage_rounding(torch.bfloat16 if dtype == ops.BF16 else torch.float16):
        loss_q, _, kl_q, grads_q, _ = O.training_grads(sd, O.ResNetCfg(), x, m, eps, beta, 1e-3, lesion)
    ts.zero_grad()
    out = ts.forward_backward(x.to(dev), m.to(dev), beta, eps.to(dev))
    torch.cuda.synchronize()
    rep = {"loss": (out["loss"].item(), loss_r.item()), "kl": (out["kl"].item(), kl_r.item()), "grads": {}}
    flat_g, flat_r = [], []
    for k, gr in grads_r.items():
        g = (ts.G[k] / ts.loss_scale).cpu()
        rep["grads"][k] = (_rel(g, gr), gr.norm().item())
        flat_g.append(g.flatten()); flat_r.append(gr.flatten())
    fg, fr = torch.cat(flat_g), torch.cat(flat_r)
    fq = torch.cat([grads_q[k].flatten() for k in grads_r])
    cos = lambda a, b: (a.double() @ b.double() / (a.double().norm() * b.double().norm())).item()
    rep["cosine"] = cos(fg, fr)
    rep["total_rel"] = _rel(fg, fr)
    rep["budget"] = {k: _rel(grads_q[k], grads_r[k]) for k in grads_r}
    rep["budget_cosine"], rep["budget_total_rel"] = cos(fq, fr), _rel(fq, fr)
    rep["vs_rounded_cosine"], rep["vs_rounded_total_rel"] = cos(fg, fq), _rel(fg, fq)
    rep["budget_loss"] = loss_q.item()
    rep["grad_norm"] = (fg.norm().item(), fr.norm().item())
    bufs = dict(model.named_buffers())
    rep["running"] = {k: _rel(bufs[k].cpu(), v) for k, v in upd_r.items()}
    # optimizer: clip to 1.0 + AdamW on the oracle gradients vs the device step on the device gradients
    params_r = {k: sd[k].clone() for k in grads_r}
    O.clip_grad_norm_(grads_r, 1.0)
    O.adamw_step(params_r, grads_r, {}, lr=1e-4, wd=1e-5)
    ts.optimizer_step()
    torch.cuda.synchronize()
    # the update of a parameter is lr * sign-like (|m/sqrt(v)| = 1 on the first step): compare the updates
    du, dr = [], []
    for k in grads_r:
        du.append((ts.P[k].cpu() - sd[k]).flatten()); dr.append((params_r[k] - sd[k]).flatten())
    du, dr = torch.cat(du), torch.cat(dr)
    rep["update_cosine"] = (du @ dr / (du.norm() * dr.norm())).item()
    rep["update_max"] = (du.abs().max().item(), dr.abs().max().item())
    return rep


@pytest.mark.parametrize("lesion,dtype_name,loss_scale", [("EX", "f16", 1024.0), ("MA", "f16", 1024.0), ("EX", "bf16", 1.0)])
def test_training_step_parity(cuda_dev, lesion, dtype_name, loss_scale):
    from vaeunet_b200 import ops
    dtype = {"bf16": ops.BF16, "f16": ops.F16}[dtype_name]
    rep = training_report(lesion, cuda_dev, dtype, loss_scale, B=8, S=64)
    f16 = dtype_name == "f16"
    got, ref = rep["loss"]
    assert abs(got - ref) <= (2e-3 if f16 else 1e-2) * abs(ref), rep["loss"]          # BASELINE: 1e-2 relative at 16 bit
    got, ref = rep["kl"]
    assert abs(got - ref) <= (2e-3 if f16 else 1e-2) * abs(ref) + 1e-5, rep["kl"]
    # full gradient: within the format budget measured by the rounded oracle (+ margin), and above an absolute floor
    assert rep["cosine"] >= rep["budget_cosine"] - (0.01 if f16 else 0.08), (rep["cosine"], rep["budget_cosine"])
    assert rep["cosine"] >= (0.985 if f16 else 0.87), rep["cosine"]
    assert rep["total_rel"] <= 1.5 * rep["budget_total_rel"] + 0.02, (rep["total_rel"], rep["budget_total_rel"])
    gmax = max(n for _, n in rep["grads"].values())
    bad = {k: (r, rep["budget"][k]) for k, (r, n) in rep["grads"].items()
           if n > 1e-3 * gmax and r > 3.0 * rep["budget"][k] + 0.05}
    assert not bad, f"tensors beyond 3x the format budget: {bad}"
    worst_run = max(rep["running"].values())
    assert worst_run <= (2e-2 if f16 else 8e-2), worst_run
    assert rep["update_cosine"] >= (0.9 if f16 else 0.7), rep["update_cosine"]        # step 1 of Adam is sign-like
    assert abs(rep["update_max"][0] - rep["update_max"][1]) <= 0.05 * rep["update_max"][1]


def test_training_loss_decreases(cuda_dev):
    """Ten steps on one fixed batch: the loss must go down and stay finite (the step is usable, not just parity-shaped)."""
    from vaeunet_b200 import ops
    O, sd, model, ts, x, m, eps = _setup("EX", cuda_dev, ops.BF16, 1.0, B=2, S=64)
    ts.lr = 1e-3
    xd, md = x.to(cuda_dev), m.to(cuda_dev)
    losses = []
    for it in range(10):
        out = ts.step(xd, md, 0.001, eps.to(cuda_dev))
        losses.append(out["loss"].item())
    assert all(l == l and abs(l) < 1e4 for l in losses), losses
    assert losses[-1] < losses[0], losses
    # after training the eval-mode module surface sees the updated weights (packed caches invalidated)
    model.eval()
    with torch.no_grad():
        lg, mu, lv = model(xd, eps=eps.to(cuda_dev))
    assert torch.isfinite(lg).all()


def test_gradient_accumulation(cuda_dev):
    """Two micro-batches with accum=2 == one batch of the concatenation for the loss terms that are means over the batch is
    NOT an identity under BatchNorm; what must hold exactly is linearity: running the same micro-batch twice with accum=2
    gives the gradient of running it once."""
    from vaeunet_b200 import ops
    O, sd, model, ts, x, m, eps = _setup("EX", cuda_dev, ops.F16, 1024.0)
    d = cuda_dev
    ts.zero_grad()
    ts.forward_backward(x.to(d), m.to(d), 0.001, eps.to(d))
    g1 = ts.flat_g.clone()
    ts.zero_grad()
    ts.forward_backward(x.to(d), m.to(d), 0.001, eps.to(d), accum=2)
    ts.forward_backward(x.to(d), m.to(d), 0.001, eps.to(d), accum=2)
    g2 = ts.flat_g.clone()
    assert ((g1 - g2).norm() / g1.norm()).item() < 2e-2      # split-K atomics + 16-bit rounding of halved gradients
