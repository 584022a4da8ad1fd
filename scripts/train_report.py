"""Training-step parity report (per-tensor gradient errors vs the CPU oracle) + a first timing of the step.
    python scripts/train_report.py [--time B S]"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
from test_training_gpu import training_report
from vaeunet_b200 import ops

dev = torch.device("cuda:0")
if "--time" not in sys.argv:
    Bs = [int(a) for a in sys.argv[1:] if a.isdigit()]
    B, S = (Bs + [8, 64])[:2] if len(Bs) >= 2 else (8, 64)
    for lesion, dtn, ls in [("EX", "f16", 1024.0), ("EX", "bf16", 1.0), ("MA", "f16", 1024.0)]:
        rep = training_report(lesion, dev, {"bf16": ops.BF16, "f16": ops.F16}[dtn], ls, B, S)
        print(f"== B={B} S={S}; budget: loss {rep['budget_loss']} cosine {rep['budget_cosine']:.5f} total_rel {rep['budget_total_rel']:.4f}; "
              f"GPU vs rounded oracle: cosine {rep['vs_rounded_cosine']:.5f} total_rel {rep['vs_rounded_total_rel']:.4f}")
        print(f"== {lesion} {dtn}: loss {rep['loss']} kl {rep['kl']} cosine {rep['cosine']:.5f} total_rel {rep['total_rel']:.4f} "
              f"grad_norm {rep['grad_norm']} update_cos {rep['update_cosine']:.4f} update_max {rep['update_max']}")
        worst = sorted(rep["grads"].items(), key=lambda kv: -kv[1][0])
        for k, (r, n) in worst[:25]:
            print(f"   {r:9.4f}  |g|={n:.3e}  {k}")
        big = sorted(rep["grads"].items(), key=lambda kv: -kv[1][1])[:12]
        print("   largest-norm tensors:")
        for k, (r, n) in big:
            print(f"   {r:9.4f}  |g|={n:.3e}  {k}")
        if "--all" in sys.argv:
            for k, (r, n) in rep["grads"].items():
                print(f"   ALL {r:9.4f}  budget {rep['budget'][k]:9.4f}  |g|={n:.3e}  {k}")
        wr = sorted(rep["running"].items(), key=lambda kv: -kv[1])[:5]
        print("   running stats worst:", [(k, round(v, 5)) for k, v in wr])
else:
    i = sys.argv.index("--time")
    B, S = int(sys.argv[i + 1]), int(sys.argv[i + 2])
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.training import TrainStep
    from vaeunet_b200.unet import UNetResNet
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
    ts = TrainStep(model)
    x, m = synthetic_fundus(2, S, S, seed=5)
    x = x.repeat(B // 2, 1, 1, 1).to(dev); m = m.repeat(B // 2, 1, 1, 1).to(dev)
    eps = torch.randn(B, 32, device=dev)
    for it in range(3):
        out = ts.step(x, m, 0.001, eps)
    torch.cuda.synchronize()
    print("loss", out["loss"].item(), "mem GB", torch.cuda.max_memory_allocated() / 2**30)
    # phase timing
    def timed(fn, n=3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(n): r = fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, r
    t_step, _ = timed(lambda: ts.step(x, m, 0.001, eps))
    t_pack, _ = timed(ts.repack)
    t_fwd, (_, _, _, Sv) = timed(lambda: ts.forward(x, eps))
    out, dl, dmu, dlv = ts.loss_and_grads(Sv, m, 0.001)
    ts.zero_grad()
    t_bwd, _ = timed(lambda: ts.backward(Sv, dl, dmu.clone(), dlv.clone()))
    t_opt, _ = timed(ts.optimizer_step)
    t0 = time.time(); ts.step(x, m, 0.001, eps); host = time.time() - t0
    torch.cuda.synchronize()
    print(f"B={B} S={S}: step {t_step:.2f} ms = {B / t_step * 1e3:.1f} img/s | repack {t_pack:.2f} fwd {t_fwd:.2f} bwd {t_bwd:.2f} opt {t_opt:.2f} | host enqueue {host*1e3:.1f} ms")
    # per conv-launch timing inside one step
    ops.CONV_TIMING = []
    ts.step(x, m, 0.001, eps)
    torch.cuda.synchronize()
    rows = [(e0.elapsed_time(e1), fl, d) for (e0, e1, fl, d) in ops.CONV_TIMING]
    ops.CONV_TIMING = None
    tot = sum(r[0] for r in rows); fl = sum(r[1] for r in rows)
    print(f"conv launches {len(rows)}: {tot:.2f} ms, {fl/1e9:.0f} GF, {fl/tot/1e9:.0f} TFLOP/s")
    agg = {}
    for t, f, d in rows:
        a = agg.setdefault(d, [0.0, 0.0, 0]); a[0] += t; a[1] += f; a[2] += 1
    for d, (t, f, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
        print(f"  {t:8.3f} ms x{n:2d} {f/t/1e9:7.0f} TF/s  {d}")
