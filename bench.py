"""Benchmarks of the B200 hot path (BASELINE.json). The default run prints ONE JSON line whose metric is MC-uncertainty
patches/s (configs[1]: 512x512 patches, T = 1.0, N = 10 latent samples, mean / std / entropy maps) and whose "train" object is
training images/s (configs[2]: VAE-UNet training step, batch 16 of 512x512 per GPU).

  python bench.py --gpus N --steps K --warmup W            # this repo's B200 path (torchrun for N > 1)
  python bench.py --impl reference ...                     # the reference algorithm on the host cores (oracle port;
                                                           #   /root/reference is not on the box)
  python bench.py --workload mc|train|unet|cfg1|cfg4|cfg5  # one workload as the line's metric (evidence lines for the other
                                                           #   BASELINE configs; see profiles/)

A step of the MC workload = `--batch` patches (default 384) through UNetResNet.mc_predict in micro-batches of 16 (encoder once
per patch, N batched decodes, fused upsample + sigmoid + uncertainty reduction), so that K = 10 steps are a ~3 s timed window
(long enough for the clocks to settle under the power cap; the roofline is printed against BOTH the burst and the sustained
cuBLAS peak of MEASURED_PEAKS.json with the SM clock median seen). Patches shard across ranks with no data-path collective
(weak scaling); time = max over ranks of CUDA-event time. `e2e` is the same loop fed from pinned host memory with the maps
copied back, all inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

PATCH, NS, TEMP = 512, 10, 1.0
MICRO = 16                   # patches per mc_predict call
F_ALG_PATCH_GF = 761.40      # F_enc + F_head + 10 * F_dec at 512^2 (dense, "reference-equivalent")
TRAIN_B = 16                 # configs[2]: batch 16 of 512x512 per GPU (weak scaling)
F_ALG_TRAIN_STEP_GF = 5289.0 # fwd + dgrad + wgrad of every conv at B=16 (no dgrad for the stem)
F_UNET_512_GF = 393.960      # basic UNet forward at 512^2
F_CFG4_PATCH_GF = 2109.6     # 72.186 + 0.032 + 15 * 135.824 at 700^2
F_CFG5_IMAGE_GF = 60412.0    # 77 * 761.40 + 1 784 (full-image encoder pass)
F_CFG1_GF = 27.651 + 9.569 + 0.0042 + 3 * 18.078   # forward + generate_predictions(N=3) at 256^2


def _peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        p = json.loads(f.read_text())
        return p.get("bf16_tflops_sustained", 1400.0), p.get("bf16_tflops", 1590.0), p.get("hbm_gbs", 6650.0), "measured"
    return 1400.0, 1590.0, 6650.0, "fallback"


def _traffic(key: str):
    """DRAM bytes per launch of the dominant kernel, from the committed `ncu --set full` summary (profiles/roofline_traffic.json)."""
    f = ROOT / "profiles" / "roofline_traffic.json"
    if f.exists():
        try:
            return json.loads(f.read_text()).get(key)
        except Exception:
            return None
    return None


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


class Ctx:
    """rank / world / device / dist of this process + the timing helpers every workload shares."""

    def __init__(self, args):
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist_
            self.dist = dist_
            dist_.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed(self, step, steps, warmup, pre_end=None):
        """W warm-up steps, then K steps between CUDA events on the current stream, barrier + synchronize on both sides,
        max over ranks. -> (ms, launches of this repo's kernels inside the timed region)."""
        from vaeunet_b200 import ops
        for i in range(warmup):
            step(i)
        self.barrier()
        l0 = ops.LAUNCHES
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            step(warmup + i)
        if pre_end is not None:
            pre_end()
        e1.record()
        self.barrier()
        ms = e0.elapsed_time(e1)
        if self.dist is not None:
            t = torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = t.item()
        return ms, ops.LAUNCHES - l0

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def _conv_roofline(step_once, n_steps, sustained, burst, src, kernel_desc, clocks, traffic_key):
    """Live per-launch timing of the tcgen05 conv kernels (CUDA events on the launch stream) over `n_steps` calls."""
    from vaeunet_b200 import ops
    ops.CONV_TIMING = []
    torch.cuda.synchronize()
    for i in range(n_steps):
        step_once(i)
    torch.cuda.synchronize()
    ev = ops.CONV_TIMING
    ops.CONV_TIMING = None
    conv_ms = sum(a.elapsed_time(b) for a, b, *_ in ev) / n_steps
    conv_flop = sum(f for _, _, f, _ in ev) / n_steps
    achieved = conv_flop / (conv_ms / 1e3) / 1e12 if conv_ms > 0 else 0.0
    tr = _traffic(traffic_key)
    return {"bound": "tensor", "kernel": kernel_desc, "achieved": achieved, "peak": sustained, "unit": "TFLOP/s",
            "frac": achieved / sustained, "frac_of_burst": achieved / burst, "peak_burst": burst,
            "peak_source": f"{src} MEASURED_PEAKS.json: bf16_tflops_sustained (frac) and bf16_tflops burst (frac_of_burst); "
                           f"SM clock median during the timed window {clocks.get('sm_mhz') if clocks else None} MHz",
            "launches_per_step": len(ev) // n_steps, "conv_ms_per_step": conv_ms, "alg_gflop_per_step_convs": conv_flop / 1e9,
            "traffic": (tr or {}).get("dram_bytes_per_launch"), "traffic_detail": tr}, ev


# ====================================================================================================== CPU baselines
def cpu_baseline(threads: int, reps: int = 3):
    """The reference algorithm (oracle port, fp32 PyTorch CPU) on ONE 512^2 patch, N=10."""
    from oracle import vaeunet_oracle as O
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    torch.set_num_threads(threads)
    sd, _ = testing.resnet_fixture()
    x, _ = synthetic_fundus(1, PATCH, seed=1234)
    eps = torch.randn(NS, 1, 32, generator=torch.Generator().manual_seed(1))
    cfg = O.ResNetCfg()
    with torch.no_grad():
        O.mc_patch_uncertainty(sd, cfg, x, TEMP, NS, eps)      # warm-up
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            O.mc_patch_uncertainty(sd, cfg, x, TEMP, NS, eps)
            ts.append(time.perf_counter() - t0)
    return 1.0 / statistics.median(ts), ts


def cpu_train_baseline(threads: int, batch: int = 2):
    """The reference training step (train.py:385-411: forward in train mode, loss + beta*KL, backward, clip, AdamW) through
    the oracle port on the host cores, ONE step at a reduced batch (per-image cost is batch-independent on the CPU)."""
    from oracle import vaeunet_oracle as O
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    torch.set_num_threads(threads)
    sd, _ = testing.resnet_fixture()
    x, m = synthetic_fundus(batch, PATCH, seed=77)
    eps = torch.randn(batch, 32, generator=torch.Generator().manual_seed(2))
    t0 = time.perf_counter()
    _, _, _, grads, _ = O.training_grads(sd, O.ResNetCfg(), x, m, eps, 0.001, 1e-3, "EX")
    params = {k: sd[k].clone() for k in grads}
    O.clip_grad_norm_(grads, 1.0)
    O.adamw_step(params, grads, {}, lr=1e-4, wd=1e-5)
    dt = time.perf_counter() - t0
    return batch / dt, dt


def gpu_eager_baseline(dev, batch: int = 4, reps: int = 3):
    """The cuDNN number the hand-written kernels must beat (SURVEY 2.2 / BASELINE.md 3): the oracle port of the reference
    (plain PyTorch ops) run on the SAME B200 - (a) fp32 with TF32 tensor cores allowed, (b) bf16 autocast; channels_last
    images; `restructured` = one encoder pass + N decodes (what this repo's engine does), `reference_loop` = the
    reference's own loop, which re-runs the encoder for every sample (visualize_vae.py:61-87 per z). patches/s."""
    from oracle import vaeunet_oracle as O
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    sd, _ = testing.resnet_fixture()
    sdc = {k: (v.to(dev).contiguous(memory_format=torch.channels_last) if v.dim() == 4 else v.to(dev)) for k, v in sd.items()}
    x, _ = synthetic_fundus(batch, PATCH, seed=4242)
    x = x.to(dev).contiguous(memory_format=torch.channels_last)
    eps = torch.randn(NS, batch, 32, generator=torch.Generator().manual_seed(3)).to(dev)
    cfg = O.ResNetCfg()
    out = {"batch": batch, "unit": "patches/s",
           "what": "oracle port (PyTorch eager -> cuDNN / cuBLAS) on the same B200, 512x512, T=1.0, N=10, uncertainty maps"}

    def restructured():
        return O.mc_patch_uncertainty(sdc, cfg, x, TEMP, NS, eps)

    def reference_loop():
        mu, lv = O.unetresnet_encode(sdc, x)
        segs = []
        for n in range(NS):
            z = mu + eps[n] * torch.exp(0.5 * lv) * TEMP
            segs.append(O.predict_full_image(sdc, cfg, x, z))
        return O.calculate_uncertainty_metrics(torch.stack(segs))

    def run(fn):
        with torch.no_grad():
            fn(); fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
        return batch * reps / (e0.elapsed_time(e1) / 1e3)

    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, torch.backends.cudnn.benchmark)
    try:
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cudnn.benchmark = True
        for name, fn in (("restructured", restructured), ("reference_loop", reference_loop)):
            try:
                out[f"tf32_{name}"] = run(fn)
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    out[f"bf16_autocast_{name}"] = run(fn)
            except Exception as e:                                        # the oracle is CPU test infrastructure first
                out[f"{name}_error"] = f"{type(e).__name__}: {e}"[:200]
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, torch.backends.cudnn.benchmark = old
    del sdc
    torch.cuda.empty_cache()
    return out


# ====================================================================================================== configs[2]: training
def bench_train(ctx: Ctx, steps: int, warmup: int, dtype_name: str):
    """configs[2]: VAE-UNet training step (CombinedLoss + beta*KL, clip 1.0, AdamW), batch 16 of 512x512 per GPU, 16-bit
    activations / fp32 master weights, data parallel with an NCCL all-reduce of the gradients (weak scaling).
    Default dtype f16 with a static 1024x loss scale: the reference trains under fp16 autocast + GradScaler, and it is the
    16-bit type whose gradients meet the tight parity bar (tests/test_training_gpu.py, tests/test_baseline_shapes_gpu.py)."""
    from vaeunet_b200 import ops, testing
    from vaeunet_b200.pipeline import StagedInput, StagedOutput
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.training import TrainStep
    from vaeunet_b200.unet import UNetResNet
    dev, rank, world, dist = ctx.dev, ctx.rank, ctx.world, ctx.dist
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw)
    model.load_state_dict(sd)
    model = model.to(dev).train()
    pg = None
    if dist is not None:
        from vaeunet_b200.distributed import gradient_group
        # 0 = the default communicator. A CTA-limited communicator was tried against SM contention with the persistent conv
        # kernels and lost: 2 GPUs 12.01 ms/step (default) vs 12.48 (2 CTAs) vs 15.04 (1 CTA), profiles/r2zv_nccl_ctas.log
        ctas = int(os.environ.get("VU_DP_NCCL_CTAS", "0"))
        pg = gradient_group(ctas) if ctas > 0 else dist.group.WORLD
    f16 = dtype_name == "f16"
    ts = TrainStep(model, lesion="EX", dtype=ops.F16 if f16 else ops.BF16, loss_scale=1024.0 if f16 else 1.0, process_group=pg)
    B, POOL = TRAIN_B, 4                       # 4 x (50 MB images + 17 MB masks) > L2; activations are GBs per step
    base, mask = synthetic_fundus(4, PATCH, seed=4321 + rank)
    base, mask = base.repeat(B // 4, 1, 1, 1), mask.repeat(B // 4, 1, 1, 1)
    host = [(base.roll(13 * i, dims=3).contiguous().pin_memory(), mask.roll(13 * i, dims=3).contiguous().pin_memory())
            for i in range(POOL)]
    devp = [(a.to(dev), b.to(dev)) for a, b in host]
    g = torch.Generator().manual_seed(7 + rank)
    eps = [torch.randn(B, 32, generator=g).to(dev) for _ in range(POOL)]
    inp, outp = StagedInput(dev, 2), StagedOutput(dev, 2)

    def step(i, e2e=False):
        j = i % POOL
        if e2e:      # images + masks from pinned host memory (next batch prefetched during this step), loss read back to the host
            x, m = inp.push(*host[j])
            inp.prefetch(*host[(i + 1) % POOL])
        else:
            x, m = devp[j]
        out = ts.step(x, m, 0.001, eps[j])
        if e2e:
            inp.release()
            outp.push({"loss": out["loss"].reshape(1)})
        return out

    def e2e_end():
        inp._pending = None
        torch.cuda.current_stream().wait_stream(outp.stream)

    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
    ms, launches = ctx.timed(lambda i: step(i), steps, warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), steps, warmup, pre_end=e2e_end)
    outp.flush()
    sustained, burst, hbm, src = _peaks()
    roof, _ = _conv_roofline(lambda i: step(i), 1, sustained, burst, src,
                             "tcgen05 conv kernels (forward + data gradient: conv_igemm / conv3x3_halo; weight gradient: conv_wgrad); "
                             "per-launch times overlap across the two streams of the step (weight gradients run on a side stream)",
                             clocks, "train")
    imgs = B * world * steps
    value = imgs / (ms / 1e3)
    step_tf = F_ALG_TRAIN_STEP_GF / (ms / steps)                               # GFLOP per ms == TFLOP/s (per GPU: weak scaling)
    roof["whole_step"] = {"alg_gflop_per_step": F_ALG_TRAIN_STEP_GF, "achieved_tflops_per_gpu": step_tf,
                          "frac_of_sustained": step_tf / sustained, "frac_of_burst": step_tf / burst}
    res = {
        "metric": "train_images_per_sec", "value": value, "unit": "img/s", "ms_per_step": ms / steps, "steps": steps, "warmup": warmup,
        "dtype": dtype_name, "scaling": "weak", "higher_is_better": True,
        "config": {"workload": "configs[2]: VAE-UNet (ResNet-34 encoder) training step, CombinedLoss + beta*KL (free bits 1e-3), "
                               f"clip 1.0, AdamW(1e-4, wd 1e-5); {dtype_name} activations / gradients"
                               + (" with a static 1024x loss scale" if f16 else "") + ", fp32 master weights",
                   "batch_per_gpu": B, "global_batch": B * world, "parallelism": f"dp{world}: bucketed NCCL all-reduce of 121.8 MB fp32 gradients"
                   if world > 1 else "dp1", "l2": "rotating pool of 4 input batches (268 MB) + multi-GB activations per step > 126 MB L2"},
        "clocks": clocks,
        "e2e": {"value": imgs / (ms_e2e / 1e3), "unit": "img/s", "ms_per_step": ms_e2e / steps,
                "h2d_bytes_per_step": B * 4 * PATCH * PATCH * 4, "d2h_bytes_per_step": 4},
        "gpu_launches": launches, "roofline": roof,
    }
    if rank == 0 and world == 1 and not ctx.args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, dt = cpu_train_baseline(threads)
        res["cpu_baseline"] = {"value": v, "unit": "img/s", "cores": threads, "kind": "port",
                               "sample": f"ONE training step (forward, loss, backward, clip, AdamW) at batch 2 of 512x512 through the "
                                         f"oracle port (fp32 PyTorch CPU autograd, {threads} threads): {dt:.1f} s"}
    del ts, model
    torch.cuda.empty_cache()
    return res


# ====================================================================================================== configs[1]: MC inference
def bench_mc(ctx: Ctx):
    from vaeunet_b200 import ops, testing
    from vaeunet_b200.engine import GraphedMC
    from vaeunet_b200.pipeline import StagedInput, StagedOutput
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.unet import UNetResNet
    args, dev, rank, world = ctx.args, ctx.dev, ctx.rank, ctx.world
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw)
    model.compute_dtype = ops.BF16 if args.dtype == "bf16" else ops.F16
    model.load_state_dict(sd)
    model = model.to(dev).eval()
    eng = model.engine()
    nmicro = max(1, args.batch // MICRO)
    B = nmicro * MICRO                        # patches per step per GPU
    POOL = 8                                  # rotating input pool: 8 x 16 x 3 MB = 403 MB > L2 (126 MB)
    base, _ = synthetic_fundus(MICRO, PATCH, seed=1234 + rank)
    host_pool = [(base.roll(17 * i, dims=3).contiguous()).pin_memory() for i in range(POOL)]
    dev_pool = [h.to(dev) for h in host_pool]
    g = torch.Generator().manual_seed(99 + rank)
    eps_pool = [torch.randn(NS, MICRO, 32, generator=g).to(dev) for _ in range(POOL)]
    OUTS = ("mean", "std", "entropy")          # configs[1]: mean / variance / entropy maps per patch
    graphed = None if not args.graph else GraphedMC(eng, MICRO, PATCH, PATCH, NS, TEMP, outs=OUTS)
    inp, outp = StagedInput(dev, 2), StagedOutput(dev, 2)

    def micro(k, e2e=False):
        j = k % POOL
        if e2e:
            # public streaming path: pinned host batch -> device on the copy-in stream (the next batch is prefetched while this
            # one computes), three maps -> pinned host buffers on the copy-out stream; all of it inside the timed region
            x, = inp.push(host_pool[j])
            inp.prefetch(host_pool[(k + 1) % POOL])
        else:
            x = dev_pool[j]
        if graphed is not None and ops.CONV_TIMING is None:
            res = graphed(x, eps_pool[j])
        else:
            with torch.no_grad():
                res = eng.mc_predict(x, TEMP, NS, eps_pool[j], outs=OUTS)
        if e2e:
            inp.release()
            outp.push({k2: res[k2] for k2 in OUTS})
        return res

    def step(i, e2e=False):
        for m in range(nmicro):
            micro(i * nmicro + m, e2e)

    def e2e_end():
        inp._pending = None                                       # drop the prefetch issued by the last step
        torch.cuda.current_stream().wait_stream(outp.stream)      # the last device->host copies are inside the timed region

    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
    ms, launches = ctx.timed(lambda i: step(i), args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), args.steps, args.warmup, pre_end=e2e_end)
    outp.flush()
    sustained, burst, hbm, src = _peaks()
    roof, ev = _conv_roofline(lambda i: micro(i), 2, sustained, burst, src,
                              "conv_igemm_pair_kernel<BN> / conv_igemm_kernel<BN> / conv3x3_halo_kernel (tcgen05 implicit GEMM, all conv layers); "
                              "per micro-batch of 16 patches", clocks, "mc")
    if args.layers and rank == 0:
        agg = {}
        for a, b, f, desc in ev:
            t = agg.setdefault(desc, [0, 0.0, 0.0])
            t[0] += 1; t[1] += a.elapsed_time(b); t[2] += f
        tot = sum(v[1] for v in agg.values())
        with open(args.layers, "w") as fh:
            fh.write(f"# per-layer conv timing over 2 micro-batches of {MICRO} patches, dtype {args.dtype} (CUDA events around each launch)\n")
            fh.write("shape,launches,ms_total,ms_share,tflops\n")
            for desc, (n, ms_l, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                fh.write(f"{desc},{n},{ms_l:.4f},{ms_l / tot:.4f},{fl / (ms_l / 1e3) / 1e12:.1f}\n")
    patches = B * world * args.steps
    value = patches / (ms / 1e3)
    per_gpu = value / world
    exec_gf_patch = roof["alg_gflop_per_step_convs"] / MICRO
    roof["whole_step"] = {
        "reference_equivalent_gflop_per_patch": F_ALG_PATCH_GF, "executed_conv_gflop_per_patch": exec_gf_patch,
        "achieved_tflops_reference_equivalent": F_ALG_PATCH_GF * per_gpu / 1e3, "achieved_tflops_executed": exec_gf_patch * per_gpu / 1e3,
        "frac_of_sustained": exec_gf_patch * per_gpu / 1e3 / sustained, "frac_of_burst": exec_gf_patch * per_gpu / 1e3 / burst,
        "frac_of_sustained_reference_equivalent": F_ALG_PATCH_GF * per_gpu / 1e3 / sustained,
        "note": "executed = FLOPs of the launched convs (the spatially constant latent channels are folded into border-class "
                "biases, weights of the gate's skip branch are shared by the N samples); reference-equivalent = the dense 761.40 "
                "GF/patch the reference executes for the same result"}
    out = {
        "metric": "mc_uncertainty_patches_per_sec", "value": value, "unit": "patches/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": "configs[1]: EX VAE-UNet (ResNet-34 encoder) MC inference, 512x512 patches, T=1.0, N=10, "
                               "mean/std/entropy maps per patch; fixture weights (seeded init + calibrated BN)",
                   "patches_per_step_per_gpu": B, "micro_batch": MICRO, "mc_samples": NS,
                   "parallelism": f"patch-sharded x{world}, no collective",
                   "l2": f"rotating pool of {POOL} input micro-batches ({POOL * MICRO * 3 * PATCH * PATCH * 4 / 1e6:.0f} MB) "
                         "+ multi-GB activation working set per micro-batch, both > 126 MB L2",
                   "timed_window_s": ms / 1e3},
        "clocks": clocks,
        "e2e": {"value": patches / (ms_e2e / 1e3), "unit": "patches/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": B * 3 * PATCH * PATCH * 4, "d2h_bytes_per_step": 3 * B * PATCH * PATCH * 4},
        "gpu_launches": launches, "roofline": roof,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, ts = cpu_baseline(threads)
        out["cpu_baseline"] = {"value": v, "unit": "patches/s", "cores": threads, "kind": "port",
                               "sample": f"median of {len(ts)} runs of ONE 512x512 patch, N=10 (oracle port of the reference, "
                                         f"fp32 PyTorch CPU, {threads} threads); run times {['%.2f' % t for t in ts]} s"}
        if not args.no_gpu_eager:
            out["gpu_eager_baseline"] = gpu_eager_baseline(dev)
    del graphed, eng, model, dev_pool, eps_pool
    torch.cuda.empty_cache()
    return out


# ====================================================================================================== other BASELINE configs
def _line(ctx, metric, unit, units_per_step, ms, ms_e2e, launches, clocks, workload, dtype, h2d, d2h, f_alg_gf_per_unit, extra=None):
    args, world = ctx.args, ctx.world
    sustained, burst, _, src = _peaks()
    total = units_per_step * world * args.steps
    value = total / (ms / 1e3)
    tf = f_alg_gf_per_unit * value / world / 1e3
    out = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
           "data": "synthetic", "config": workload, "clocks": clocks,
           "e2e": {"value": total / (ms_e2e / 1e3), "unit": unit, "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": d2h},
           "gpu_launches": launches,
           "roofline": {"bound": "tensor", "kernel": "whole step (all tcgen05 conv launches + memory-bound glue)", "achieved": tf,
                        "peak": sustained, "unit": "TFLOP/s", "frac": tf / sustained, "frac_of_burst": tf / burst, "peak_burst": burst,
                        "alg_gflop_per_unit": f_alg_gf_per_unit, "peak_source": f"{src} MEASURED_PEAKS.json", "traffic": None}}
    if extra:
        out.update(extra)
    return out


def bench_unet(ctx: Ctx):
    """basic attention U-Net (unet/unet_model.py:6-36) forward at 512x512, batch 8: images/s (394 GF/image)."""
    from vaeunet_b200 import ops, testing
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.unet import UNet
    args, dev = ctx.args, ctx.dev
    sd = testing.unet_fixture(bilinear=False)
    model = UNet(3, 1, False); model.load_state_dict(sd); model = model.to(dev).eval()
    model.compute_dtype = args.dtype
    B, POOL = 8, 8
    base, _ = synthetic_fundus(B, PATCH, seed=77 + ctx.rank)
    host = [base.roll(11 * i, dims=3).contiguous().pin_memory() for i in range(POOL)]
    devp = [h.to(dev) for h in host]
    hout = torch.empty((B, 1, PATCH, PATCH)).pin_memory()

    def step(i, e2e=False):
        with torch.no_grad():
            x = host[i % POOL].to(dev, non_blocking=True) if e2e else devp[i % POOL]
            y = model(x)
            if e2e:
                hout.copy_(y, non_blocking=True)

    sampler = ClockSampler(ctx.local); sampler.start() if ctx.rank == 0 else None
    ms, launches = ctx.timed(lambda i: step(i), args.steps, args.warmup)
    clocks = sampler.stop() if ctx.rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), args.steps, args.warmup)
    extra = None
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline:
        from oracle import vaeunet_oracle as O
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        xc = base[:1]
        with torch.no_grad():
            O.unet_forward(sd, xc, False)
            t0 = time.perf_counter(); O.unet_forward(sd, xc, False); dt = time.perf_counter() - t0
        extra = {"cpu_baseline": {"value": 1.0 / dt, "unit": "img/s", "cores": threads, "kind": "port",
                                  "sample": f"ONE 512x512 image through the oracle port of UNet.forward (fp32 PyTorch CPU): {dt:.2f} s"}}
    return _line(ctx, "unet_forward_images_per_sec", "img/s", B, ms, ms_e2e, launches, clocks,
                 {"workload": "basic attention UNet (ConvTranspose2d up-sampling) forward, 512x512, batch 8 per GPU, logits out",
                  "l2": "rotating pool of 8 input batches (201 MB) + GBs of activations > 126 MB L2"},
                 args.dtype, B * 3 * PATCH * PATCH * 4, B * PATCH * PATCH * 4, F_UNET_512_GF, extra)


def bench_cfg1(ctx: Ctx):
    """configs[0]: UNetResNet forward + N=3 MC latent samples at T=1.0 on ONE 3x256x256 patch (the reference's CPU-runnable
    case): this repo on the B200 vs the reference algorithm on the host cores, same call sequence."""
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.unet import UNetResNet
    from vaeunet_b200.utils.vae_utils import generate_predictions
    args, dev = ctx.args, ctx.dev
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).eval()
    model.compute_dtype = args.dtype
    POOL = 64
    base, _ = synthetic_fundus(1, 256, seed=256)
    host = [base.roll(3 * i, dims=3).contiguous().pin_memory() for i in range(POOL)]
    devp = [h.to(dev) for h in host]
    g = torch.Generator().manual_seed(5)
    e1 = torch.randn(1, 32, generator=g).to(dev); e3 = torch.randn(3, 1, 32, generator=g).to(dev)
    hout = torch.empty((1, 1, 256, 256)).pin_memory(); hpred = torch.empty((1, 1, 128, 128)).pin_memory()
    REP = 32                                  # calls per step (one call = 0.1 TFLOP: far too short to time alone)

    def step(i, e2e=False):
        with torch.no_grad():
            for r in range(REP):
                k = (i * REP + r) % POOL
                x = host[k].to(dev, non_blocking=True) if e2e else devp[k]
                out, mu, lv = model(x, eps=e1)
                pred = generate_predictions(model, x, TEMP, 3, eps=e3)
                if e2e:
                    hout.copy_(out, non_blocking=True); hpred.copy_(pred, non_blocking=True)

    sampler = ClockSampler(ctx.local); sampler.start() if ctx.rank == 0 else None
    ms, launches = ctx.timed(lambda i: step(i), args.steps, args.warmup)
    clocks = sampler.stop() if ctx.rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), args.steps, args.warmup)
    extra = None
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline:
        from oracle import vaeunet_oracle as O
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        cfg = O.ResNetCfg()
        ts = []
        with torch.no_grad():
            for _ in range(4):
                t0 = time.perf_counter()
                O.unetresnet_forward(sd, cfg, base, e1.cpu())
                O.generate_predictions(sd, cfg, base, TEMP, 3, e3.cpu())
                ts.append(time.perf_counter() - t0)
        dt = statistics.median(ts[1:])
        extra = {"cpu_baseline": {"value": 1.0 / dt, "unit": "calls/s", "cores": threads, "kind": "port",
                                  "sample": f"median of 3 runs of forward + generate_predictions(N=3) on one 256x256 patch (oracle port, "
                                            f"fp32 PyTorch CPU, {threads} threads): {dt * 1e3:.0f} ms"}}
    return _line(ctx, "cfg1_forward_plus_mc3_calls_per_sec", "calls/s", REP, ms, ms_e2e, launches, clocks,
                 {"workload": "configs[0]: VAE-UNet forward + generate_predictions(N=3, T=1.0) on one synthetic 3x256x256 patch per call",
                  "calls_per_step": REP, "l2": "64 rotating inputs; the working set of one call (~60 MB) fits L2 by nature of the config"},
                 args.dtype, REP * 3 * 256 * 256 * 4, REP * (256 * 256 + 128 * 128) * 4, F_CFG1_GF, extra)


def bench_cfg4(ctx: Ctx):
    """configs[3]: MA inference on 700x700 patches, T=2.0, N=15. Throughput line: a STREAM of patches, patch-sharded over the
    ranks with all 15 samples local (no collective: the natural sharding, weak scaling). `sample_sharded` (N > 1 ranks): the
    minimum-latency form for ONE patch - samples split 2,2,2,2,2,2,2,1, encoder replicated, one all-gather of the per-pixel
    moments + Chan merge - with its latency next to the single-GPU latency of the same patch."""
    from vaeunet_b200 import distributed as D
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    from vaeunet_b200.unet import UNetResNet
    args, dev, world = ctx.args, ctx.dev, ctx.world
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).eval()
    model.compute_dtype = args.dtype
    S, N, T, MB, NM, POOL = 700, 15, 2.0, 4, 16, 8      # 16 micro-batches of 4 patches per step
    base, _ = synthetic_fundus(MB, S, seed=700 + ctx.rank, lesion="MA")
    host = [base.roll(7 * i, dims=3).contiguous().pin_memory() for i in range(POOL)]
    devp = [h.to(dev) for h in host]
    g = torch.Generator().manual_seed(15 + ctx.rank)
    eps = [torch.randn(N, MB, 32, generator=g).to(dev) for _ in range(POOL)]
    OUTS = ("mean", "std", "entropy")
    hmaps = {k: torch.empty((MB, S, S)).pin_memory() for k in OUTS}

    def step(i, e2e=False):
        with torch.no_grad():
            for m in range(NM):
                k = (i * NM + m) % POOL
                x = host[k].to(dev, non_blocking=True) if e2e else devp[k]
                res = model.mc_predict(x, T, N, eps=eps[k], outs=OUTS)
                if e2e:
                    for kk in OUTS:
                        hmaps[kk].copy_(res[kk], non_blocking=True)

    sampler = ClockSampler(ctx.local); sampler.start() if ctx.rank == 0 else None
    ms, launches = ctx.timed(lambda i: step(i), args.steps, args.warmup)
    clocks = sampler.stop() if ctx.rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), args.steps, args.warmup)
    # single-patch latency: local (all samples on this GPU) vs sample-sharded over the ranks
    x1 = devp[0][:1].contiguous()
    e15 = torch.randn(N, 1, 32, generator=torch.Generator().manual_seed(4)).to(dev)
    lat = {}

    def lat_local(i):
        with torch.no_grad():
            model.mc_predict(x1, T, N, eps=e15, outs=OUTS)

    def lat_sharded(i):
        D.mc_predict_sample_sharded(model, x1, T, N, eps=e15, group=None)

    ms_l, _ = ctx.timed(lat_local, 20, 5)
    lat["one_patch_all_samples_local_ms"] = ms_l / 20
    if world > 1:
        ms_s, _ = ctx.timed(lat_sharded, 20, 5)
        lat["one_patch_samples_sharded_ms"] = ms_s / 20
        lat["split"] = D.shard_counts(N, world)
        lat["note"] = ("one 700^2 patch is 2 110 GF (~2-3 ms of work): split 8 ways the per-rank launch sequence (~70 kernels) and the "
                       "all-gather latency dominate; a stream of patches should be patch-sharded (the line's value)")
    extra = {"sample_sharded": lat}
    if ctx.rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import vaeunet_oracle as O
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        with torch.no_grad():
            t0 = time.perf_counter()
            O.mc_patch_uncertainty(sd, O.ResNetCfg(), base[:1], T, N, e15.cpu())
            dt = time.perf_counter() - t0
        extra["cpu_baseline"] = {"value": 1.0 / dt, "unit": "patches/s", "cores": threads, "kind": "port",
                                 "sample": f"ONE 700x700 patch, N=15, T=2.0 through the oracle port (fp32 PyTorch CPU): {dt:.2f} s"}
    return _line(ctx, "cfg4_mc_patches_per_sec", "patches/s", MB * NM, ms, ms_e2e, launches, clocks,
                 {"workload": "configs[3]: MA VAE-UNet MC inference, 700x700 patches (features 350/175/88/44/22), T=2.0, N=15, "
                              "mean/std/entropy maps", "patches_per_step_per_gpu": MB * NM, "micro_batch": MB,
                  "parallelism": f"patch-sharded x{world} (samples local), no collective",
                  "l2": "rotating pool of 8 micro-batches (188 MB) + GBs of activations > 126 MB L2"},
                 args.dtype, MB * NM * 3 * S * S * 4, 3 * MB * NM * S * S * 4, F_CFG4_PATCH_GF, extra)


def bench_cfg5(ctx: Ctx):
    """configs[4]: one full-scale 4288x2848 fundus image tiled into 77 patches of 512 (overlap 100), N=10: images/s.
    The 8-bit image goes to the device once (36.6 MB); tiles are cut by a gather kernel; every tile batch is blended by one
    launch; N ranks deal the tiles round-robin and exchange the per-sample numerators with ONE reduce-scatter by row strips
    (488 MB fp32 in total) + an all-gather of the maps (distributed.tiled_image_mc)."""
    from vaeunet_b200 import distributed as D
    from vaeunet_b200 import ops, testing
    from vaeunet_b200.testing import f4_inputs
    from vaeunet_b200.unet import UNetResNet
    args, dev, world = ctx.args, ctx.dev, ctx.world
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).eval()
    model.compute_dtype = args.dtype
    H, W, N = 2848, 4288, 10
    img_u8, _ = f4_inputs(4288, H, W)
    host = torch.from_numpy(img_u8).pin_memory()
    dev_u8 = host.to(dev)
    eps = torch.randn(N, 1, 32, generator=torch.Generator().manual_seed(9)).to(dev)
    zero = torch.zeros((1, 2), dtype=torch.int32, device=dev)
    OUTS = ("mean", "std", "entropy")
    hmaps = {k: torch.empty((H, W)).pin_memory() for k in OUTS}
    t_comm = {}

    def step(i, e2e=False):
        u8 = host.to(dev, non_blocking=True) if e2e else dev_u8
        img = ops.tile_gather(u8, zero, H, W, 255.0)                     # uint8 HWC -> fp32 [1,3,H,W] on the device
        res = D.tiled_image_mc(model, img, N, 512, 100, 1.0, batch_size=args.tile_batch, eps=eps, group=None)
        if e2e and ctx.rank == 0:
            # one image = ONE set of result maps: every rank holds them after the all-gather, rank 0 hands them to the host
            # (all ranks copying 147 MB each over a shared host link doubled the step at 8 GPUs: 15.0 -> 29.4 ms)
            for k in OUTS:
                hmaps[k].copy_(res[k], non_blocking=True)

    sampler = ClockSampler(ctx.local); sampler.start() if ctx.rank == 0 else None
    ms, launches = ctx.timed(lambda i: step(i), args.steps, args.warmup)
    clocks = sampler.stop() if ctx.rank == 0 else None
    ms_e2e, _ = ctx.timed(lambda i: step(i, True), args.steps, args.warmup)
    extra = {}
    if world > 1:                       # the exchange step alone: reduce-scatter of [Hp, N, W] + [Hp, W] fp32 and the map all-gather
        rows = D.strip_rows(H, world)
        num = torch.zeros((rows * world, N, W), device=dev); den = torch.zeros((rows * world, W), device=dev)

        def comm(i):
            a = D.reduce_scatter_rows(num); b = D.reduce_scatter_rows(den)
            D.all_gather_stack(torch.zeros((5, rows, W), device=dev))

        ms_c, _ = ctx.timed(comm, 10, 3)
        extra["exchange"] = {"reduce_scatter_plus_all_gather_ms": ms_c / 10, "reduce_scatter_bytes_total": rows * world * (N + 1) * W * 4,
                             "tiles_per_rank": [len(D.deal_round_robin(77, r, world)) for r in range(world)],
                             "note": "the full-image encoder pass for mu / logvar (1 784 GF) runs on rank 0 only: the serial part of an image"}
    if ctx.rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import vaeunet_oracle as O
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        x = torch.from_numpy(img_u8[1168:1680, 1888:2400].astype("float32") / 255.0).permute(2, 0, 1).unsqueeze(0)
        with torch.no_grad():
            t0 = time.perf_counter()
            O.mc_patch_uncertainty(sd, O.ResNetCfg(), x, 1.0, N, eps.cpu())
            dt = time.perf_counter() - t0
        est = dt * 77 * (F_CFG5_IMAGE_GF / (77 * F_ALG_PATCH_GF))
        extra["cpu_baseline"] = {"value": 1.0 / est, "unit": "images/s", "cores": threads, "kind": "port",
                                 "sample": f"ONE of the 77 tiles (512x512, N=10) through the oracle port: {dt:.2f} s; scaled by the image's "
                                           f"60 412 GF / 761.4 GF per tile -> {est:.0f} s per image (extrapolated, stated as such)"}
    return _line(ctx, "cfg5_tiled_images_per_sec", "images/s", 1.0 / world, ms, ms_e2e, launches, clocks,
                 {"workload": "configs[4]: UNetResNet on one synthetic 4288x2848 fundus image, 77 tiles of 512 (overlap 100), N=10, "
                              "mean/std/entropy maps of the whole image", "tile_batch": args.tile_batch,
                  "parallelism": f"tiles dealt round-robin over {world} rank(s); reduce-scatter by row strips + all-gather" if world > 1
                  else "1 GPU", "scaling_note": "STRONG scaling: one image is split over the ranks (value = images/s of the job)",
                  "l2": "one image = 146 MB fp32 + 488 MB of per-sample accumulators > 126 MB L2"},
                 args.dtype, H * W * 3, 3 * H * W * 4, F_CFG5_IMAGE_GF, dict(extra, scaling="strong"))


# ====================================================================================================== reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    from oracle import vaeunet_oracle as O
    from vaeunet_b200 import testing
    from vaeunet_b200.synthetic import synthetic_fundus
    sd, _ = testing.resnet_fixture()
    cfg = O.ResNetCfg()
    x, _ = synthetic_fundus(1, PATCH, seed=1234)
    eps = torch.randn(NS, 1, 32, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        for _ in range(max(1, min(args.warmup, 2))):
            O.mc_patch_uncertainty(sd, cfg, x, TEMP, NS, eps)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            O.mc_patch_uncertainty(sd, cfg, x, TEMP, NS, eps)
        dt = time.perf_counter() - t0
    v = args.steps / dt
    sample = f"{args.steps} steps x 1 synthetic 512x512 patch, N=10, fp32 PyTorch CPU (oracle port of the reference algorithm)"
    print(json.dumps({
        "impl": "reference", "metric": "mc_uncertainty_patches_per_sec", "value": v, "unit": "patches/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: EX VAE-UNet MC inference, 512x512 patches, T=1.0, N=10, mean/std/entropy maps",
                   "patches_per_step": 1},
        "cpu_baseline": {"value": v, "unit": "patches/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "patches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=384, help="MC workload: patches per step per GPU (micro-batches of 16)")
    ap.add_argument("--train-steps", type=int, default=200, help="timed optimizer steps of the 'train' object (>= 2.5 s window)")
    ap.add_argument("--train-dtype", default="f16", choices=["bf16", "f16"])
    ap.add_argument("--tile-batch", type=int, default=8, help="cfg5: tiles per network call")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-eager", action="store_true", help="skip the cuDNN-eager leg of the MC line")
    ap.add_argument("--layers", default=None, help="write a per-conv-shape timing table (csv) to this path")
    ap.add_argument("--dtype", default="f16", choices=["bf16", "f16"])
    ap.add_argument("--graph", action="store_true", help="replay the MC micro-batch from a CUDA graph (engine.GraphedMC)")
    ap.add_argument("--workload", default="both", choices=["both", "mc", "train", "unet", "cfg1", "cfg4", "cfg5"],
                    help="both: MC line + 'train' object (the driver's run); the others print that workload as the line's metric")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    args.warmup = max(args.warmup, 3)
    ctx = Ctx(args)
    if args.workload == "train":
        res = bench_train(ctx, args.steps, args.warmup, args.train_dtype)
        res.update({"n_gpus": ctx.world, "vs_baseline": None, "data": "synthetic"})
    elif args.workload in ("unet", "cfg1", "cfg4", "cfg5"):
        res = {"unet": bench_unet, "cfg1": bench_cfg1, "cfg4": bench_cfg4, "cfg5": bench_cfg5}[args.workload](ctx)
    else:
        res = bench_mc(ctx)
        if args.workload == "both":
            res["train"] = bench_train(ctx, args.train_steps, args.warmup, args.train_dtype)
    if ctx.rank == 0:
        print(json.dumps(res))
    ctx.close()


if __name__ == "__main__":
    main()
