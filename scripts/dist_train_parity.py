"""Data-parallel training step over NCCL (run under torchrun on N GPUs of one box):
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 scripts/dist_train_parity.py
Every rank trains on its own batch. Checks (max over ranks): (1) after the bucketed all-reduce the flat gradient equals
the SUM over ranks of the local gradients (computed by a twin TrainStep without a process group + one plain all-reduce);
(2) after the optimizer step every rank holds identical parameters; (3) times the step with CUDA events."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, S = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4, 128)


def make(pg):
    sd, kw = testing.resnet_fixture()
    m = UNetResNet(**kw)
    m.load_state_dict(sd)
    return TrainStep(m.to(dev).train(), process_group=pg)


x, m = synthetic_fundus(B, S, S, seed=100 + rank)
x, m = x.to(dev), m.to(dev)
eps = torch.randn(B, 32, generator=torch.Generator().manual_seed(5 + rank)).to(dev)
ts_dp, ts_local = make(dist.group.WORLD), make(None)
from vaeunet_b200.training import BUCKET_ORDER
# local gradients without any collective, then the SAME buffer through the bucketed all-reduce path vs one plain all-reduce
pg, ts_dp.pg = ts_dp.pg, None
ts_dp.zero_grad(); ts_dp.forward_backward(x, m, 0.001, eps)
ts_dp.pg = pg
want = ts_dp.flat_g.clone()
dist.all_reduce(want)
for b in BUCKET_ORDER:
    ts_dp._start_all_reduce(b)
ts_dp._wait_all_reduce()
torch.cuda.synchronize()
err = ((ts_dp.flat_g - want).norm() / want.norm()).reshape(1)
dist.all_reduce(err, op=dist.ReduceOp.MAX)
# run-to-run spread of two separate backward passes on the same data (split-K atomics reorder fp32 sums; bf16 amplifies)
ts_local.zero_grad(); ts_local.forward_backward(x, m, 0.001, eps)
g1 = ts_local.flat_g.clone()
ts_local.zero_grad(); ts_local.forward_backward(x, m, 0.001, eps)
spread = ((ts_local.flat_g - g1).norm() / g1.norm()).reshape(1)
dist.all_reduce(spread, op=dist.ReduceOp.MAX)
# gradient accumulation under DP (train.py:398,406: accumulation_steps = 2): only the LAST micro-batch communicates
# (sync=False on the others), and the result equals one plain all-reduce of the accumulated local gradients
x2, m2 = x.roll(7, dims=3).contiguous(), m.roll(7, dims=3).contiguous()
ts_dp.zero_grad()
ts_dp.forward_backward(x, m, 0.001, eps, accum=2, sync=False)
ts_dp.forward_backward(x2, m2, 0.001, eps, accum=2, sync=True)
ts_dp._wait_all_reduce()
torch.cuda.synchronize()
got_acc = ts_dp.flat_g.clone()
ts_local.zero_grad()
ts_local.forward_backward(x, m, 0.001, eps, accum=2)
ts_local.forward_backward(x2, m2, 0.001, eps, accum=2)
want_acc = ts_local.flat_g.clone()
dist.all_reduce(want_acc)
err_acc = ((got_acc - want_acc).norm() / want_acc.norm()).reshape(1)
dist.all_reduce(err_acc, op=dist.ReduceOp.MAX)
ts_dp.optimizer_step()
p0 = ts_dp.flat_p.clone()
dist.broadcast(p0, 0)
same = (ts_dp.flat_p - p0).abs().max().reshape(1)
dist.all_reduce(same, op=dist.ReduceOp.MAX)
# timing (bucketed vs one all-reduce at the end)
res = {}
for name, bucketed in (("bucketed", True), ("single", False)):
    ts_dp.bucket_all_reduce = bucketed
    for _ in range(3):
        ts_dp.step(x, m, 0.001, eps)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ts_dp.step(x, m, 0.001, eps)
    e1.record()
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 5], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[name] = t.item()
for _ in range(3):
    ts_local.step(x, m, 0.001, eps)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ts_local.step(x, m, 0.001, eps)
e1.record(); torch.cuda.synchronize()
if rank == 0:
    ok = err.item() < 1e-6 and same.item() == 0.0 and err_acc.item() < 3.0 * spread.item() + 1e-6
    print(f"DP + accumulation (2 micro-batches, sync on the last only) vs plain all-reduce of the accumulated local gradients: "
          f"rel {err_acc.item():.1e} (budget 3 x run-to-run spread {spread.item():.1e})", flush=True)
    print(f"DP training world {world} B={B}x{S}x{S} per rank: bucketed all-reduce vs one plain all-reduce of the same local gradients rel {err.item():.1e} "
          f"(run-to-run spread of two backward passes {spread.item():.1e}); "
          f"max |param - rank0 param| after the step {same.item():.1e} {'PASS' if ok else 'FAIL'}; "
          f"step {res['bucketed']:.2f} ms bucketed / {res['single']:.2f} ms single all-reduce / {e0.elapsed_time(e1) / 5:.2f} ms no DP", flush=True)
dist.destroy_process_group()
