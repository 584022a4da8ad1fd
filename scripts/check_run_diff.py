"""Debug: record the outputs of every ops / train_ops call of two identical training steps and list where they diverge."""
import sys, types
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vaeunet_b200 import ops, train_ops as T
from test_training_gpu import _setup

B, S = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
O, sd, model, ts, x, m, eps = _setup("EX", dev, ops.BF16, 1.0, B, S)
rec = []

def snap(v):
    if isinstance(v, ops.Act):
        return v.t[..., v.coff:v.coff + v.C].clone()
    if isinstance(v, torch.Tensor):
        return v.clone()
    if isinstance(v, (tuple, list)):
        return [snap(u) for u in v]
    if isinstance(v, dict):
        return [snap(u) for u in v.values()]
    return None

def wrap(mod, name):
    f = getattr(mod, name)
    def g(*a, **k):
        pre = snap([list(a), list(k.values())])
        r = f(*a, **k)
        post = snap([list(a), list(k.values())])
        shp = [tuple(u.t.shape) + (u.coff, u.C) if isinstance(u, ops.Act) else tuple(u.shape) if isinstance(u, torch.Tensor) else u for u in a]
        rec.append((f"{mod.__name__.split('.')[-1]}.{name}", str(shp)[:150], snap(r), post, pre))
        return r
    setattr(mod, name, g)

skip = {"check", "load", "require_device", "new_act", "_count", "_stream", "_ptr", "_f32", "_npix", "_sums"}
for mod in (T, ops):
    for name, f in list(vars(mod).items()):
        if isinstance(f, types.FunctionType) and f.__module__ == mod.__name__ and name not in skip and not name.startswith("_"):
            wrap(mod, name)

def flat(v, out):
    if isinstance(v, torch.Tensor):
        out.append(v)
    elif isinstance(v, list):
        for u in v:
            flat(u, out)
    return out

runs, gs = [], []
for run in range(4):
    rec.clear()
    ts.zero_grad()
    out = ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev))
    torch.cuda.synchronize()
    runs.append(list(rec))
    gs.append(ts.flat_g.clone())
gd = [((gs[0] - g).norm() / gs[0].norm()).item() for g in gs]
print("gradient diff vs run 0:", gd)
other = max(range(4), key=lambda j: gd[j])
print(len(runs[0]), len(runs[other]), "calls; comparing run 0 with run", other)
shown = 0
for i, (a, b) in enumerate(zip(runs[0], runs[other])):
    def worst_of(xa, xb):
        worst = 0.0
        for u, v in zip(flat(xa, []), flat(xb, [])):
            if u.shape == v.shape and u.numel():
                worst = max(worst, (u.double() - v.double()).norm().item() / (v.double().norm().item() + 1e-30))
        return worst
    w_in, w_out = worst_of(a[4], b[4]), max(worst_of(a[2], b[2]), worst_of(a[3], b[3]))
    if max(w_in, w_out) > 1e-6 and not a[0].endswith(('bn_train_fwd', 'vec_layer_fwd', 'map_bn_sigmoid_fwd')):
        print(f"{i:4d} {a[0]:28s} inputs {w_in:.3e} outputs {w_out:.3e}  {a[1][:60]}")
        shown += 1
        if shown >= 25:
            break
