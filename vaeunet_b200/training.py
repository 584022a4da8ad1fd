"""The VAE-UNet training step on B200 (reference train.py:385-411).

    seg, mu, logvar = model(images)                       # train mode: batch-statistics BatchNorm
    loss = criterion(seg, masks) + beta * kl_with_free_bits(mu, logvar, free_bits)
    loss.backward(); clip_grad_norm_(params, 1.0); AdamW.step()

`TrainStep` runs exactly that for the reference's default UNetResNet configuration (ResNet-34 encoder,
latent_injection='all', attention gates, skips) without autograd: an explicit forward that keeps what the
backward needs, and an explicit backward in reverse order. Every arithmetic operation is one of this repo's
kernels: tcgen05 implicit-GEMM convolutions for forward / data gradient (`vu_conv2d_fwd` on dY with rotated or
parity filter banks) / weight gradient (`vu_conv2d_wgrad_oihw`), and vectorised memory-bound kernels for
train-mode BatchNorm, pooling, bilinear resize, the attention gate pieces, the loss and AdamW.

Storage: fp32 master parameters live in ONE flat buffer (the module's nn.Parameters are views of it, so
`state_dict()` is unchanged), gradients in a second flat buffer (`p.grad` views; one memset, one squared-norm
reduction, one AdamW launch, and — data parallel — bucketed NCCL all-reduces of contiguous ranges that overlap the
rest of the backward). Activations and activation gradients are NHWC 16-bit (bf16 by default: the reference's
fp16 autocast needs a GradScaler, bf16 does not).

Spatially constant maps (z_spatial, z_initial(z), z_proj(z), W_g of a constant map) stay per-image
vectors; their train-mode BatchNorm statistics over B*H*W pixels equal the statistics over the B rows, and their
gradients are per-image sums over pixels.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops
from . import train_ops as T
from ._lib import VaeUnetError
from .engine import RESNET34_STAGES
from .ops import Act
from ._nvtx import rng

import os

# BatchNorm forward statistics accumulated in the epilogue of the producing conv (VU_FUSE_BN_STATS=0: separate pass)
FUSE_BN_STATS = os.environ.get("VU_FUSE_BN_STATS", "1") != "0"
# Weight gradients on a second CUDA stream (VU_TRAIN_STREAMS=0: everything on the current stream). They are leaves of the
# backward (only the optimizer reads them), so they overlap the BatchNorm-backward / data-gradient chain instead of
# sitting inside it: tensor-pipe work under the memory-bound kernels of the chain.
WGRAD_STREAM = os.environ.get("VU_TRAIN_STREAMS", "1") != "0"
_SKIP_WGRAD = os.environ.get("VU_DEBUG_SKIP_WGRAD", "0") == "1"

_ALIGN = 64   # floats: every parameter starts on a 256-byte boundary of the flat buffers


BUCKET_ORDER = ("decoder", "latent", "layer4", "layer3", "rest")     # the order the backward finishes them
_BUCKET_PREFIXES = {
    "decoder": ("decoder_blocks.", "final_conv."),
    "latent": ("mu_head.", "logvar_head.", "z_initial."),
    "layer4": ("encoder.layer4.",),
    "layer3": ("encoder.layer3.",),
    "rest": ("encoder.conv1.", "encoder.bn1.", "encoder.layer1.", "encoder.layer2."),
}


def param_layout(named_shapes):
    """Flat-buffer layout of the parameters (host logic, no device): every parameter starts on a 256-byte boundary, in
    `named_parameters()` order. -> (offsets {name: float offset}, total floats, buckets {name: (lo, hi)}): the buckets are
    contiguous, disjoint ranges that cover the whole buffer; each is all-reduced as soon as the backward has finished it."""
    offs, total = {}, 0
    for n, shape in named_shapes:
        k = 1
        for d in shape:
            k *= int(d)
        offs[n] = total
        total += (k + _ALIGN - 1) // _ALIGN * _ALIGN
    names = [n for n, _ in named_shapes]
    ends = {n: (offs[names[i + 1]] if i + 1 < len(names) else total) for i, n in enumerate(names)}
    buckets = {}
    for b, prefixes in _BUCKET_PREFIXES.items():
        ks = [n for n in names if n.startswith(prefixes)]
        if ks:
            buckets[b] = (min(offs[n] for n in ks), max(ends[n] for n in ks))
    return offs, total, buckets


class _Conv:
    """A tensor-core convolution of the model: live 16-bit filter bank + data-gradient banks (views of one flat 16-bit
    buffer, repacked from the fp32 masters by one batched launch per step)."""
    __slots__ = ("name", "k", "stride", "w16", "banks_t", "banks", "n_fwd", "n_bank")

    def __init__(self, name, w, k, stride):
        O, I = w.shape[:2]
        self.name, self.k, self.stride = name, k, stride
        nb = 4 if (stride == 2 and k == 3) else 1
        self.n_fwd = O * I * k * k
        self.n_bank = nb * self.n_fwd

    def bind(self, flat16, off_fwd, off_bank, w):
        O, I = w.shape[:2]
        k = self.k
        nb = self.n_bank // self.n_fwd
        self.w16 = flat16[off_fwd:off_fwd + self.n_fwd].view(O, k * k, I)
        self.banks_t = flat16[off_bank:off_bank + self.n_bank].view(nb, I, k * k, O)
        self.banks = [self.banks_t[b] for b in range(nb)]


class TrainStep:
    """One optimisation step of train.py for `UNetResNet`.

    step(images [B,3,H,W] fp32 NCHW on the device, masks [B,1,H,W], beta, eps [B,L]) -> dict of device scalars
    (loss, recon, kl, grad_norm); nothing synchronises with the host.
    """

    def __init__(self, model, lr: float = 1e-4, weight_decay: float = 1e-5, betas=(0.9, 0.999), adam_eps: float = 1e-8,
                 max_grad_norm: float = 1.0, lesion: str = "EX", free_bits: float = 1e-3, dtype=ops.BF16,
                 loss_scale: float = 1.0, process_group=None, bucket_all_reduce: bool = True):
        ops.require_device()
        if not (model.latent_injection == "all" and model.use_attention and model.use_skip and model.n_classes == 1):
            raise VaeUnetError("TrainStep covers the reference's default UNetResNet (latent_injection='all', attention, "
                               "skips, n_classes=1)")
        self.model = model
        self.dev = next(model.parameters()).device
        if self.dev.type != "cuda":
            raise VaeUnetError("TrainStep: move the model to the B200 first (no CPU fallback)")
        self.dt = dtype
        self.lr, self.wd, self.betas, self.adam_eps = lr, weight_decay, betas, adam_eps
        self.max_grad_norm, self.lesion, self.free_bits = max_grad_norm, lesion, free_bits
        self.loss_scale = float(loss_scale)
        self.pg = process_group
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
        self.bucket_all_reduce = bucket_all_reduce
        self.step_count = 0
        self.L = model.latent_dim
        self._flatten()
        self._register_convs()
        self.sumsq = torch.zeros(1, dtype=torch.float64, device=self.dev)
        self._pending = []
        self.side = torch.cuda.Stream(device=self.dev) if WGRAD_STREAM else None
        self._side_keep = []        # operands of weight gradients in flight on the side stream (kept alive until the join)
        self._side_used = False
        self._sync_grads = True
        self._reduced = False       # the flat gradient already holds an all-reduced sum (no further local accumulation)

    # ------------------------------------------------------------------ parameter storage
    def _flatten(self):
        named = [(n, p) for n, p in self.model.named_parameters()]
        offs, total, self.buckets = param_layout([(n, tuple(p.shape)) for n, p in named])
        dev = self.dev
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.P: Dict[str, torch.Tensor] = {}
        self.G: Dict[str, torch.Tensor] = {}
        self.offs = offs
        for n, p in named:
            o, k = offs[n], p.numel()
            view = self.flat_p[o:o + k].view(p.shape)
            view.copy_(p.data.float())
            p.data = view
            g = self.flat_g[o:o + k].view(p.shape)
            p.grad = g
            self.P[n], self.G[n] = view, g
        self.Bf = dict(self.model.named_buffers())
        self.nbt = [b for n, b in self.Bf.items() if n.endswith("num_batches_tracked")]

    def _register_convs(self):
        self.convs: Dict[str, _Conv] = {}
        for n, w in self.P.items():
            if not n.endswith(".weight") or w.dim() != 4:
                continue
            k = w.shape[2]
            if n in ("encoder.conv1.weight", "final_conv.weight") or n.startswith(("mu_head", "logvar_head", "z_initial")):
                continue
            if ".z_proj." in n or ".psi." in n or n == "decoder_blocks.0.attention.W_g.0.weight":
                continue                                      # vector layers / fp32 per-pixel dot products
            stride = 1
            if n.startswith("encoder.layer"):
                li, b = int(n.split(".")[1][5:]), int(n.split(".")[2])
                if li > 1 and b == 0 and (".conv1." in n or ".downsample." in n):
                    stride = 2
            self.convs[n[:-len(".weight")]] = _Conv(n, w, k, stride)
        ws = self.P["encoder.conv1.weight"]
        kp = (ws[0].numel() + 7) // 8 * 8
        self.stem_w16 = torch.empty((ws.shape[0], 1, kp), dtype=ops._TORCH_DT[self.dt], device=self.dev)
        import struct
        a16 = lambda v: (v + 127) // 128 * 128                  # 256-byte aligned 16-bit offsets (TMA base alignment)
        tot16, recs = 0, []
        for name, c in self.convs.items():
            off_fwd = tot16; tot16 += a16(c.n_fwd)
            off_bank = tot16; tot16 += a16(c.n_bank)
            w = self.P[c.name]
            recs.append((c, off_fwd, off_bank, struct.pack("<qqqiiii", self.offs[c.name], off_fwd, off_bank, w.shape[0],
                                                           w.shape[1], c.k, c.stride)))
        self.flat16 = torch.zeros(tot16, dtype=ops._TORCH_DT[self.dt], device=self.dev)
        for c, off_fwd, off_bank, _ in recs:
            c.bind(self.flat16, off_fwd, off_bank, self.P[c.name])
        self.pack_table = torch.frombuffer(bytearray(b"".join(r[3] for r in recs)), dtype=torch.uint8).to(self.dev)
        # The wgrad kernel accumulates in its packed layout [O][k*k][I] (contiguous 16-byte vector reductions); one batched
        # launch per all-reduce bucket transposes those into the parameters' OIHW .grad views.
        total = 0
        goff = {}
        for name, c in self.convs.items():
            goff[name] = total
            total += (self.P[c.name].numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.flat_gp = torch.zeros(total, dtype=torch.float32, device=self.dev)
        # one fp64 (sum, sum of squares) pair per output channel of every conv that feeds a BatchNorm
        self.stat_ws = torch.zeros(2 * (64 + sum(c.w16.shape[0] for c in self.convs.values())), dtype=torch.float64,
                                   device=self.dev)
        self._stat_off = 0
        self.GP = {}
        for name, c in self.convs.items():
            O, I = self.P[c.name].shape[:2]
            self.GP[name] = self.flat_gp[goff[name]:goff[name] + O * I * c.k * c.k].view(O, c.k * c.k, I)
        import struct
        self.unpack_tables = {}
        for bname, (lo, hi) in self.buckets.items():
            recs, max_row = [], 1
            for name, c in self.convs.items():
                o = self.offs[c.name]
                if lo <= o < hi:
                    O, I = self.P[c.name].shape[:2]
                    recs.append(struct.pack("<qqiiii", goff[name], o, O, c.k * c.k, I, 0))
                    max_row = max(max_row, c.k * c.k * I)
            if recs:
                tab = torch.frombuffer(bytearray(b"".join(recs)), dtype=torch.uint8).to(self.dev)
                self.unpack_tables[bname] = (tab, len(recs), max_row)

    def repack(self):
        """fp32 master filters -> 16-bit tensor-core banks (forward + data-gradient): one batched launch per step."""
        ops.check(ops.load().vu_pack_weights_batched(self.flat_p.data_ptr(), self.flat16.data_ptr(), self.pack_table.data_ptr(),
                                                     len(self.convs), self.dt, ops._stream()), "vu_pack_weights_batched")
        ws = self.P["encoder.conv1.weight"]
        T.pack_rows_padded(ws, self.stem_w16.shape[2], self.dt, out=self.stem_w16)
        ops._count(2)

    # ------------------------------------------------------------------ small helpers
    def _stat_slot(self, cout: int) -> torch.Tensor:
        """fp64 [2*cout] slice of the per-step statistics workspace (zeroed once at the start of forward())."""
        o = self._stat_off
        self._stat_off = o + 2 * cout
        assert self._stat_off <= self.stat_ws.numel(), "statistics workspace too small"
        return self.stat_ws[o:o + 2 * cout]

    def _conv(self, x: Act, name: str, bias: Optional[torch.Tensor] = None, out: Optional[Act] = None,
              stats: bool = False) -> Act:
        """stats: the conv epilogue also accumulates the BatchNorm statistics of its output (picked up by _bn)."""
        c = self.convs[name]
        sums = self._stat_slot(c.w16.shape[0]) if stats and FUSE_BN_STATS else None
        y = ops.conv2d(x, c.w16, c.k, c.stride, c.k // 2, None, bias, False, out=out, stat_sums=sums)
        y.stat_sums = sums
        return y

    def _bn(self, x_raw: Act, bn: str, relu: bool, residual: Optional[Act] = None, out: Optional[Act] = None):
        P, Bf = self.P, self.Bf
        y, mean, invstd = T.bn_train_fwd(x_raw, P[bn + "weight"], P[bn + "bias"], Bf[bn + "running_mean"],
                                         Bf[bn + "running_var"], relu, residual, out,
                                         sums=getattr(x_raw, "stat_sums", None))
        return y, (mean, invstd)

    def _bn_bwd(self, x_raw: Act, dy: Act, y_relu: Optional[Act], st, bn: str, want_dz=False, dx: Optional[Act] = None,
                mask_from_x: bool = False):
        """mask_from_x: the layer was BN + ReLU with no residual -> the ReLU mask is recomputed from x_raw (y is not read)."""
        dx, _, _, dz = T.bn_train_bwd(x_raw, dy, None if mask_from_x else y_relu, st[0], st[1], self.P[bn + "weight"], want_dz,
                                      dx, self.G[bn + "weight"], self.G[bn + "bias"],
                                      relu_beta=self.P[bn + "bias"] if mask_from_x else None)
        return dx, dz

    def _wgrad(self, x: Act, dy: Act, name: str):
        c = self.convs[name]
        if _SKIP_WGRAD:                # timing experiment only (scripts/train_breakdown.py): the main chain without its leaves
            return
        if self.side is None:
            ops.conv2d_wgrad(x, dy, c.k, c.stride, c.k // 2, out=self.GP[name])
            return
        # fork: the side stream waits for everything enqueued so far (dy is final), then runs the weight gradient
        # concurrently with the rest of the backward chain
        main = torch.cuda.current_stream()
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            ops.conv2d_wgrad(x, dy, c.k, c.stride, c.k // 2, out=self.GP[name])
        self._side_keep.append((x, dy))
        self._side_used = True

    def _join_side(self):
        """The current stream waits for the weight gradients in flight (before their packed buffer is read)."""
        if self.side is not None and self._side_used:
            torch.cuda.current_stream().wait_stream(self.side)
            self._side_keep.clear()     # safe: later allocations are ordered after the wait on this stream
            self._side_used = False

    def _dgrad(self, dy: Act, name: str, H: int, W: int, out: Optional[Act] = None, accumulate=False) -> Act:
        c = self.convs[name]
        return ops.conv2d_dgrad(dy, c.banks, c.k, c.stride, H, W, out=out, accumulate=accumulate)

    def _vec(self, z, lin: str, bn: str, relu: bool, count: float):
        P, Bf = self.P, self.Bf
        w = P[lin + "weight"]
        return T.vec_layer_fwd(z, w.view(w.shape[0], -1), P[lin + "bias"], P[bn + "weight"], P[bn + "bias"],
                               Bf[bn + "running_mean"], Bf[bn + "running_var"], relu, count)

    def _vec_bwd(self, dy, saved, lin: str, bn: str, relu: bool, dz):
        P, G = self.P, self.G
        w = P[lin + "weight"]
        T.vec_layer_bwd(dy, saved, w.view(w.shape[0], -1), P[bn + "weight"], relu, dz, True,
                        dw=G[lin + "weight"], db=G[lin + "bias"], dg=G[bn + "weight"], dbeta=G[bn + "bias"])

    # ------------------------------------------------------------------ forward (train mode)
    def forward(self, images: torch.Tensor, eps: torch.Tensor):
        P, dt, dev, L = self.P, self.dt, self.dev, self.L
        images = images.float().contiguous()
        B, _, H, W = images.shape
        S: dict = {"images": images, "B": B, "H": H, "W": W}
        # stem 7x7 s2 on the tensor cores: im2col patches (147 -> 152 columns) x 1x1 filter bank
        xcol = T.im2col_small_cin(images, 7, 2, 3, dt)
        self.stat_ws.zero_()
        self._stat_off = 0
        sums0 = self._stat_slot(self.stem_w16.shape[0]) if FUSE_BN_STATS else None
        x0raw = ops.conv2d(xcol, self.stem_w16, 1, 1, 0, None, None, False, stat_sums=sums0)
        x0raw.stat_sums = sums0
        S["xcol"] = xcol
        f0, st0 = self._bn(x0raw, "encoder.bn1.", True)
        pool, pool_code = T.maxpool_idx_fwd(f0, 3, 2, 1)
        S.update(x0raw=x0raw, f0=f0, st0=st0, pool_code=pool_code)
        x = pool
        feats: List[Act] = [f0]
        blocks = []
        for li, nb in enumerate(RESNET34_STAGES, start=1):
            for b in range(nb):
                q = f"encoder.layer{li}.{b}."
                r = {"q": q, "x": x}
                if (q + "downsample.0") in self.convs:
                    r["idraw"] = self._conv(x, q + "downsample.0", stats=True)
                    idt, r["st_ds"] = self._bn(r["idraw"], q + "downsample.1.", False)
                else:
                    idt = x
                r["y1raw"] = self._conv(x, q + "conv1", stats=True)
                r["y1"], r["st1"] = self._bn(r["y1raw"], q + "bn1.", True)
                r["y2raw"] = self._conv(r["y1"], q + "conv2", stats=True)
                r["out"], r["st2"] = self._bn(r["y2raw"], q + "bn2.", True, residual=idt)
                x = r["out"]
                blocks.append(r)
            feats.append(x)
        S["blocks"], S["feats"] = blocks, feats
        f4 = feats[4]
        # latent heads + reparameterisation (unet_resnet.py:205-217)
        wm, wl = P["mu_head.0.weight"], P["logvar_head.0.weight"]
        eps = eps.reshape(1, B, L).float().contiguous()
        mu, lv, z = ops.latent_head(f4, wm.view(L, -1), P["mu_head.0.bias"], wl.view(L, -1), P["logvar_head.0.bias"], eps, 1.0, 1)
        z = z[0]
        S.update(mu=mu, lv=lv, z=z, eps=eps)
        # z_initial on the constant map (unet_resnet.py:221-229)
        xv, S["zi"] = self._vec(z, "z_initial.0.", "z_initial.1.", True, float(B * f4.H * f4.W))
        x: object = xv
        dec = []
        for i in range(4):
            p = f"decoder_blocks.{i}."
            skip = feats[3 - i]
            Hs, Ws = skip.H, skip.W
            const = i == 0
            cx = xv.shape[1] if const else x.C
            cs = skip.C
            cat = ops.new_act(B, Hs, Ws, cx + cs + L, dt, dev)
            xs = cat.slice(0, cx)
            r = {"p": p, "x": x, "skip": skip, "cat": cat, "cx": cx, "cs": cs, "const": const}
            a = p + "attention."
            if const:
                ops.broadcast_vec(xv, xs)
                g1v, r["g_saved"] = self._vec(xv, a + "W_g.0.", a + "W_g.1.", False, float(B * Hs * Ws))
                g1n = ops.broadcast_vec(g1v, ops.new_act(B, Hs, Ws, g1v.shape[1], dt, dev))
            else:
                ops.bilinear_ac(x, Hs, Ws, out=xs)
                # W_g is a 1x1 conv + bias and bilinear weights sum to 1: W_g(up(x)) + b == up(W_g x + b), so the GEMM (and, by
                # linearity, its weight / data gradients) runs at the LOW resolution of x: a quarter of the pixels
                g1low = self._conv(x, a + "W_g.0", bias=P[a + "W_g.0.bias"])
                r["g1raw"] = ops.bilinear_ac(g1low, Hs, Ws)
                g1n, r["st_g"] = self._bn(r["g1raw"], a + "W_g.1.", False)
            r["x1raw"] = self._conv(skip, a + "W_x.0", bias=P[a + "W_x.0.bias"], stats=True)
            r["s"], r["st_x"] = self._bn(r["x1raw"], a + "W_x.1.", True, residual=g1n)        # relu(g1 + x1)
            r["psi_raw"] = ops.final_conv1(r["s"], P[a + "psi.0.weight"].view(-1), 0.0, bias_dev=P[a + "psi.0.bias"])
            r["psi"], r["psi_stat"] = T.map_bn_sigmoid_fwd(r["psi_raw"], P[a + "psi.1.weight"], P[a + "psi.1.bias"],
                                                           self.Bf[a + "psi.1.running_mean"], self.Bf[a + "psi.1.running_var"])
            T.scale_by_map(skip, r["psi"], cat.slice(cx, cs))
            zv, r["zp_saved"] = self._vec(z, p + "z_proj.0.", p + "z_proj.1.", True, float(B * Hs * Ws))
            ops.broadcast_vec(zv, cat.slice(cx + cs, L))
            r["y1raw"] = self._conv(cat, p + "conv1.0", stats=True)
            r["y1"], r["st1"] = self._bn(r["y1raw"], p + "conv1.1.", True)
            r["y2raw"] = self._conv(r["y1"], p + "conv2.0", stats=True)
            r["y2"], r["st2"] = self._bn(r["y2raw"], p + "conv2.1.", True)
            x = r["y2"]
            dec.append(r)
        S["dec"] = dec
        lg_low = ops.final_conv1(x, P["final_conv.weight"].view(-1), 0.0, bias_dev=P["final_conv.bias"])
        logits = ops.resize_logits(lg_low, H, W, ops.SIG_NONE)
        S["logits"] = logits
        torch._foreach_add_(self.nbt, 1)
        self._invalidate_packed()          # the running statistics moved: eval-mode folded BN caches are stale
        return logits.view(B, 1, H, W), mu, lv, S

    # ------------------------------------------------------------------ loss (utils/loss.py; train.py:312-316,387-398)
    def loss_and_grads(self, S, masks: torch.Tensor, beta: float, accum: int = 1):
        logits = S["logits"]
        t = masks.reshape(logits.shape).float().contiguous()
        if self.lesion == "MA":
            w_bce, w_dice, w_focal, alpha, gamma = 0.0, 0.5, 0.5, 0.9, 2.0
        else:
            w_bce, w_dice, w_focal, alpha, gamma = 0.5, 0.5, 0.0, 0.8, 2.0
        sums = ops.seg_loss_sums(logits, t, False, alpha, gamma)
        gs = self.loss_scale / accum
        dlogits = ops.seg_loss_bwd(logits, t, sums, w_bce, w_dice, w_focal, alpha, gamma, 1.0, gs)
        kl, dmu, dlv = ops.kl_free_bits(S["mu"], S["lv"], self.free_bits, want_grad=True, grad_scale=beta * gs)
        n = float(logits.numel())
        Pm, Tm = torch.clamp(sums[0], min=0.5), torch.clamp(sums[1], min=0.5)
        dice = 1.0 - (2.0 * sums[2] + 1.0) / (Pm + Tm + 1.0)
        recon = (w_bce * sums[3] / n + w_dice * dice + w_focal * sums[4] / n).float()
        return {"loss": recon + beta * kl, "recon": recon, "kl": kl}, dlogits, dmu, dlv

    # ------------------------------------------------------------------ backward
    def backward(self, S, dlogits: torch.Tensor, dmu: torch.Tensor, dlv: torch.Tensor):
        P, G, dt, dev, L = self.P, self.G, self.dt, self.dev, self.L
        B = S["B"]
        feats, dec = S["feats"], S["dec"]
        self.flat_gp.zero_()
        y2 = dec[3]["y2"]
        dlow = T.resize_map_bwd(dlogits, y2.H, y2.W)
        G["final_conv.weight"].view(-1).add_(T.weighted_channel_sum(y2, dlow, per_image=False))
        G["final_conv.bias"].add_(T.map_sum(dlow))
        dy = T.map_expand(dlow, P["final_conv.weight"].view(-1), None, ops.new_act(B, y2.H, y2.W, y2.C, dt, dev))
        dz = torch.zeros((B, L), dtype=torch.float32, device=dev)
        dfeat: List[Optional[Act]] = [None] * 5
        dxv = None
        for i in (3, 2, 1, 0):
            r = dec[i]
            p, a = r["p"], r["p"] + "attention."
            cat, cx, cs, skip = r["cat"], r["cx"], r["cs"], r["skip"]
            Hs, Ws = skip.H, skip.W
            dy2raw, _ = self._bn_bwd(r["y2raw"], dy, r["y2"], r["st2"], p + "conv2.1.", mask_from_x=True)
            self._wgrad(r["y1"], dy2raw, p + "conv2.0")
            dy1 = self._dgrad(dy2raw, p + "conv2.0", Hs, Ws)
            dy1raw, _ = self._bn_bwd(r["y1raw"], dy1, r["y1"], r["st1"], p + "conv1.1.", mask_from_x=True)
            self._wgrad(cat, dy1raw, p + "conv1.0")
            dcat = self._dgrad(dy1raw, p + "conv1.0", Hs, Ws)
            # z_proj slice: per-image sums -> vector layer backward (accumulates into dz)
            dzv = T.weighted_channel_sum(dcat.slice(cx + cs, L), None, per_image=True)
            self._vec_bwd(dzv, r["zp_saved"], p + "z_proj.0.", p + "z_proj.1.", True, dz)
            # attention gate (unet_resnet.py:24-29)
            dsa = dcat.slice(cx, cs)
            dskip = T.scale_by_map(dsa, r["psi"], ops.new_act(B, Hs, Ws, cs, dt, dev))
            dpsi = T.channel_dot(dsa, skip)
            dpsi_raw, _, _ = T.map_bn_sigmoid_bwd(r["psi_raw"], dpsi, r["psi"], r["psi_stat"], P[a + "psi.1.weight"],
                                                  G[a + "psi.1.weight"], G[a + "psi.1.bias"])
            s = r["s"]
            G[a + "psi.0.weight"].view(-1).add_(T.weighted_channel_sum(s, dpsi_raw, per_image=False))
            # psi.0.bias, W_x.0.bias, W_g.0.bias feed a train-mode BatchNorm: their gradient is identically zero
            ds = T.map_expand(dpsi_raw, P[a + "psi.0.weight"].view(-1), s, ops.new_act(B, Hs, Ws, s.C, dt, dev))
            dx1raw, dg1n = self._bn_bwd(r["x1raw"], ds, s, r["st_x"], a + "W_x.1.", want_dz=True)
            self._wgrad(skip, dx1raw, a + "W_x.0")
            self._dgrad(dx1raw, a + "W_x.0", Hs, Ws, out=dskip, accumulate=True)
            dxs = dcat.slice(0, cx)
            if r["const"]:
                dg1v = T.weighted_channel_sum(dg1n, None, per_image=True)
                dxv = T.weighted_channel_sum(dxs, None, per_image=True)
                self._vec_bwd(dg1v, r["g_saved"], a + "W_g.0.", a + "W_g.1.", False, dxv)
            else:
                dg1raw, _ = self._bn_bwd(r["g1raw"], dg1n, None, r["st_g"], a + "W_g.1.")
                xin = r["x"]
                dg1low = T.bilinear_ac_bwd(dg1raw, xin.H, xin.W)          # adjoint of the upsample, then W_g at low resolution
                self._wgrad(xin, dg1low, a + "W_g.0")
                dy = T.bilinear_ac_bwd(dxs, xin.H, xin.W)
                self._dgrad(dg1low, a + "W_g.0", xin.H, xin.W, out=dy, accumulate=True)
            dfeat[3 - i] = dskip
        # z_initial, reparameterisation, heads
        self._vec_bwd(dxv, S["zi"], "z_initial.0.", "z_initial.1.", True, dz)
        self._all_reduce_bucket("decoder")
        T.reparam_bwd(dz.view(1, B, L), S["eps"], S["lv"], 1.0, dmu, dlv)
        f4 = feats[4]
        hw = float(f4.H * f4.W)
        gap = T.weighted_channel_sum(f4, None, per_image=True) / hw                # [B, 512]
        dgap = torch.zeros_like(gap)
        for head, dh in (("mu_head.0.", dmu), ("logvar_head.0.", dlv)):
            w = P[head + "weight"]
            T.vec_layer_bwd(dh, (gap, dh, dh, None, None), w.view(L, -1), None, False, dgap, True,
                            dw=G[head + "weight"], db=G[head + "bias"])
        dfeat[4] = ops.broadcast_vec(dgap / hw, ops.new_act(B, f4.H, f4.W, f4.C, dt, dev))
        self._all_reduce_bucket("latent")
        # encoder, last block first
        blocks = S["blocks"]
        dout = dfeat[4]
        stage_end = {}
        idx = 0
        for li, nb in enumerate(RESNET34_STAGES, start=1):
            idx += nb
            stage_end[idx - 1] = li
        for bi in range(len(blocks) - 1, -1, -1):
            r = blocks[bi]
            q, x = r["q"], r["x"]
            if bi in stage_end and stage_end[bi] < 4:
                T.add_(dout, dfeat[stage_end[bi]])                      # this block's output is also a decoder skip
            dy2raw, dz_id = self._bn_bwd(r["y2raw"], dout, r["out"], r["st2"], q + "bn2.", want_dz=True)
            self._wgrad(r["y1"], dy2raw, q + "conv2")
            dy1 = self._dgrad(dy2raw, q + "conv2", r["y1"].H, r["y1"].W)
            dy1raw, _ = self._bn_bwd(r["y1raw"], dy1, r["y1"], r["st1"], q + "bn1.", mask_from_x=True)
            self._wgrad(x, dy1raw, q + "conv1")
            if "idraw" in r:
                dx = self._dgrad(dy1raw, q + "conv1", x.H, x.W)
                didraw, _ = self._bn_bwd(r["idraw"], dz_id, None, r["st_ds"], q + "downsample.1.")
                self._wgrad(x, didraw, q + "downsample.0")
                T.add_(dx, self._dgrad(didraw, q + "downsample.0", x.H, x.W))
            else:
                dx = self._dgrad(dy1raw, q + "conv1", x.H, x.W, out=dz_id, accumulate=True)
            dout = dx
            if q == "encoder.layer4.0.":
                self._all_reduce_bucket("layer4")
            elif q == "encoder.layer3.0.":
                self._all_reduce_bucket("layer3")
        # stem: max-pool, BN + ReLU, 7x7 conv
        f0 = S["f0"]
        df0 = T.maxpool_idx_bwd(S["pool_code"], dout, f0.H, f0.W, 3, 2, 1)
        T.add_(df0, dfeat[0])
        dx0raw, _ = self._bn_bwd(S["x0raw"], df0, f0, S["st0"], "encoder.bn1.", mask_from_x=True)
        xcol = S["xcol"]
        dw = ops.conv2d_wgrad(xcol, dx0raw, 1, 1, 0)                                   # [64, 1, 152] packed
        gw = G["encoder.conv1.weight"]
        gw.view(gw.shape[0], -1).add_(dw.view(gw.shape[0], -1)[:, :gw[0].numel()])
        self._all_reduce_bucket("rest")

    # ------------------------------------------------------------------ data parallel (NCCL all-reduce of gradients)
    def _all_reduce_bucket(self, name: str):
        """A bucket of parameters has its final gradients: unpack its conv weight gradients, then start its all-reduce."""
        self._join_side()
        if name in self.unpack_tables:
            tab, n, max_row = self.unpack_tables[name]
            ops.check(ops.load().vu_unpack_wgrad_batched(self.flat_gp.data_ptr(), self.flat_g.data_ptr(), tab.data_ptr(), n,
                                                         max_row, ops._stream()), "vu_unpack_wgrad_batched")
            ops._count()
        self._start_all_reduce(name)

    def _start_all_reduce(self, name: str):
        """Asynchronous NCCL all-reduce (sum) of one contiguous range of the flat gradient; the 1/world average is folded
        into the AdamW kernel's gradient scale. With gradient accumulation only the LAST micro-batch communicates
        (`forward_backward(sync=False)` = DDP's no_sync): the buckets then hold the local sum over all micro-batches."""
        if self.pg is None or self.world == 1 or not self._sync_grads:
            return
        import torch.distributed as dist
        if not self.bucket_all_reduce:
            if name == "rest":
                self._pending.append(dist.all_reduce(self.flat_g, group=self.pg, async_op=True))
            return
        lo, hi = self.buckets[name]
        self._pending.append(dist.all_reduce(self.flat_g[lo:hi], group=self.pg, async_op=True))

    def _wait_all_reduce(self):
        for w in self._pending:
            w.wait()
        self._pending = []

    # ------------------------------------------------------------------ the step
    def zero_grad(self):
        self._wait_all_reduce()
        self.flat_g.zero_()
        self._reduced = False

    def forward_backward(self, images, masks, beta: float, eps: torch.Tensor, accum: int = 1, sync: bool = True):
        """Gradients of one (micro-)batch accumulated into the flat gradient buffer (train.py:387-403).
        accum: number of micro-batches per optimizer step (the loss is divided by it, train.py:398).
        sync: data parallel only - pass False for every micro-batch but the last one of an accumulation window; the
        gradient all-reduces then start on the last micro-batch only, on the accumulated local gradients."""
        if self._reduced:
            raise VaeUnetError("forward_backward after the gradients of this window were all-reduced: call optimizer_step() "
                               "(or zero_grad()) first, and pass sync=False for all micro-batches but the last")
        self._sync_grads = bool(sync)
        self._reduced = bool(sync) and self.pg is not None and self.world > 1
        with rng("vaeunet.train.forward"):
            self.repack()
            _, _, _, S = self.forward(images, eps)
        with rng("vaeunet.train.loss"):
            out, dlogits, dmu, dlv = self.loss_and_grads(S, masks, beta, accum)
        with rng("vaeunet.train.backward"):
            self.backward(S, dlogits, dmu, dlv)
        return out

    def optimizer_step(self):
        """clip_grad_norm_(max_norm) + AdamW (train.py:405-408) on the flat buffers; averages over the DP group."""
        with rng("vaeunet.train.optimizer"):
            self._optimizer_step()

    def _optimizer_step(self):
        self._wait_all_reduce()
        self._reduced = False
        self.step_count += 1
        gscale = 1.0 / (self.loss_scale * self.world)
        T.sumsq(self.flat_g, self.sumsq, True)
        T.adamw_step(self.flat_p, self.flat_g, self.flat_m, self.flat_v, self.lr, self.betas, self.adam_eps, self.wd,
                     self.step_count, self.sumsq if self.max_grad_norm > 0 else None, self.max_grad_norm, gscale)
        self._invalidate_packed()

    def _invalidate_packed(self):
        for m in self.model.modules():
            m.__dict__.pop("_unit_cache", None)

    def step(self, images, masks, beta: float, eps: Optional[torch.Tensor] = None):
        B = images.shape[0]
        if eps is None:
            eps = torch.randn(B, self.L, device=self.dev)
        self.zero_grad()
        out = self.forward_backward(images, masks, beta, eps)
        self.optimizer_step()
        out["grad_norm"] = (self.sumsq.sqrt() / (self.loss_scale * self.world)).float()
        return out
