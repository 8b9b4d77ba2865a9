"""Batched-client functional networks — the torch backend (CPU + oracle).

The reference trains its ``no_models`` clients one after another through a single ``local_model`` object
(image_train.py:21-32).  Here the K clients of a rank are a *batch dimension*: their parameters are rows of a flat
slab ``P[K, Dp]`` (buffers ``S[K, Db]``), activations are ``[B, K*C, H, W]`` and every convolution is one grouped
convolution with ``groups=K``; linears are ``bmm``.  Clients never mix, so autograd on the slab yields the K
independent gradients in one ``[K, Dp]`` tensor.  Ragged clients (different batch sizes / step counts) are handled
by a per-sample validity mask that enters the BatchNorm statistics and the loss.

This module is the framework's own engine on flat buffers (SURVEY §7.1 "CPU backend"), not the reference's code;
the native sm_100a executor implements the same networks with hand-written kernels and is tested against it.
Architectures: models/resnet_cifar.py:67-104, models/resnet_tinyimagenet.py:122-238, models/MnistNet.py:7-31,
models/loan_model.py:10-27 in the reference.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .. import config as C
from .spec import ParamSpec

BN_MOMENTUM = 0.1
BN_EPS = 1e-5


class BatchedNet:
    def __init__(self, kind: str, spec: ParamSpec) -> None:
        self.kind = kind
        self.spec = spec
        self.has_bn = spec.Db > 0
        self.has_dropout = kind == C.TYPE_LOAN
        self.dropout_p = 0.5          # LoanNet's nn.Dropout(0.5); 0 disables it (deterministic tests)

    # -------------------------------------------------------------------------------------------- public
    def forward(self, P: torch.Tensor, S: Optional[torch.Tensor], cnt: Optional[torch.Tensor], x: torch.Tensor,
                training: bool, shared_input: bool, mask: Optional[torch.Tensor] = None,
                active: Optional[torch.Tensor] = None, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        """Returns logits ``[K, B, n_classes]``.

        P [K, Dp]; S [K, Db] (BN running stats, updated in place when ``training``); cnt [K, n_counters] int64.
        x: ``[B, ...]`` when ``shared_input`` (all K models see the same samples) else ``[K, B, ...]``.
        mask [K, B] bool = which samples are real (None = all); active [K] bool = which clients take this step
        (inactive clients' BN statistics / counters are left untouched).
        """
        ctx = _Ctx(self.spec, P, S, cnt, training, mask, active, generator)
        ctx.dropout_p = self.dropout_p
        if self.kind == C.TYPE_LOAN:
            return self._loan(ctx, x, shared_input)
        if self.kind == C.TYPE_MNIST:
            return self._mnist(ctx, x, shared_input)
        return self._resnet(ctx, x, shared_input, tiny=self.kind == C.TYPE_TINYIMAGENET)

    # --------------------------------------------------------------------------------------------- nets
    def _resnet(self, c: "_Ctx", x, shared, tiny: bool):
        head = "fc" if tiny else "linear"
        skip = "downsample" if tiny else "shortcut"
        y = c.conv_in(x, "conv1.weight", None, stride=2 if tiny else 1, pad=3 if tiny else 1, shared=shared)
        y = F.relu(c.bn(y, "bn1"))
        if tiny:
            y = F.max_pool2d(y, 3, 2, 1)
        for li in range(1, 5):
            for bi in range(2):
                pre = f"layer{li}.{bi}"
                stride = 2 if (li > 1 and bi == 0) else 1
                z = F.relu(c.bn(c.conv(y, f"{pre}.conv1.weight", None, stride, 1), f"{pre}.bn1"))
                z = c.bn(c.conv(z, f"{pre}.conv2.weight", None, 1, 1), f"{pre}.bn2")
                if f"{pre}.{skip}.0.weight" in c.spec.entries:
                    y = c.bn(c.conv(y, f"{pre}.{skip}.0.weight", None, stride, 0), f"{pre}.{skip}.1")
                y = F.relu(z + y)
        y = F.avg_pool2d(y, 4) if not tiny else F.adaptive_avg_pool2d(y, 1)
        return c.linear(c.to_kbf(y), f"{head}.weight", f"{head}.bias")

    def _mnist(self, c: "_Ctx", x, shared):
        y = F.max_pool2d(F.relu(c.conv_in(x, "conv1.weight", "conv1.bias", 1, 0, shared)), 2, 2)
        y = F.max_pool2d(F.relu(c.conv(y, "conv2.weight", "conv2.bias", 1, 0)), 2, 2)
        y = F.relu(c.linear(c.to_kbf(y), "fc1.weight", "fc1.bias"))
        y = c.linear(y, "fc2.weight", "fc2.bias")
        return F.log_softmax(y, dim=-1)

    def _loan(self, c: "_Ctx", x, shared):
        K = c.K
        y = x.unsqueeze(0).expand(K, *x.shape) if shared else x
        y = F.relu(c.dropout(c.linear(y, "layer1.0.weight", "layer1.0.bias")))
        y = F.relu(c.dropout(c.linear(y, "layer2.0.weight", "layer2.0.bias")))
        return c.linear(y, "layer3.0.weight", "layer3.0.bias")


class _Ctx:
    """One forward pass' worth of bound state (slabs, mode, masks)."""

    def __init__(self, spec, P, S, cnt, training, mask, active, generator) -> None:
        self.spec, self.P, self.S, self.cnt = spec, P, S, cnt
        self.training, self.mask, self.active, self.gen = training, mask, active, generator
        self.K = P.shape[0]

    def w(self, name):
        return self.spec.view(self.P, name)

    def buf(self, name):
        e = self.spec.entries[name]
        o = e.offset - self.spec.Dp
        return self.S[:, o:o + e.numel]

    # first conv: x either shared [B, Cin, H, W] or per client [K, B, Cin, H, W] -> [B, K*Cout, Ho, Wo]
    def conv_in(self, x, wname, bname, stride, pad, shared):
        w = self.w(wname)
        K, Cout = w.shape[:2]
        b = self.w(bname).reshape(-1) if bname else None
        if shared:
            return F.conv2d(x, w.reshape(K * Cout, *w.shape[2:]), b, stride, pad)
        Kx, B = x.shape[:2]
        xg = x.permute(1, 0, 2, 3, 4).reshape(B, Kx * x.shape[2], *x.shape[3:])
        return F.conv2d(xg, w.reshape(K * Cout, *w.shape[2:]), b, stride, pad, groups=K)

    def conv(self, x, wname, bname, stride, pad):
        w = self.w(wname)
        K, Cout = w.shape[:2]
        b = self.w(bname).reshape(-1) if bname else None
        if x.is_cuda and K > 1:
            # library grouped convolutions are pathologically slow for some shapes (stride 2); this torch path is
            # only the oracle / fallback on GPUs, so K plain convolutions are fine
            Cin = w.shape[2]
            outs = [F.conv2d(x[:, k * Cin:(k + 1) * Cin], w[k], None if b is None else b[k * Cout:(k + 1) * Cout],
                             stride, pad) for k in range(K)]
            return torch.cat(outs, 1)
        return F.conv2d(x, w.reshape(K * Cout, *w.shape[2:]), b, stride, pad, groups=K)

    def bn(self, x, prefix):
        K = self.K
        B, KC, H, W = x.shape
        Cn = KC // K
        g = self.w(prefix + ".weight").reshape(1, KC, 1, 1)
        b = self.w(prefix + ".bias").reshape(1, KC, 1, 1)
        rm, rv = self.buf(prefix + ".running_mean"), self.buf(prefix + ".running_var")
        if not self.training:
            scale = g * torch.rsqrt(rv.reshape(1, KC, 1, 1) + BN_EPS)
            return (x - rm.reshape(1, KC, 1, 1)) * scale + b
        if self.mask is None:
            n = torch.full((K,), float(B * H * W), device=x.device)
            mean = x.mean(dim=(0, 2, 3))
            var = x.var(dim=(0, 2, 3), unbiased=False)
        else:
            m = self.mask.t().to(x.dtype)                       # [B, K]
            n = (m.sum(0) * (H * W)).clamp_(min=1.0)            # [K]
            mk = m.reshape(B, K, 1, 1, 1)
            xv = x.reshape(B, K, Cn, H, W)
            mean = ((xv * mk).sum(dim=(0, 3, 4)) / n[:, None]).reshape(KC)
            var = ((((xv - mean.reshape(1, K, Cn, 1, 1)) ** 2) * mk).sum(dim=(0, 3, 4)) / n[:, None])
            # a client with no valid sample in this step (idle slot of the lock-step schedule) must not blow its
            # activations up by 1/sqrt(eps) per layer: normalise it with unit variance (its output is unused)
            var = torch.where((m.sum(0) > 0)[:, None], var, torch.ones_like(var)).reshape(KC)
        with torch.no_grad():
            act = self.active if self.active is not None else torch.ones(K, dtype=torch.bool, device=x.device)
            a = act[:, None].to(x.dtype)
            unb = (n / (n - 1).clamp(min=1.0))[:, None]
            mom = BN_MOMENTUM * a
            rm.mul_(1 - mom).add_(mom * mean.detach().reshape(K, Cn))
            rv.mul_(1 - mom).add_(mom * var.detach().reshape(K, Cn) * unb)
            if self.cnt is not None and self.cnt.numel():
                ci = self.spec.counter_index(prefix + ".num_batches_tracked")
                self.cnt[:, ci] += act.long()
        xhat = (x - mean.reshape(1, KC, 1, 1)) * torch.rsqrt(var.reshape(1, KC, 1, 1) + BN_EPS)
        return xhat * g + b

    def to_kbf(self, x):
        """[B, K*C, H, W] -> [K, B, C*H*W] (the per-model flatten order is C,H,W as in ``x.view(B, -1)``)."""
        B, KC, H, W = x.shape
        return x.reshape(B, self.K, KC // self.K, H, W).permute(1, 0, 2, 3, 4).reshape(self.K, B, -1)

    def linear(self, x, wname, bname):
        w, b = self.w(wname), self.w(bname)
        return torch.baddbmm(b.unsqueeze(1), x, w.transpose(1, 2))

    def dropout(self, x, p: float = None):
        p = getattr(self, "dropout_p", 0.5) if p is None else p
        if not self.training or p <= 0.0:
            return x
        keep = torch.rand(x.shape, device=x.device, generator=self.gen) >= p
        return x * keep.to(x.dtype) / (1.0 - p)
