"""Torch backend: the engine's compute interface implemented with plain torch ops.

Runs on CPU (BASELINE config 1, and every non-GPU test) and is the fp32 oracle for the native sm_100a backend
(``cuda_backend.py``), which implements the same interface with hand-written kernels.  Interface:

  train(schedule, W, cnt, anchor, want_grad_sum) -> TrainResult     K clients, T lock-step SGD steps
  evaluate(models, split, trig_row)              -> (loss_sum[M], correct[M], n)
  + the flat elementwise ops (scale / norms)

Semantics reproduced (reference file:line):
  * SGD(momentum, weight_decay), fresh per client per round: image_train.py:33-35,63-65  (g += wd*w; v = mu*v + g;
    w -= lr*v; first step v = g)
  * poisoned batches = first ``poisoning_per_batch`` samples get trigger + ``poison_label_swap`` (Q13):
    image_helper.py:298-326, loan_train.py:99-107
  * poison loss ``alpha*CE + (1-alpha)*||w - w_anchor||_2`` over parameters: image_train.py:85-90, helper.py:111-123
    (Q25: the reference back-propagates the distance term even when alpha==1, where it contributes exactly 0)
  * FoolsGold client feature = sum over all steps of the raw gradients (before weight decay): image_train.py:94-100
  * logged train loss = sum of batch-mean losses / n_samples (Q6): image_train.py:221,247
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from .. import config as C
from ..models import BatchedNet, ParamSpec
from .schedule import RoundSchedule


@dataclass
class TrainResult:
    grad_sum: Optional[torch.Tensor]   # [K, Dp] or None
    loss_sum: torch.Tensor             # [K, E_max] sum of batch-mean losses
    correct: torch.Tensor              # [K, E_max]
    count: torch.Tensor                # [K, E_max] samples seen
    poison_count: torch.Tensor         # [K, E_max] poisoned samples seen
    nan_flag: torch.Tensor             # [K] bool: non-finite logits seen (replaces LoanNet's per-forward host sync)
    step_loss: Optional[torch.Tensor] = None   # [K, T] per-batch loss   (only when vis_train_batch_loss)
    step_dist: Optional[torch.Tensor] = None   # [K, T] distance to the round's start model after each batch
                                               # (only when batch_track_distance; image_train.py:107-117,225-236)


class TorchBackend:
    name = "torch"

    def __init__(self, params, spec: ParamSpec, data, device: torch.device) -> None:
        self.params, self.spec, self.data, self.device = params, spec, data, device
        self.kind = params["type"]
        self.net = BatchedNet(self.kind, spec)
        self.is_image = params.is_image
        self.label_swap = int(params["poison_label_swap"])
        self.eval_chunk = 512 if device.type == "cpu" else 2048
        self.launches = 0  # native kernel launches (none on this backend)

    # ------------------------------------------------------------------------------------------ batches
    def _gather(self, ids: torch.Tensor, train: bool) -> Tuple[torch.Tensor, torch.Tensor]:
        d = self.data
        xs, ys = (d.train_x, d.train_y) if train else (d.test_x, d.test_y)
        safe = ids.clamp(min=0)
        x = xs[safe]
        y = ys[safe]
        if self.is_image:
            x = d.to_float_nchw(x)
        else:
            x = x.float()
        return x, y

    def _poison_(self, x: torch.Tensor, y: torch.Tensor, sel: torch.Tensor, trig_row: torch.Tensor) -> None:
        """In place.  x [K, B, ...], y [K, B], sel [K, B] bool which samples to poison, trig_row [K]."""
        d = self.data
        if self.is_image:
            m = d.trigger_masks[trig_row]                                    # [K, H, W]
            full = sel[:, :, None, None, None] & m[:, None, None, :, :]       # [K, B, 1, H, W]
            x.masked_fill_(full.expand_as(x), 1.0)
        else:
            m = d.trigger_mask[trig_row][:, None, :] & sel[:, :, None]        # [K, B, F]
            v = d.trigger_val[trig_row][:, None, :].expand_as(x)
            x.copy_(torch.where(m, v, x))
        y.masked_fill_(sel, self.label_swap)

    # ------------------------------------------------------------------------------------------ training
    def train(self, sch: RoundSchedule, W: torch.Tensor, cnt: torch.Tensor, anchor: torch.Tensor,
              want_grad_sum: bool, generator: Optional[torch.Generator] = None) -> TrainResult:
        """W [K, Df] local states (updated in place), cnt [K, nc], anchor [K, Df] = states at segment start."""
        p, spec, dev = self.params, self.spec, self.device
        K, Dp = W.shape[0], spec.Dp
        B = int(p["batch_size"])
        mu, wd, alpha = float(p["momentum"]), float(p["decay"]), float(p["alpha_loss"])
        E = max(sch.E_max, 1)
        res = TrainResult(
            torch.zeros(K, Dp, device=dev) if want_grad_sum else None,
            torch.zeros(K, E, device=dev), torch.zeros(K, E, dtype=torch.long, device=dev),
            torch.zeros(K, E, dtype=torch.long, device=dev), torch.zeros(K, E, dtype=torch.long, device=dev),
            torch.zeros(K, dtype=torch.bool, device=dev))
        if sch.T == 0:
            return res
        if p["vis_train_batch_loss"]:
            res.step_loss = torch.zeros(K, sch.T, device=dev)
        if p["batch_track_distance"]:
            res.step_dist = torch.zeros(K, sch.T, device=dev)
        V = torch.zeros(K, Dp, device=dev)
        P = W[:, :Dp]                      # view: updated in place
        S = W[:, Dp:]
        poisoning = sch.poison_n > 0
        use_dist = alpha != 1.0 and bool(poisoning.any())
        ar = torch.arange(B, device=dev)
        kk = torch.arange(K, device=dev)
        for t in range(sch.T):
            ids, act = sch.idx[:, t], sch.active[:, t]
            mask = ids >= 0
            x, y = self._gather(ids, train=True)
            if sch.flip is not None:
                f = sch.flip[:, t]
                x = torch.where(f[:, :, None, None, None], x.flip(-1), x)
            sel = (ar[None, :] < sch.poison_n[:, None]) & mask
            if bool(sel.any()):
                self._poison_(x, y, sel, sch.trig_row)
            Pg = P.detach().clone().requires_grad_(True)
            logits = self.net.forward(Pg, S, cnt, x, training=True, shared_input=False, mask=mask, active=act,
                                      generator=generator)
            ce = F.cross_entropy(logits.reshape(K * B, -1), y.reshape(-1), reduction="none").reshape(K, B)
            n_k = mask.sum(1)
            loss_k = (ce * mask).sum(1) / n_k.clamp(min=1)
            if use_dist:
                dist = (Pg - anchor[:, :Dp]).norm(dim=1)
                loss_k = torch.where(poisoning, alpha * loss_k + (1 - alpha) * dist, loss_k)
            torch.where(act, loss_k, torch.zeros_like(loss_k)).sum().backward()
            g = Pg.grad
            with torch.no_grad():
                self._sgd_update(W, V, g, sch.lr[:, t], act, res.grad_sum, mu, wd)
                e = sch.epoch_of[:, t]
                pred = logits.argmax(-1)
                res.loss_sum[kk, e] += torch.where(act, loss_k.detach(), torch.zeros_like(loss_k))
                res.correct[kk, e] += (((pred == y) & mask).sum(1)) * act
                res.count[kk, e] += n_k * act
                res.poison_count[kk, e] += sel.sum(1) * act
                res.nan_flag |= ~torch.isfinite(logits).reshape(K, -1).all(1) & act
                if res.step_loss is not None:
                    res.step_loss[:, t] = torch.where(act, loss_k.detach(), torch.zeros_like(loss_k))
                if res.step_dist is not None:
                    res.step_dist[:, t] = (P - anchor[:, :Dp]).norm(dim=1) * act
        self.W_last = W
        return res

    def _sgd_update(self, W, V, g, lr_t, act, grad_sum, mu, wd) -> None:
        """v = mu*v + (g + wd*w); w -= lr*v on the rows with act[k]; grad_sum += g (raw gradient, Q10)."""
        Dp = self.spec.Dp
        P = W[:, :Dp]
        a = act[:, None].to(g.dtype)
        g = torch.where(act[:, None], g, torch.zeros_like(g))
        if grad_sum is not None:
            grad_sum += g
        g = g + wd * P
        V.mul_(1 - a + a * mu).add_(g * a)
        P.sub_(V * (lr_t[:, None] * a))

    # ---------------------------------------------------------------------------------------- evaluation
    @torch.no_grad()
    def evaluate_jobs(self, model: torch.Tensor, jobs, shard=None):
        """One model on several (split, trigger) input sets; returns [(loss_sum, correct, n)] for this rank's shard."""
        out = []
        for split, trig in jobs:
            l, c, n = self.evaluate(model.reshape(1, -1), split, trig, shard=shard)
            out.append((float(l[0]), int(c[0]), n))
        return out

    @torch.no_grad()
    def evaluate(self, models: torch.Tensor, split: str, trig_row: int, shard=None
                 ) -> Tuple[torch.Tensor, torch.Tensor, int]:
        """models [M, Df].  split 'clean' (whole test set, true labels) or 'poison' (poison test subset, every
        sample stamped with trigger ``trig_row`` and relabelled; test.py:54-239).  Returns per-model
        (sum of CE, number correct) and the sample count."""
        d, spec, dev = self.data, self.spec, self.device
        M = models.shape[0]
        ids_all = d.all_test_idx if split == "clean" else d.poison_test_idx
        if shard is not None:
            r, w = shard
            per = (int(ids_all.numel()) + w - 1) // w
            ids_all = ids_all[r * per:(r + 1) * per]
        loss = torch.zeros(M, dtype=torch.float64, device=dev)
        correct = torch.zeros(M, dtype=torch.long, device=dev)
        P, S = models[:, :spec.Dp], models[:, spec.Dp:]
        n = int(ids_all.numel())
        for s in range(0, n, self.eval_chunk):
            ids = ids_all[s:s + self.eval_chunk]
            x, y = self._gather(ids, train=False)
            if split == "poison":
                row = torch.tensor([trig_row], device=dev)
                xb, yb = x.unsqueeze(0), y.unsqueeze(0)
                self._poison_(xb, yb, torch.ones_like(yb, dtype=torch.bool), row)
                x, y = xb[0], yb[0]
            logits = self.net.forward(P, S, None, x, training=False, shared_input=True)   # [M, b, cls]
            b = y.numel()
            ce = F.cross_entropy(logits.reshape(M * b, -1), y.repeat(M), reduction="none").reshape(M, b)
            loss += ce.sum(1).double()
            correct += (logits.argmax(-1) == y[None, :]).sum(1)
        return loss, correct, n

    # ------------------------------------------------------------------------------------------ flat ops
    @staticmethod
    def scale_(W: torch.Tensor, anchor: torch.Tensor, scale: torch.Tensor) -> None:
        """W <- anchor + (W - anchor) * scale[k]  over the whole state (image_train.py:166-171)."""
        W.copy_(anchor + (W - anchor) * scale[:, None])

    def dist_norm(self, W: torch.Tensor, anchor: torch.Tensor) -> torch.Tensor:
        """helper.py:67-71: L2 distance over *parameters* only.  Returns [K]."""
        Dp = self.spec.Dp
        return (W[:, :Dp] - anchor[:, :Dp]).norm(dim=1)

    def global_norm(self, W: torch.Tensor) -> torch.Tensor:
        """helper.py:60-64: L2 norm over parameters.  [K]"""
        return W[:, :self.spec.Dp].norm(dim=1)

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
