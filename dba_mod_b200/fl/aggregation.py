"""Server-side aggregation on flat slabs: FedAvg (+DP noise), RFA / Weiszfeld geometric median, FoolsGold.

Reference semantics (helper.py):
  * FedAvg ``average_shrink_models`` 240-257:  G += (eta / no_models) * sum_c delta_c  (+ N(0, sigma) per element when
    ``diff_privacy``), over *all* state_dict entries incl. BN running stats (Q9, Q23: unweighted).
  * RFA ``geometric_median_update`` 295-373:  alphas = n_c / sum n;  m = sum alpha_c delta_c;  up to ``maxiter`` times:
    w_c = alpha_c / max(eps, ||m - delta_c||), normalise, m = sum w_c delta_c, objective sum alpha_c ||m - delta_c||,
    stop when |prev - obj| < ftol * obj;  finally  G += eta * m  (eta, not eta/n: Q23).  Reported: ``wv`` = weights
    of the last non-converged iteration (Q8) and ``alphas`` = final distances.
  * FoolsGold ``foolsgold_update`` 259-293 + class FoolsGold 527-607: per-client history of the last-layer-weight
    gradient (Q12), cosine similarity, pardoning, logit -> wv;  g = sum_c wv_c * gradsum_c / n;  one step of a fresh
    SGD(lr, momentum, decay) on the global parameters with grad = eta * g, i.e.  w -= lr * (eta*g + decay*w)  (Q11);
    BN buffers untouched (Q9).

Every function takes the *local* shard of client rows plus a ``comm`` object; with ``world == 1`` the collectives are
identities.  The torch ops here are the oracle / CPU path; on B200 the same steps run as the fused kernels of
csrc/ (local K-way reduce + peer-memory reduce + update + broadcast in one launch) through ``CudaBackend``.

Decision on Q3 (int64 ``num_batches_tracked``): modern torch makes the reference crash when it adds a float update
into these counters; the torch<=1.4 behaviour was integer truncation.  FedAvg reproduces the truncating add; RFA and
FoolsGold leave the counters alone and exclude them from distances (they carry no model information).
"""
from __future__ import annotations

import logging
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

logger = logging.getLogger("logger")


# ------------------------------------------------------------------------------------------------ comm
class LocalComm:
    """world == 1."""
    rank, world = 0, 1

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def all_gather_rows(self, t: torch.Tensor, counts: List[int]) -> torch.Tensor:
        return t

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        return t

    def barrier(self) -> None:
        pass


class TorchDistComm:
    """``torch.distributed`` collectives (gloo on CPU, NCCL on GPUs).  On B200 this is the measured *baseline*
    path for the aggregation collectives; the product path is the fused peer-memory kernels."""

    def __init__(self, group=None) -> None:
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def all_reduce_sum_(self, t):
        self.dist.all_reduce(t, group=self.group)
        return t

    def all_gather_rows(self, t, counts):
        """Ragged gather along dim 0: rank r contributes counts[r] rows."""
        mx = max(counts)
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        out = [torch.empty_like(pad) for _ in range(self.world)]
        self.dist.all_gather(out, pad, group=self.group)
        return torch.cat([o[:c] for o, c in zip(out, counts)], 0)

    def broadcast_(self, t, src=0):
        self.dist.broadcast(t, src=src, group=self.group)
        return t

    def barrier(self):
        self.dist.barrier(group=self.group)


# ---------------------------------------------------------------------------------------------- FedAvg
def dp_noise_like(t: torch.Tensor, sigma: float, generator=None) -> torch.Tensor:
    """helper.py:187-191 (Q22: the reference needs CUDA for this; here any device works)."""
    return torch.randn(t.shape, device=t.device, dtype=t.dtype, generator=generator) * sigma


def fedavg_update_(G: torch.Tensor, G_cnt: torch.Tensor, delta: torch.Tensor, delta_cnt: Optional[torch.Tensor],
                   eta: float, no_models: int, comm=LocalComm(), sigma: Optional[float] = None,
                   generator=None) -> None:
    """G [Df] += eta/no_models * sum over all ranks' rows of delta [K_local, Df]."""
    coef = eta / no_models
    acc = delta.sum(0) if delta.shape[0] else torch.zeros_like(G)
    comm.all_reduce_sum_(acc)
    upd = acc * coef
    if sigma is not None:
        noise = dp_noise_like(G, sigma, generator)
        comm.broadcast_(noise, 0)     # every replica of G must see the same noise
        upd = upd + noise
    G.add_(upd)
    if delta_cnt is not None and G_cnt.numel() and delta_cnt.numel():
        c = delta_cnt.sum(0).double() if delta_cnt.shape[0] else torch.zeros_like(G_cnt, dtype=torch.double)
        comm.all_reduce_sum_(c)
        G_cnt.add_((c * coef).trunc().long())   # Q3: torch<=1.4 truncating add


# ------------------------------------------------------------------------------------------------- RFA
@dataclass
class RFAResult:
    num_oracle_calls: int
    wv: List[float]
    distances: List[float]
    median: torch.Tensor
    objective: float
    iters: int


def _rows_dist(m: torch.Tensor, pts: torch.Tensor) -> torch.Tensor:
    return (pts - m[None, :]).double().pow(2).sum(1).sqrt() if pts.shape[0] else torch.zeros(0, dtype=torch.double,
                                                                                              device=m.device)


def geometric_median(points: torch.Tensor, num_samples: torch.Tensor, all_num_samples: List[float],
                     counts: List[int], maxiter: int = 4, eps: float = 1e-5, ftol: float = 1e-6,
                     comm=LocalComm()) -> RFAResult:
    """Weiszfeld over the union of all ranks' rows.  points [K_local, Df]; num_samples [K_local] (this rank's
    rows); all_num_samples = sample counts of ALL clients in global slot order; counts[r] = rows on rank r."""
    dev = points.device
    tot = float(sum(all_num_samples))
    alphas = (num_samples.double() / tot).float()            # reference: float64 -> float32

    def _wavg(w_local: torch.Tensor, w_total: torch.Tensor) -> torch.Tensor:
        part = ((w_local / w_total).float()[:, None] * points).sum(0) if points.shape[0] else \
            torch.zeros(points.shape[1], device=dev)
        return comm.all_reduce_sum_(part)

    def _objective(m) -> Tuple[float, torch.Tensor]:
        d = _rows_dist(m, points)
        o = (alphas.double() * d).sum().reshape(1)
        comm.all_reduce_sum_(o)
        return float(o.item()), d

    a_tot = comm.all_reduce_sum_(alphas.sum().reshape(1).clone())
    median = _wavg(alphas, a_tot)
    calls = 1
    obj, dist = _objective(median)
    wv_local: Optional[torch.Tensor] = None
    last_w = alphas / a_tot
    it = 0
    for it in range(maxiter):
        prev_obj = obj
        w = alphas / torch.clamp(dist.float(), min=eps)
        w_tot = comm.all_reduce_sum_(w.sum().reshape(1).clone())
        w = w / w_tot
        median = _wavg(w, torch.ones(1, device=dev))
        calls += 1
        obj, dist = _objective(median)
        last_w = w
        if abs(prev_obj - obj) < ftol * obj:
            break
        wv_local = w.clone()
    if wv_local is None:          # Q8: the reference crashes here (wv is None); report the last weights instead
        wv_local = last_w
    wv_all = comm.all_gather_rows(wv_local.reshape(-1, 1), counts).reshape(-1)
    d_all = comm.all_gather_rows(dist.float().reshape(-1, 1), counts).reshape(-1)
    return RFAResult(calls, [float(v) for v in wv_all.tolist()], [float(v) for v in d_all.tolist()], median, obj,
                     it + 1)


def rfa_update_(G: torch.Tensor, res: RFAResult, eta: float, sigma: Optional[float] = None, comm=LocalComm(),
                generator=None, max_update_norm: Optional[float] = None) -> bool:
    update_norm = float(res.median.double().pow(2).sum().sqrt())
    if max_update_norm is not None and update_norm >= max_update_norm:
        logger.info("\t\t\tUpdate norm = {} is too large. Update rejected".format(update_norm))
        return False
    upd = res.median * eta
    if sigma is not None:
        noise = dp_noise_like(G, sigma, generator)
        comm.broadcast_(noise, 0)
        upd = upd + noise
    G.add_(upd)
    return True


# ------------------------------------------------------------------------------------------- FoolsGold
def foolsgold_weights(feats: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """feats [n, L] (any float dtype) -> (wv [n], alpha [n]) in float64; helper.py:574-607."""
    x = feats.double()
    n = x.shape[0]
    norms = x.norm(dim=1, keepdim=True)
    norms = torch.where(norms == 0, torch.ones_like(norms), norms)      # sklearn.normalize leaves zero rows
    xn = x / norms
    cs = xn @ xn.t() - torch.eye(n, dtype=torch.double, device=x.device)
    maxcs = cs.max(dim=1).values
    ratio = maxcs[:, None] / maxcs[None, :]
    pardon = maxcs[:, None] < maxcs[None, :]
    cs = torch.where(pardon, cs * ratio, cs)
    wv = 1 - cs.max(dim=1).values
    wv = wv.clamp(0, 1)
    alpha = cs.max(dim=1).values
    mx = wv.max()
    wv = wv / mx if float(mx) != 0.0 else wv
    wv = torch.where(wv == 1, torch.full_like(wv, 0.99), wv)
    wv = torch.log(wv / (1 - wv)) + 0.5
    wv = torch.where(torch.isnan(wv), torch.zeros_like(wv), wv)
    wv = torch.where(wv > 1, torch.ones_like(wv), wv)
    wv = torch.where(wv < 0, torch.zeros_like(wv), wv)
    return wv, alpha


@dataclass
class FoolsGoldState:
    """Cross-round history ``memory_dict[name] += grad_lastlayer`` (helper.py:545-549); float64 like numpy.
    Not part of the reference's checkpoint (§3.5), kept that way."""
    use_memory: bool = False
    memory: Dict[object, torch.Tensor] = field(default_factory=dict)
    wv_history: List[List[float]] = field(default_factory=list)

    def features(self, names: List, feats: torch.Tensor) -> torch.Tensor:
        """feats [n, L] = this round's last-layer gradient sums in ``names`` order.  Updates the history and
        returns the matrix FoolsGold should look at."""
        rows = []
        f64 = feats.double()
        for i, nm in enumerate(names):
            if nm in self.memory:
                self.memory[nm] = self.memory[nm] + f64[i]
            else:
                self.memory[nm] = f64[i].clone()
            rows.append(self.memory[nm])
        return torch.stack(rows) if self.use_memory else f64


def foolsgold_update_(G: torch.Tensor, Dp: int, grad_sum: torch.Tensor, wv_local: torch.Tensor, n_clients: int,
                      eta: float, lr: float, decay: float, comm=LocalComm()) -> None:
    """G[:Dp] -= lr * (eta * sum_c wv_c * gradsum_c / n + decay * G[:Dp])."""
    part = (wv_local.float()[:, None] * grad_sum).sum(0) if grad_sum.shape[0] else \
        torch.zeros(Dp, device=G.device)
    comm.all_reduce_sum_(part)
    g = part * (eta / n_clients)
    P = G[:Dp]
    P.sub_(lr * (g + decay * P))
