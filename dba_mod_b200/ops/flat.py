"""Tensor-level wrappers of the flat-slab kernels (csrc/flat_ops.cu)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import native

_scratch = {}


def _scratch_for(dev: torch.device, elems: int) -> torch.Tensor:
    t = _scratch.get(dev)
    if t is None or t.numel() < elems:
        t = torch.empty(max(elems, 1 << 16), dtype=torch.float64, device=dev)
        _scratch[dev] = t
    return t


def sgd_step_(W: torch.Tensor, V: torch.Tensor, Gr: torch.Tensor, lr: torch.Tensor, active: Optional[torch.Tensor],
              Dp: int, mu: float, wd: float, grad_sum: Optional[torch.Tensor] = None) -> None:
    """W [K, ld] (parameters in the first Dp columns), V / Gr / grad_sum [K, Dp], lr [K] fp32, active [K] uint8."""
    lib = native.load()
    K = W.shape[0]
    assert V.stride(0) == Gr.stride(0) and (grad_sum is None or grad_sum.stride(0) == V.stride(0))
    native.check(lib.dba_sgd_step(native.ptr(W), W.stride(0), native.ptr(V), native.ptr(Gr), native.ptr(grad_sum),
                                  V.stride(0), native.ptr(lr), native.ptr(active), K, Dp, mu, wd, native.stream()))


def scale_rows_(W: torch.Tensor, A: torch.Tensor, s: torch.Tensor) -> None:
    lib = native.load()
    native.check(lib.dba_scale_rows(native.ptr(W), W.stride(0), native.ptr(A), A.stride(0), native.ptr(s),
                                    W.shape[0], W.shape[1], native.stream()))


def rows_dist(X: torch.Tensor, Y: Optional[torch.Tensor] = None, Z: Optional[torch.Tensor] = None, n: int = None,
              sqrt: bool = True) -> torch.Tensor:
    """out[k] = || X[k,:n] - Y[k or 0,:n] - Z[:n] ||  (fp64 result).  Y with one row is broadcast."""
    lib = native.load()
    K = X.shape[0]
    n = n or X.shape[1]
    out = torch.empty(K, dtype=torch.float64, device=X.device)
    need = int(lib.dba_rows_sqdist_scratch_elems(K))
    sc = _scratch_for(X.device, need)
    ldy = 0 if (Y is None or Y.shape[0] == 1) else Y.stride(0)
    native.check(lib.dba_rows_sqdist(native.ptr(X), X.stride(0), native.ptr(Y), ldy, native.ptr(Z), K, n,
                                     native.ptr(out), int(sqrt), native.ptr(sc), sc.numel(), native.stream()))
    return out


def fedavg_update_(G: torch.Tensor, W: torch.Tensor, coef: float, sigma: float = 0.0, seed: int = 0,
                   bcast: bool = False) -> None:
    lib = native.load()
    native.check(lib.dba_fedavg_update(native.ptr(G), native.ptr(W), W.stride(0) if W.numel() else G.numel(),
                                       W.shape[0], G.numel(), coef, sigma, seed, int(bcast), native.stream()))


def weiszfeld(W: torch.Tensor, G: torch.Tensor, alphas: torch.Tensor, maxiter: int, eps: float = 1e-5,
              ftol: float = 1e-6):
    """Device-side RFA loop over the deltas (W[c] - G).  Returns (median [D], dist [n] f64, wv [n], obj [1], state)."""
    lib = native.load()
    n, D = W.shape[0], G.numel()
    dev = W.device
    median = torch.empty(D, dtype=torch.float32, device=dev)
    dist = torch.zeros(n, dtype=torch.float64, device=dev)
    wts = torch.zeros(n, dtype=torch.float32, device=dev)
    wv = torch.zeros(n, dtype=torch.float32, device=dev)
    obj = torch.zeros(1, dtype=torch.float64, device=dev)
    state = torch.zeros(4, dtype=torch.int32, device=dev)
    sc = _scratch_for(dev, int(lib.dba_rows_sqdist_scratch_elems(n)))
    native.check(lib.dba_weiszfeld(native.ptr(W), W.stride(0), native.ptr(G), n, D, native.ptr(alphas), maxiter, eps,
                                   ftol, native.ptr(median), native.ptr(dist), native.ptr(wts), native.ptr(wv),
                                   native.ptr(obj), native.ptr(state), native.ptr(sc), sc.numel(), native.stream()))
    return median, dist, wv, obj, state


def foolsgold_weights(history: torch.Tensor, ids: torch.Tensor, feat: torch.Tensor, use_memory: bool
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    """history [N_total, L] f64 (updated in place), ids [n] int32 rows of this round's clients, feat [n, L] fp32."""
    lib = native.load()
    n, L = feat.shape
    dev = feat.device
    work = torch.empty(n, L, dtype=torch.float64, device=dev)
    gram = torch.empty(n, n, dtype=torch.float64, device=dev)
    cs = torch.empty(n, n, dtype=torch.float64, device=dev)
    wv = torch.empty(n, dtype=torch.float64, device=dev)
    alpha = torch.empty(n, dtype=torch.float64, device=dev)
    native.check(lib.dba_foolsgold_weights(native.ptr(history), history.stride(0), native.ptr(ids), native.ptr(feat),
                                           feat.stride(0), n, L, int(use_memory), native.ptr(work), native.ptr(gram),
                                           native.ptr(cs), native.ptr(wv), native.ptr(alpha), native.stream()))
    return wv, alpha


def foolsgold_apply_(G: torch.Tensor, S: torch.Tensor, wv: torch.Tensor, Dp: int, eta_over_n: float, lr: float,
                     decay: float) -> None:
    lib = native.load()
    native.check(lib.dba_foolsgold_apply(native.ptr(G), native.ptr(S), S.stride(0) if S.numel() else Dp,
                                         native.ptr(wv), S.shape[0], Dp, eta_over_n, lr, decay, native.stream()))
