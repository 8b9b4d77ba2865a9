"""Auxiliary Helper functionality of the reference that is not on the ``main.py`` call path, restated
on flat parameter vectors so a user of the reference still finds it:

  model_max_values(_var)      helper.py:74-85    per-tensor max |w - w_ref|
  get_one_vec                 helper.py:88-108   flatten the parameters (already the native representation here)
  cos_sim_loss                helper.py:125-137  1e3 * (1 - cos(scale*(w - w_ref) + w_ref, w_ref))
  model_cosine_similarity     helper.py:139-163  per-tensor cosine of the 100x-scaled update, averaged
  accum_similarity            helper.py:165-184  per-tensor cosine between two accumulators
  update/save_epoch_submit_dict helper.py:437-454  JSON dump of the submitted updates per global epoch
  estimate_fisher / consolidate / ewc_loss  helper.py:456-525  elastic weight consolidation

All functions take flat vectors ``[Dp]`` (or ``[K, Dp]``) laid out by :class:`ParamSpec`.
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from ..models import BatchedNet, ParamSpec


def _tensors(spec: ParamSpec, flat: torch.Tensor):
    for name in spec.param_names:
        e = spec.entries[name]
        yield name, flat[..., e.offset:e.offset + e.numel]


def model_max_values(spec: ParamSpec, flat: torch.Tensor, ref: torch.Tensor) -> List[torch.Tensor]:
    return [(a - b).abs().amax(dim=-1) for (_, a), (_, b) in zip(_tensors(spec, flat), _tensors(spec, ref))]


def model_max_values_var(spec: ParamSpec, flat: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    return sum(model_max_values(spec, flat, ref))


def get_one_vec(spec: ParamSpec, flat: torch.Tensor) -> torch.Tensor:
    return flat[..., :spec.Dp]


def cos_sim_loss(spec: ParamSpec, flat: torch.Tensor, target_vec: torch.Tensor, scale_weights: float) -> torch.Tensor:
    w, t = flat[..., :spec.Dp], target_vec[..., :spec.Dp]
    cs = F.cosine_similarity(scale_weights * (w - t) + t, t, dim=-1)
    return 1e3 * (1 - cs)


def model_cosine_similarity(spec: ParamSpec, flat: torch.Tensor, ref: torch.Tensor, scale: float = 100.0
                            ) -> torch.Tensor:
    cs = []
    for (_, a), (_, b) in zip(_tensors(spec, flat), _tensors(spec, ref)):
        cs.append(F.cosine_similarity(scale * (a - b) + b, b, dim=-1))
    mean_cs = sum(cs) / len(cs)
    return 1e3 * (1 - mean_cs)


def accum_similarity(spec: ParamSpec, last_acc: torch.Tensor, new_acc: torch.Tensor) -> torch.Tensor:
    cs = [F.cosine_similarity(a, b, dim=-1) for (_, a), (_, b) in zip(_tensors(spec, last_acc),
                                                                     _tensors(spec, new_acc))]
    return 1 - sum(cs) / len(cs)


# ---------------------------------------------------------------------------------------- submitted updates (JSON)
def update_epoch_submit_dict(spec: ParamSpec, deltas: torch.Tensor, names: List, epoch: int,
                             store: Optional[Dict] = None) -> Dict:
    """store[epoch][client] = {tensor name: nested list} (what the reference keeps after ``.tolist()``, Q21)."""
    store = {} if store is None else store
    per = {}
    for k, nm in enumerate(names):
        sd = spec.unflatten(deltas[k].detach().cpu(), None)
        per[str(nm)] = {n: v.tolist() for n, v in sd.items()}
    store[int(epoch)] = per
    return store


def save_epoch_submit_dict(folder_path: str, store: Dict) -> str:
    path = os.path.join(folder_path, "epoch_submit_update.json")
    with open(path, "w") as f:
        json.dump(store, f, ensure_ascii=False, indent=1)
    return path


# --------------------------------------------------------------------------------- elastic weight consolidation
class EWC:
    """Fisher information of the per-sample log-likelihoods and the quadratic EWC penalty (helper.py:456-525)."""

    def __init__(self, kind: str, spec: ParamSpec) -> None:
        self.spec = spec
        self.net = BatchedNet(kind, spec)
        self.mean: Optional[torch.Tensor] = None
        self.fisher: Optional[torch.Tensor] = None

    def estimate_fisher(self, flat: torch.Tensor, batches) -> torch.Tensor:
        """batches: iterable of (x [B, ...] float, y [B] long).  Returns the squared gradient of the mean
        log-likelihood, as a flat ``[Dp]`` vector."""
        spec = self.spec
        P = flat[:spec.Dp].detach().clone().requires_grad_(True)
        S = flat[spec.Dp:].detach().clone()[None] if spec.Db else flat.new_zeros(1, 0)
        lls = []
        for x, y in batches:
            logits = self.net.forward(P[None], S, None, x, training=False, shared_input=True)[0]
            lls.append(F.log_softmax(logits, dim=1)[torch.arange(y.numel()), y])
        ll = torch.cat(lls).mean(0)
        (g,) = torch.autograd.grad(ll, P)
        return g ** 2

    def consolidate(self, flat: torch.Tensor, fisher: torch.Tensor) -> None:
        self.mean = flat[:self.spec.Dp].detach().clone()
        self.fisher = fisher.detach().clone()

    def ewc_loss(self, flat: torch.Tensor, lamda: float) -> torch.Tensor:
        if self.mean is None:
            return flat.new_zeros(1)
        return (lamda / 2) * (self.fisher * (flat[..., :self.spec.Dp] - self.mean) ** 2).sum(-1)
