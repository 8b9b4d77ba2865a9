"""Round schedule: everything a round's local training needs, decided on the host ONCE and laid out as device
tensors, so the K clients of a rank then run ``T`` lock-step optimiser steps with no host logic in between.

Replaces the reference's nested Python loops (image_train.py:21-271, loan_train.py:17-196): which client poisons
(``is_poison`` and adversary and global epoch in its ``{i}_poison_epochs``), its epoch count
(``internal_poison_epochs`` vs ``internal_epochs``), its LR per internal epoch (``poison_lr`` + the
``MultiStepLR([0.2N, 0.8N], 0.1)`` semantics, Q4/Q18; LOAN's ASR-adaptive poison LR, loan_train.py:67-75), which
trigger it stamps (global pattern when ``len(adversary_list) == 1``) and the per-epoch sample order
(``SubsetRandomSampler`` / ``shuffle=True`` -> a fresh device-side permutation per client and epoch).

Step t of client k is batch ``t % nb_k`` of internal epoch ``t // nb_k``; clients shorter than ``T`` are masked.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import torch

from .. import config as C


@dataclass
class ClientPlan:
    name: object
    slot: int
    adv_index: int            # index into adversary_list, -1 = benign participant
    poisoning: bool           # trains with poisoned batches in this global epoch
    trig_row: int             # row of the trigger table (0 = global pattern, j+1 = local j); -1 = none
    n_samples: int
    nb: int
    n_epochs: int
    lr_per_epoch: List[float]
    poison_per_batch: int
    scale: float              # scale_weights_poison applied after training (1.0 = none)


@dataclass
class RoundSchedule:
    epoch: int
    plans: List[ClientPlan]
    idx: torch.Tensor          # [K, T, B] int64 dataset index, -1 = padding
    lr: torch.Tensor           # [K, T] float32
    active: torch.Tensor       # [K, T] bool
    epoch_of: torch.Tensor     # [K, T] int64 internal-epoch id (0-based)
    poison_n: torch.Tensor     # [K] int64: leading samples of every batch to poison
    trig_row: torch.Tensor     # [K] int64 (clamped to >= 0)
    flip: Optional[torch.Tensor]  # [K, T, B] bool random horizontal flips (Tiny-ImageNet train transform)
    T: int = 0
    E_max: int = 0
    steps_host: List[int] = field(default_factory=list)


def plan_clients(params, data, names: List, epoch: int, poison_lr_override: Optional[dict] = None
                 ) -> List[ClientPlan]:
    plans = []
    for slot, name in enumerate(names):
        adv_index = params.adversary_index(name) if (params["is_poison"] and name in params["adversary_list"]) else -1
        poisoning = adv_index >= 0 and epoch in params.poison_epochs_of(adv_index)
        trig = -1
        if adv_index >= 0:
            trig = 0 if params.is_centralized_attack else adv_index + 1
        n = data.client_size(name)
        nb = data.batches_per_epoch(name)
        if poisoning:
            n_ep = int(params["internal_poison_epochs"])
            base = float(params["poison_lr"])
            if poison_lr_override and name in poison_lr_override:
                base = poison_lr_override[name]
            if params["poison_step_lr"]:
                fac = C.multistep_lr_factors(n_ep, 0.1, params["multistep_lr_compat"],
                                             step_first=(params["type"] == C.TYPE_LOAN))
            else:
                fac = [1.0] * n_ep
            lrs = [base * f for f in fac]
            scale = 1.0 if params["baseline"] else float(params["scale_weights_poison"])
            ppb = int(params["poisoning_per_batch"])
        else:
            n_ep = int(params["internal_epochs"])
            lrs = [float(params["lr"])] * n_ep
            scale, ppb = 1.0, 0
        plans.append(ClientPlan(name, slot, adv_index, poisoning, trig, n, nb, n_ep, lrs, ppb, scale))
    return plans


def _stable_id(name) -> int:
    import zlib
    return int(name) if isinstance(name, int) else zlib.crc32(str(name).encode())


def build_schedule(params, data, plans: List[ClientPlan], epoch: int, device: torch.device,
                   generator: Optional[torch.Generator] = None) -> RoundSchedule:
    """Sample orders are drawn from a stream keyed by (seed, global epoch, client name), so a client sees the same
    batches no matter which rank / slot it lands on (world-size invariance of the whole simulation)."""
    K = len(plans)
    if generator is None:
        generator = torch.Generator(device=device)
    base_seed = (int(params["seed"]) * 1000003 + int(epoch)) * 1000003
    B = int(params["batch_size"])
    steps = [p.n_epochs * p.nb for p in plans]
    T = max(steps) if steps else 0
    E_max = max((p.n_epochs for p in plans), default=0)
    idx = torch.full((K, max(T, 1), B), -1, dtype=torch.long, device=device)
    lr = torch.zeros(K, max(T, 1), dtype=torch.float32, device=device)
    active = torch.zeros(K, max(T, 1), dtype=torch.bool, device=device)
    epoch_of = torch.zeros(K, max(T, 1), dtype=torch.long, device=device)
    for k, p in enumerate(plans):
        if p.n_samples == 0 or p.n_epochs == 0:
            continue
        table = data.client_index_tensor(p.name)
        generator.manual_seed((base_seed + _stable_id(p.name)) % (2 ** 63 - 1))
        keys = torch.rand(p.n_epochs, p.n_samples, device=device, generator=generator)
        order = table[keys.argsort(dim=1)]                                    # [E, n] fresh order per epoch
        pad = p.nb * B - p.n_samples
        if pad:
            order = torch.cat([order, torch.full((p.n_epochs, pad), -1, dtype=torch.long, device=device)], 1)
        tk = p.n_epochs * p.nb
        idx[k, :tk] = order.reshape(tk, B)
        active[k, :tk] = True
        e_ids = torch.arange(p.n_epochs, device=device).repeat_interleave(p.nb)
        epoch_of[k, :tk] = e_ids
        lr[k, :tk] = torch.tensor(p.lr_per_epoch, dtype=torch.float32, device=device)[e_ids]
    poison_n = torch.tensor([p.poison_per_batch if p.poisoning else 0 for p in plans], dtype=torch.long,
                            device=device)
    trig_row = torch.tensor([max(p.trig_row, 0) for p in plans], dtype=torch.long, device=device)
    flip = None
    if getattr(data, "train_flip", False):
        flip = torch.zeros(K, max(T, 1), B, dtype=torch.bool, device=device)
        for k, p in enumerate(plans):
            generator.manual_seed((base_seed + 7 + _stable_id(p.name)) % (2 ** 63 - 1))
            flip[k] = torch.rand(max(T, 1), B, device=device, generator=generator) < 0.5
    return RoundSchedule(epoch, plans, idx, lr, active, epoch_of, poison_n, trig_row, flip, T, E_max, steps)
