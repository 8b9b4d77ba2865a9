"""ParamSpec: a model as ONE flat fp32 vector (SURVEY §7.1 "flat everything").

The reference handles a model as a ``state_dict`` of 6..122 tensors and loops over it in Python for every copy,
delta, norm and aggregation step (helper.py:60-71,193-257,295-418; image_train.py:26-27,166-171,302-306).  Here the
model state is ``flat[Df] = [ all parameters (Dp) | all float buffers (Db: BN running mean / var) ]`` plus a tiny
int64 side vector for the ``num_batches_tracked`` counters, so each of those steps is a single kernel over
``[K, Df]`` slabs.  The ``state_dict`` form (reference key names, order, shapes and dtypes) exists only at the
checkpoint / oracle boundary: :meth:`flatten` and :meth:`unflatten`.

Internal storage order of 4-D convolution weights is ``[Cout, kh, kw, Cin]`` (channels-last), not PyTorch's
``[Cout, Cin, kh, kw]``: that is the order in which the tensor-core weight-gradient kernel produces its tiles
(contiguous runs of Cin per (Cout, tap)) and from which the packed bf16 operand layouts are cut.  All flat-slab
maths (SGD, deltas, aggregation, norms) is element-wise and therefore indifferent to the order; ``state_dict``s
going in or out are permuted here, and :meth:`view` hands out ``[Cout, Cin, kh, kw]``-shaped (strided) views.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch


@dataclass(frozen=True)
class Entry:
    name: str
    shape: Tuple[int, ...]
    kind: str          # 'param' | 'buffer' | 'counter'
    offset: int        # into flat (param / buffer) or into the counter vector
    numel: int
    ohwi: bool = False  # 4-D conv weights are STORED [Cout, kh, kw, Cin] in the flat vector (see ParamSpec)


class ParamSpec:
    def __init__(self, module: torch.nn.Module) -> None:
        sd = module.state_dict()
        pnames = [n for n, _ in module.named_parameters()]
        pset = set(pnames)
        self.order: List[str] = list(sd.keys())
        self.entries: Dict[str, Entry] = {}
        off = 0
        for n in pnames:
            t = sd[n]
            self.entries[n] = Entry(n, tuple(t.shape), "param", off, t.numel(), ohwi=(t.dim() == 4))
            off += t.numel()
        self.Dp = off
        nc = 0
        for n in self.order:
            if n in pset:
                continue
            t = sd[n]
            if t.dtype.is_floating_point:
                self.entries[n] = Entry(n, tuple(t.shape), "buffer", off, t.numel())
                off += t.numel()
            else:
                self.entries[n] = Entry(n, tuple(t.shape), "counter", nc, t.numel())
                nc += t.numel()
        self.Df = off
        self.Db = self.Df - self.Dp
        self.n_counters = nc
        self.D = self.Df + nc
        self.param_names = pnames
        self.buffer_names = [n for n in self.order if self.entries[n].kind == "buffer"]
        self.counter_names = [n for n in self.order if self.entries[n].kind == "counter"]
        # FoolsGold feature = named_parameters()[-2] (helper.py:537,544; Q12)
        e = self.entries[pnames[-2]]
        self.fg_slice = slice(e.offset, e.offset + e.numel)
        self.fg_len = e.numel

    # --------------------------------------------------------------------------------------- conversions
    def flatten(self, state_dict, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
        flat = torch.empty(self.Df, dtype=torch.float32)
        cnt = torch.zeros(max(self.n_counters, 1), dtype=torch.int64)
        missing = [n for n in self.order if n not in state_dict]
        if missing:
            raise KeyError(f"state_dict lacks keys {missing[:4]}...")
        for n in self.order:
            e = self.entries[n]
            t = state_dict[n].detach().cpu()
            if tuple(t.shape) != e.shape:
                raise ValueError(f"{n}: shape {tuple(t.shape)} != expected {e.shape}")
            if e.kind == "counter":
                cnt[e.offset:e.offset + e.numel] = t.reshape(-1).long()
            else:
                if e.ohwi:
                    t = t.permute(0, 2, 3, 1)
                flat[e.offset:e.offset + e.numel] = t.reshape(-1).float()
        if device is not None:
            flat, cnt = flat.to(device), cnt.to(device)
        return flat, cnt

    def unflatten(self, flat: torch.Tensor, counters: torch.Tensor = None) -> "OrderedDict[str, torch.Tensor]":
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for n in self.order:
            e = self.entries[n]
            if e.kind == "counter":
                v = counters[e.offset:e.offset + e.numel] if counters is not None else torch.zeros(e.numel,
                                                                                                    dtype=torch.int64)
                out[n] = v.reshape(e.shape).clone().long()
            elif e.ohwi:
                o, i, kh, kw = e.shape
                out[n] = flat[e.offset:e.offset + e.numel].reshape(o, kh, kw, i).permute(0, 3, 1, 2).contiguous()
            else:
                out[n] = flat[e.offset:e.offset + e.numel].reshape(e.shape).clone()
        return out

    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        """flat [..., Df] (or [..., Dp]) -> [..., *shape] view of one float entry (no copy)."""
        e = self.entries[name]
        v = flat[..., e.offset:e.offset + e.numel]
        if e.ohwi:
            o, i, kh, kw = e.shape
            nd = flat.dim() - 1
            v = v.reshape(*flat.shape[:-1], o, kh, kw, i)
            return v.permute(*range(nd), nd, nd + 3, nd + 1, nd + 2)
        return v.reshape(*flat.shape[:-1], *e.shape)

    def counter_index(self, name: str) -> int:
        return self.entries[name].offset
