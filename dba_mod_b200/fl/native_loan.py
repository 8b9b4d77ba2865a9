"""Native executor for the LOAN MLP: one persistent CTA per client trains the whole round (csrc/loan_mlp.cu)."""
from __future__ import annotations

import torch

from ..ops import native


class NativeLoan:
    def __init__(self, params, spec, data, device) -> None:
        self.params, self.spec, self.data, self.device = params, spec, data, device
        self.lib = native.load()
        assert spec.Dp == 5529 and data.n_features == 91
        self.label_swap = int(params["poison_label_swap"])
        self.mu, self.wd = float(params["momentum"]), float(params["decay"])
        self.trig_mask = data.trigger_mask.to(torch.uint8).contiguous()
        self.trig_val = data.trigger_val.float().contiguous()
        self.seed = int(params["seed"])
        self.calls = 0
        self.dropout = True
        self.can_train = self.can_eval = True

    def train(self, sch, W, cnt, anchor, want_grad_sum):
        from .torch_backend import TrainResult
        dev, d, lib = self.device, self.data, self.lib
        K, B, T, E = W.shape[0], int(self.params["batch_size"]), sch.T, max(sch.E_max, 1)
        gs = torch.zeros(K, self.spec.Dp, device=dev) if want_grad_sum else None
        loss = torch.zeros(K, E, device=dev)
        corr = torch.zeros(K, E, dtype=torch.int32, device=dev)
        cntm = torch.zeros(K, E, dtype=torch.int32, device=dev)
        pc = torch.zeros(K, E, dtype=torch.int32, device=dev)
        nan = torch.zeros(K, dtype=torch.uint8, device=dev)
        if T > 0:
            self.calls += 1
            native.check(lib.dba_loan_train(
                native.ptr(W), W.stride(0), native.ptr(gs), gs.stride(0) if gs is not None else 0,
                native.ptr(d.train_x), native.ptr(d.train_y), native.ptr(sch.idx.contiguous()),
                native.ptr(sch.lr.contiguous()), native.ptr(sch.active.to(torch.uint8).contiguous()),
                native.ptr(sch.epoch_of.contiguous()), native.ptr(sch.poison_n), native.ptr(sch.trig_row),
                native.ptr(self.trig_mask), native.ptr(self.trig_val), native.ptr(loss), native.ptr(corr),
                native.ptr(cntm), native.ptr(pc), native.ptr(nan), K, T, B, E, self.label_swap, self.mu, self.wd,
                (self.seed * 2654435761 + self.calls * 40503) & 0xFFFFFFFF, int(self.dropout), native.stream()))
        return TrainResult(gs, loss, corr.long(), cntm.long(), pc.long(), nan.bool())

    @torch.no_grad()
    def evaluate(self, models, split, trig_row, shard=None):
        dev, d, lib = self.device, self.data, self.lib
        models = models.contiguous()
        M = models.shape[0]
        ids = d.all_test_idx if split == "clean" else d.poison_test_idx
        if shard is not None:
            r, w = shard
            per = (int(ids.numel()) + w - 1) // w
            ids = ids[r * per:(r + 1) * per]
        ids = ids.contiguous()
        n = int(ids.numel())
        loss = torch.zeros(M, dtype=torch.float64, device=dev)
        corr = torch.zeros(M, dtype=torch.long, device=dev)
        poison = split == "poison"
        native.check(lib.dba_loan_eval(
            native.ptr(models), models.stride(0), M, native.ptr(d.test_x), native.ptr(d.test_y), native.ptr(ids), n,
            native.ptr(self.trig_mask[trig_row]) if poison else None,
            native.ptr(self.trig_val[trig_row]) if poison else None, self.label_swap, native.ptr(loss),
            native.ptr(corr), native.stream()))
        return loss, corr, n
