"""Native tensor-core executor for MnistNet (reference models/MnistNet.py:7-30), K client replicas at once.

    conv1 5x5 (1->20)  : fused input stage writes the 5x5 im2col patch (25 -> 32 columns), the convolution is a 1x1
                         tcgen05 GEMM with bias + ReLU epilogue on the 24x24 grid
    max-pool 2x2       : csrc/mnist_ops.cu -> 12x12 grid
    conv2 5x5 (20->50) : 25 row-shifted taps of the 'same' convolution on the 12x12 grid, stored on its centre 8x8
                         (= the unpadded convolution), bias + ReLU epilogue
    max-pool 2x2       : -> 4x4, written FLATTENED: channel n = pos*64 + c of a 1x1 grid
    fc1 800->500       : 1x1 tcgen05 convolution over those 1024 channels (weights re-indexed from torch's c*16 + pos
                         order by the pack kernel), bias + ReLU epilogue
    fc2 500->10 + CE   : the fused head kernel (log_softmax followed by cross_entropy == cross_entropy, Q20)

Backward mirrors it: head -> ReLU/bias -> weight gradient (flattened-linear index mode) + input gradient (transposed
pack) -> pool backward fused with ReLU mask and bias gradient -> 25-tap weight gradient + input gradient -> ... ->
folded weight gradient of conv1; then the flat SGD step.  No BatchNorm, no int64 counters.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch

from .. import config as C
from ..ops import flat, native
from ..ops.layout import Geom
from .lockstep import LockStepTrainer


class NativeMnist(LockStepTrainer):
    # measured on B200 (tools/gpu_step_cost.py mnist): 0.27 ms for one client ... 0.61 ms for ten
    _STEP_COST = (0.23, 0.038)
    _CUT_COST = 0.1

    def __init__(self, params, spec, data, device) -> None:
        if params["type"] != C.TYPE_MNIST:
            raise ImportError("NativeMnist covers the MNIST network only")
        self.params, self.spec, self.data, self.device = params, spec, data, device
        self.lib = native.load()
        self.compute_dtype = str(params.get("compute_dtype", "tf32"))
        self.adt, self.V = native.act_layout(self.compute_dtype)      # activation dtype, channels per 16-byte chunk
        self.B = int(params["batch_size"])
        self.label_swap = int(params["poison_label_swap"])
        self.mu, self.wd = float(params["momentum"]), float(params["decay"])
        self.eval_chunk = int(params.get("native_eval_chunk_mnist", 1000))
        self.can_eval = True
        self.can_train = bool(params.get("native_train", True))
        self.supports_dist_loss = True
        e = spec.entries
        self.e = e
        # pack layout (elements per client, bf16 or fp32)
        self.o1 = 0                                   # conv1 folded [1][32/8][32][8]
        self.o2 = self.o1 + 32 * 32                   # conv2 [25][32/8][64][8]
        self.o2t = self.o2 + 25 * 32 * 64             # conv2 transposed [25][64/8][32][8]
        self.o3 = self.o2t + 25 * 32 * 64             # fc1 [16][64/8][512][8] == [1][1024/8][512][8]
        self.o3t = self.o3 + 16 * 64 * 512            # fc1 transposed [1][512/8][1024][8]
        self.pack_elems = self.o3t + 16 * 64 * 512
        rows = [[e["conv1.weight"].offset, 20, 1, 25, self.o1, -1, 32, 32, 1],
                [e["conv2.weight"].offset, 50, 20, 25, self.o2, self.o2t, 32, 64, 0],
                [e["fc1.weight"].offset, 500, 50, 16, self.o3, self.o3t, 64, 512, 2]]
        self.pack_table = torch.tensor(rows, dtype=torch.int32, device=device).contiguous()
        # padded bias rows [32 | 64 | 512] gathered from the flat state in one index_select
        gidx, mask = [], []
        for name, n, npad in (("conv1.bias", 20, 32), ("conv2.bias", 50, 64), ("fc1.bias", 500, 512)):
            off = e[name].offset
            gidx += [off + i if i < n else off for i in range(npad)]
            mask += [1.0 if i < n else 0.0 for i in range(npad)]
        self.bias_gidx = torch.tensor(gidx, dtype=torch.long, device=device)
        self.bias_mask = torch.tensor(mask, dtype=torch.float32, device=device)
        self.boff = (0, 32, 96)
        self.trig_u8 = data.trigger_masks.to(torch.uint8).contiguous()
        self._train_bufs: Dict[int, dict] = {}
        self._eval_bufs: Dict[int, tuple] = {}

    # ------------------------------------------------------------------------------------------ helpers
    def _zbuf(self, K: int, C_: int, g: Geom, slack_rows: int = 0) -> torch.Tensor:
        """Planar zero buffer [K, C/V, R, V]; slack_rows zero rows before the first and after the last plane (the
        5x5 operand blocks reach 2*(W+1)+2 rows beyond a pass, see dba_conv_tc_valid)."""
        V = self.V
        n = K * (C_ // V) * g.rows * V
        pad = slack_rows * V
        base = torch.zeros(n + 2 * pad, dtype=self.adt, device=self.device)
        v = base[pad:pad + n].view(K, C_ // V, g.rows, V)
        v._dba_base = base          # keep the allocation alive
        return v

    def _zsplit(self, K: int, C_: int, g: Geom, slack_rows: int = 0):
        """tf32 mode: the bf16 hi/lo split copy [K, 2, C/8, R, 8] of a tensor the weight-gradient kernel reads
        (None in bf16 mode, where the tensor itself is the operand)."""
        if self.compute_dtype != "tf32":
            return None
        n = K * 2 * (C_ // 8) * g.rows * 8
        pad = slack_rows * 8
        base = torch.zeros(n + 2 * pad, dtype=torch.bfloat16, device=self.device)
        v = base[pad:pad + n].view(K, 2, C_ // 8, g.rows, 8)
        v._dba_base = base
        return v

    def pack_weights(self, W: torch.Tensor, out=None) -> torch.Tensor:
        self._enter()
        K = W.shape[0]
        if out is None:
            out = torch.zeros(K, self.pack_elems, dtype=self.adt, device=self.device)
        native.check(self.lib.dba_pack_weights(native.ptr(W), W.stride(0), native.ptr(self.pack_table), 3, K,
                                               native.ptr(out), out.stride(0), native.stream()))
        return out

    def _bias(self, W: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        torch.index_select(W, 1, self.bias_gidx, out=out)
        out.mul_(self.bias_mask)
        return out

    def _conv(self, x, pack, off, bias, bcol, K, B, H, cin, cout, ks, out, relu, vlo=0, vhi=0, x_shared=False,
              w_shared=False) -> None:
        wt = pack[:, off:]
        sh = bias[:, bcol:] if bias is not None else None
        native.check(self.lib.dba_conv_tc_valid(
            native.ptr(x), native.ptr(wt), 0 if w_shared else pack.stride(0), native.ptr(out), None, None,
            native.ptr(sh), (0 if w_shared else bias.stride(0)) if bias is not None else 0, None, 0, K, B, H, H, cin,
            cout, ks, int(relu), int(x_shared), vlo, vhi, native.stream()))

    def _geoms(self, B: int):
        return Geom(B, 24, 24), Geom(B, 12, 12), Geom(B, 1, 1)

    def _forward(self, K, B, t, pack, bias, d_x, d_y, idx, ld_idx, trow, pn, flip, labels, w_shared=False) -> None:
        """Input stage ... fc1 activations into t['h'] (shared by training and evaluation)."""
        lib, d = self.lib, self.data
        g24, g12, g1 = self._geoms(B)
        native.check(lib.dba_input_stage_image_split(
            native.ptr(d_x), native.ptr(d_y), native.ptr(idx), ld_idx, native.ptr(self.trig_u8), native.ptr(trow),
            native.ptr(pn), native.ptr(flip) if flip is not None else None, native.ptr(t["xin"]),
            native.ptr(t.get("xin_s")), native.ptr(labels),
            K, B, d.H, d.W, d.C, 5, 1, 0, 24, 24, 32, self.label_swap, g24.rows, native.stream()))
        self._conv(t["xin"], pack, self.o1, bias, self.boff[0], K, B, 24, 32, 32, 1, t["a1"], True, w_shared=w_shared)
        native.check(lib.dba_maxpool2_fwd_split(native.ptr(t["a1"]), native.ptr(t["p1"]), native.ptr(t.get("p1_s")), K,
                                                32, B, 24, 24, 0, 12, 12, 0, g24.rows, g12.rows, native.stream()))
        self._conv(t["p1"], pack, self.o2, bias, self.boff[1], K, B, 12, 32, 64, 5, t["a2"], True, 2, 10,
                   w_shared=w_shared)
        native.check(lib.dba_maxpool2_fwd_split(native.ptr(t["a2"]), native.ptr(t["f"]), native.ptr(t.get("f_s")), K,
                                                64, B, 12, 12, 2, 4, 4, 1, g12.rows, g1.rows, native.stream()))
        self._conv(t["f"], pack, self.o3, bias, self.boff[2], K, B, 1, 1024, 512, 1, t["h"], True, w_shared=w_shared)

    def _act_bufs(self, K: int, B: int, train: bool) -> dict:
        g24, g12, g1 = self._geoms(B)
        t = {"xin": self._zbuf(K, 32, g24), "a1": self._zbuf(K, 32, g24), "p1": self._zbuf(K, 32, g12, g12.Wp + 2),
             "a2": self._zbuf(K, 64, g12), "f": self._zbuf(K, 1024, g1), "h": self._zbuf(K, 512, g1),
             "labels": torch.zeros(K, B, dtype=torch.int32, device=self.device),
             "bias": torch.zeros(K, 608, device=self.device)}
        if train:
            t.update({"dh": self._zbuf(K, 512, g1), "dz3": self._zbuf(K, 512, g1), "df": self._zbuf(K, 1024, g1),
                      "dz2": self._zbuf(K, 64, g12, g12.Wp + 2), "dp1": self._zbuf(K, 32, g12),
                      "dz1": self._zbuf(K, 32, g24)})
            # tf32 mode: split copies of the weight-gradient operands (None in bf16 mode)
            t.update({"xin_s": self._zsplit(K, 32, g24), "p1_s": self._zsplit(K, 32, g12, g12.Wp + 2),
                      "f_s": self._zsplit(K, 1024, g1), "dz3_s": self._zsplit(K, 512, g1),
                      "dz2_s": self._zsplit(K, 64, g12, g12.Wp + 2), "dz1_s": self._zsplit(K, 32, g24)})
        return t

    def _wop(self, t: dict, name: str) -> torch.Tensor:
        """Weight-gradient operand form of tensor ``name``: the tensor itself (bf16 mode) or its split copy (tf32)."""
        s = t.get(name + "_s")
        return s if s is not None else t[name]

    # ---------------------------------------------------------------------------------------- evaluation
    @staticmethod
    def _shard(ids: torch.Tensor, shard) -> torch.Tensor:
        if shard is None:
            return ids
        r, w = shard
        n = ids.numel()
        per = (n + w - 1) // w
        return ids[r * per:min(n, (r + 1) * per)]

    def _eval_general(self, models: torch.Tensor, jobs, w_shared: bool, shard=None):
        self._enter()
        d, dev, lib = self.data, self.device, self.lib
        J, Bt = len(jobs), self.eval_chunk
        models = models if models.stride(-1) == 1 else models.contiguous()
        id_sets = [self._shard(d.all_test_idx if sp == "clean" else d.poison_test_idx, shard) for sp, _ in jobs]
        ns = [int(x.numel()) for x in id_sets]
        order = sorted(range(J), key=lambda j: -ns[j])        # longest input sets first (see NativeResNet)
        if order != list(range(J)):
            jobs = [jobs[j] for j in order]
            id_sets = [id_sets[j] for j in order]
            ns = [ns[j] for j in order]
            if not w_shared:
                models = models[torch.tensor(order, device=dev)]
        pack = self.pack_weights(models)
        # one set of activation buffers per pass size, with room for the largest job list seen so far (an attack round
        # adds backdoor / per-trigger evaluations: allocating a new set per job count put 0.1 s of torch.zeros into that
        # round); the kernels address rows [0, Js) only
        ent = self._eval_bufs.get(Bt)
        if ent is None or ent[0] < J:
            cap = max(J, 24 if ent is None else 2 * ent[0])
            ent = (cap, self._act_bufs(cap, Bt, train=False))
            self._eval_bufs[Bt] = ent
        t = ent[1]
        Mw = models.shape[0]
        bias = self._bias(models, t["bias"][:Mw])
        nmax = max(ns) if ns else 0
        mf = torch.zeros(J, 1, device=dev)
        mi = torch.zeros(J, 1, 2, dtype=torch.int32, device=dev)
        trow = torch.tensor([max(tr, 0) for _, tr in jobs], dtype=torch.long, device=dev)
        pn = torch.tensor([Bt if sp == "poison" else 0 for sp, _ in jobs], dtype=torch.long, device=dev)
        idx_all = torch.full((J, (nmax + Bt - 1) // Bt * Bt if nmax else Bt), -1, dtype=torch.long, device=dev)
        for j, ids in enumerate(id_sets):
            idx_all[j, :ids.numel()] = ids
        e = self.e
        g1 = Geom(Bt, 1, 1)
        for s in range(0, nmax, Bt):
            Js = sum(1 for n in ns if n > s)
            self._forward(Js, Bt, t, pack, bias, d.test_x, d.test_y, idx_all[:, s:s + Bt], idx_all.stride(0), trow, pn,
                          None, t["labels"], w_shared=w_shared)
            native.check(lib.dba_head_ex(native.ptr(t["h"]), native.ptr(models), 0 if w_shared else models.stride(0),
                                      e["fc2.weight"].offset, e["fc2.bias"].offset, None, 0, None,
                                      native.ptr(t["labels"]), 0, None, None, native.ptr(mf), native.ptr(mi), 1, None,
                                      Js, Bt, 500, 512, 1, 1, 10, g1.rows, native.stream()))
        mf_c, mi_c = mf.cpu(), mi.cpu()
        out = [None] * J
        for pos, j in enumerate(order):
            out[j] = (float(mf_c[pos, 0]), int(mi_c[pos, 0, 0]), ns[pos])
        return out

    @torch.no_grad()
    def evaluate_jobs(self, model: torch.Tensor, jobs, shard=None):
        """ONE model on several input sets (the round's global evaluations, main.py:198-231)."""
        return self._eval_general(model.reshape(1, -1), jobs, True, shard)

    @torch.no_grad()
    def evaluate_pairs(self, models: torch.Tensor, jobs, shard=None):
        """P (model, input set) pairs in one batched pass (image_train.py:148-164,268-295)."""
        return self._eval_general(models, jobs, False, shard)

    @torch.no_grad()
    def evaluate(self, models: torch.Tensor, split: str, trig_row: int, shard=None):
        res = self._eval_general(models, [(split, trig_row)] * models.shape[0], False, shard)
        loss = torch.tensor([r[0] for r in res], dtype=torch.float64, device=self.device)
        corr = torch.tensor([r[1] for r in res], dtype=torch.long, device=self.device)
        return loss, corr, (res[0][2] if res else 0)

    # ------------------------------------------------------------------------------------------ training
    def _tbufs(self, K: int) -> dict:
        if K in self._train_bufs:
            return self._train_bufs[K]
        dev, B, sp = self.device, self.B, self.spec
        t = self._act_bufs(K, B, train=True)
        ld = (sp.Dp + 3) // 4 * 4
        t["pack"] = torch.zeros(K, self.pack_elems, dtype=self.adt, device=dev)
        for key in ("Gr", "V", "gs", "anchor"):
            t[key] = torch.zeros(K, ld, device=dev)
        t["W"] = torch.zeros(K, (sp.Df + 3) // 4 * 4, device=dev)
        t["cnt"] = torch.zeros(K, sp.n_counters, dtype=torch.long, device=dev)
        t["mf"] = torch.zeros(K, self._metric_slots(), device=dev)
        t["mf_dist"] = torch.zeros(K, self._metric_slots(), device=dev)
        t["mi"] = torch.zeros(K, self._metric_slots(), 2, dtype=torch.int32, device=dev)
        t["nan"] = torch.zeros(K, dtype=torch.uint8, device=dev)
        t["dist"] = torch.zeros(K, dtype=torch.float64, device=dev)
        t["dist_scratch"] = torch.zeros(int(self.lib.dba_rows_sqdist_scratch_elems(max(K, 1))), dtype=torch.float64,
                                        device=dev)
        t["st_idx"] = torch.full((K, B), -1, dtype=torch.long, device=dev)
        t["st_flip"] = torch.zeros(K, B, dtype=torch.uint8, device=dev)
        t["st_nvalid"] = torch.zeros(K, dtype=torch.int32, device=dev)
        t["st_active"] = torch.zeros(K, dtype=torch.uint8, device=dev)
        t["st_slot"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["st_lr"] = torch.zeros(K, device=dev)
        t["trig_row"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["poison_n"] = torch.zeros(K, dtype=torch.long, device=dev)
        self._train_bufs[K] = t
        return t

    def _step_body(self, K: int, t: dict, want_grad_sum: bool, with_flip: bool, use_dist: bool = False) -> None:
        lib, d, B, e = self.lib, self.data, self.B, self.e
        Dp = self.spec.Dp
        g24, g12, g1 = self._geoms(B)
        W, Gr, V = t["W"], t["Gr"], t["V"]
        ldg = Gr.stride(0)
        pack = self.pack_weights(W, t["pack"])
        bias = self._bias(W, t["bias"])
        Gr.zero_()
        self._forward(K, B, t, pack, bias, d.train_x, d.train_y, t["st_idx"], t["st_idx"].stride(0), t["trig_row"],
                      t["poison_n"], t["st_flip"] if with_flip else None, t["labels"])
        native.check(lib.dba_head_ex(native.ptr(t["h"]), native.ptr(W), W.stride(0), e["fc2.weight"].offset,
                                  e["fc2.bias"].offset, native.ptr(Gr), ldg, native.ptr(t["dh"]),
                                  native.ptr(t["labels"]), 0, native.ptr(t["st_nvalid"]), native.ptr(t["st_slot"]),
                                  native.ptr(t["mf"]), native.ptr(t["mi"]), t["mf"].shape[1], native.ptr(t["nan"]), K, B,
                                  500, 512, 1, 1, 10, g1.rows, native.stream()))
        # fc1
        native.check(lib.dba_relu_bwd_bias_split(native.ptr(t["dh"]), native.ptr(t["h"]), native.ptr(t["dz3"]),
                                                 native.ptr(t.get("dz3_s")), native.ptr(Gr), ldg, e["fc1.bias"].offset,
                                                 500, K, 512, B, 1, 1, g1.rows, native.stream()))
        native.check(lib.dba_wgrad_linear_chw(native.ptr(self._wop(t, "dz3")), native.ptr(self._wop(t, "f")),
                                              native.ptr(Gr), ldg, e["fc1.weight"].offset, K, B, 512, 500, 64, 50, 16,
                                              native.stream()))
        self._conv(t["dz3"], pack, self.o3t, None, 0, K, B, 1, 512, 1024, 1, t["df"], False)
        # pool2 + conv2
        native.check(lib.dba_maxpool2_bwd_relu_bias_split(
            native.ptr(t["df"]), native.ptr(t["a2"]), native.ptr(t["dz2"]), native.ptr(t.get("dz2_s")), native.ptr(Gr),
            ldg, e["conv2.bias"].offset, 50, K, 64, B, 12, 12, 2, 4, 4, 1, g12.rows, g1.rows, native.stream()))
        native.check(lib.dba_wgrad_tc(native.ptr(self._wop(t, "dz2")), native.ptr(self._wop(t, "p1")), native.ptr(Gr),
                                      ldg, e["conv2.weight"].offset, K, B, 12, 12, 32, 64, 20, 50, 5, native.stream()))
        self._conv(t["dz2"], pack, self.o2t, None, 0, K, B, 12, 64, 32, 5, t["dp1"], False)
        # pool1 + conv1
        native.check(lib.dba_maxpool2_bwd_relu_bias_split(
            native.ptr(t["dp1"]), native.ptr(t["a1"]), native.ptr(t["dz1"]), native.ptr(t.get("dz1_s")), native.ptr(Gr),
            ldg, e["conv1.bias"].offset, 20, K, 32, B, 24, 24, 0, 12, 12, 0, g24.rows, g12.rows, native.stream()))
        native.check(lib.dba_wgrad_tc(native.ptr(self._wop(t, "dz1")), native.ptr(self._wop(t, "xin")), native.ptr(Gr),
                                      ldg, e["conv1.weight"].offset, K, B, 24, 24, 32, 32, 25, 20, 1, native.stream()))
        if use_dist:   # K11: distance-regularised poison loss (image_train.py:87-90)
            sc = t["dist_scratch"]
            native.check(lib.dba_rows_sqdist(native.ptr(W), W.stride(0), native.ptr(t["anchor"]), t["anchor"].stride(0),
                                             None, K, Dp, native.ptr(t["dist"]), 1, native.ptr(sc), sc.numel(),
                                             native.stream()))
            native.check(lib.dba_dist_grad(native.ptr(Gr), ldg, native.ptr(W), W.stride(0), native.ptr(t["anchor"]),
                                           t["anchor"].stride(0), native.ptr(t["dist"]), native.ptr(t["poison_n"]),
                                           native.ptr(t["st_active"]), native.ptr(t["st_slot"]),
                                           native.ptr(t["mf_dist"]), t["mf_dist"].shape[1],
                                           float(self.params["alpha_loss"]), K, Dp, native.stream()))
        flat.sgd_step_(W, V, Gr, t["st_lr"], t["st_active"], Dp, self.mu, self.wd, t["gs"] if want_grad_sum else None)
