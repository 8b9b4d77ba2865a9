"""Native executor for the CIFAR ResNet-18 family: evaluation and training of K client replicas (or M models
under test) as chains of hand-written sm_100a kernels.

  input stage (uint8 gather, /255, trigger, stem im2col)      csrc/net_ops.cu   dba_input_stage_image
  every convolution, fwd and input-gradient                   csrc/conv_tc.cu   dba_conv_tc   (tcgen05 / TMEM / UBLKCP)
  weight gradients                                            csrc/wgrad_tc.cu  dba_wgrad_tc  (tcgen05, MN-major operands)
  BatchNorm (train: stats from the conv epilogue; apply+residual+ReLU; backward)   csrc/bn_ops.cu
  avg-pool + linear + cross-entropy + metrics (+ backward)    csrc/net_ops.cu   dba_head
  SGD / FoolsGold grad-sum                                    csrc/flat_ops.cu  dba_sgd_step

Reference behaviour: models/resnet_cifar.py:67-104 run per client by image_train.py:191-271 (training) and
test.py:7-239 (evaluation).  Evaluation folds BatchNorm into the convolution epilogue (scale/shift + residual + ReLU),
so a forward pass is 20 convolution launches + input stage + head for ALL models at once.
"""
from __future__ import annotations

import os

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from .. import config as C
from ..ops import conv as convop
from ..ops import native
from ..ops.layout import Geom
from .lockstep import LockStepTrainer

BN_EPS = 1e-5


@dataclass
class ConvL:
    name: str          # weight key
    bn: str            # BatchNorm prefix
    cin: int
    cout: int
    ks: int
    H: int             # input resolution (square)
    sub: int
    fold: bool = False
    cin_p: int = 0
    cout_p: int = 0
    w_off: int = 0     # forward pack offset (elements, per client)
    wt_off: int = -1   # transposed pack offset
    bn_off: int = 0    # offset into the per-model scale/shift rows
    d2s: bool = False  # 3x3 stride-2 layer of a space-to-depth block: transposed slot holds the depth-to-space pack


class NativeResNet(LockStepTrainer):
    def __init__(self, params, spec, data, device) -> None:
        self.params, self.spec, self.data, self.device = params, spec, data, device
        self.lib = native.load()
        self.kind = params["type"]
        # activation element mode: tf32 = fp32 activations / TF32 tensor-core math (the reference's precision class),
        # bf16 = the fast mode (half the bytes, twice the MMA rate)
        self.compute_dtype = str(params.get("compute_dtype", "tf32"))
        self.adt, self.V = native.act_layout(self.compute_dtype)
        if self.kind == C.TYPE_CIFAR:       # models/resnet_cifar.py:67-104
            self.widths, self.ress, self.ncls = [32, 64, 128, 256], [32, 16, 8, 4], 10
            self.stem = dict(ks=3, stride=1, pad=1, res=32)
            self.pool, self.head, self.skip = False, "linear", "shortcut"
        elif self.kind == C.TYPE_TINYIMAGENET:   # models/resnet_tinyimagenet.py:122-238 (torchvision layout)
            self.widths, self.ress, self.ncls = [64, 128, 256, 512], [16, 8, 4, 2], 200
            self.stem = dict(ks=7, stride=2, pad=3, res=32)
            self.pool, self.head, self.skip = True, "fc", "downsample"
        else:
            raise ImportError("the native tensor-core executor covers the ResNet-18 families (CIFAR, Tiny-ImageNet)")
        self.B = int(params["batch_size"])
        self.label_swap = int(params["poison_label_swap"])
        self.head_C, self.head_H = self.widths[-1], self.ress[-1]
        self.use_s2d = bool(params.get("native_s2d", True))
        self._plan()
        self.can_eval = True
        self.supports_dist_loss = True
        self.can_train = bool(params.get("native_train", True))
        self._train_bufs: Dict[int, dict] = {}
        self.mu, self.wd = float(params["momentum"]), float(params["decay"])
        self.eval_chunk = int(params.get("native_eval_chunk", 500))
        self.overlap_wgrad = bool(params.get("overlap_wgrad", True)) and device.type == "cuda"
        self._side = torch.cuda.Stream(device=device) if self.overlap_wgrad else None
        # BatchNorm-backward reduction sums accumulated in the epilogue of the input-gradient convolution that produces
        # the gradient (13 of the 20 layers of the CIFAR net: the stride-2 space-to-depth blocks keep the reduce kernel)
        # Off by default: measured slower (K = 1: 1.57 vs 1.49 ms / step, K = 10: 6.78 vs 5.93) - the two extra 16-byte
        # loads per stored chunk sit on the epilogue's dependent chain and cost more than the reduce kernel they save.
        self.fuse_bn_sums = (bool(params.get("fuse_bn_sums", False)) or os.environ.get("DBA_FUSE_BN_SUMS") == "1") \
            and os.environ.get("DBA_FUSE_BN_SUMS") != "0"
        self._sums_ready = set()
        self._eval_bufs: Dict[Tuple[int, int], dict] = {}
        d = data
        self.trig_u8 = d.trigger_masks.to(torch.uint8).contiguous()

    # ----------------------------------------------------------------------------------------------- plan
    def _plan(self) -> None:
        sp = self.spec
        st = self.stem
        L: List[ConvL] = [ConvL("conv1.weight", "bn1", 3, self.widths[0], st["ks"], st["res"], 1, fold=True)]
        widths = self.widths
        res = self.ress[0]
        cin = widths[0]
        self.blocks = []
        for li, w in enumerate(widths, start=1):
            for bi in range(2):
                pre = f"layer{li}.{bi}"
                sub = 2 if (li > 1 and bi == 0) else 1
                c1 = ConvL(f"{pre}.conv1.weight", f"{pre}.bn1", cin, w, 3, res, sub)
                c2 = ConvL(f"{pre}.conv2.weight", f"{pre}.bn2", w, w, 3, res // sub, 1)
                sc = None
                if f"{pre}.{self.skip}.0.weight" in sp.entries:
                    sc = ConvL(f"{pre}.{self.skip}.0.weight", f"{pre}.{self.skip}.1", cin, w, 1, res, sub)
                L += [c1, c2] + ([sc] if sc else [])
                self.blocks.append((c1, c2, sc))
                res //= sub
                cin = w
        for bi, (c1, _, _) in enumerate(self.blocks):
            c1.d2s = self._s2d_block(bi)
        off = 0
        bn_off = 0
        rows = []
        fold_rows = []
        for l in L:
            taps = l.ks * l.ks
            if l.fold:
                kk = taps * l.cin
                l.cin_p = 32 if kk <= 32 else (kk + 63) // 64 * 64       # 27 -> 32 (CIFAR), 147 -> 192 (Tiny 7x7)
                l.cout_p = (l.cout + 31) // 32 * 32
                n_f = l.cin_p * l.cout_p
            else:
                l.cin_p, l.cout_p = (l.cin + 15) // 16 * 16, (l.cout + 31) // 32 * 32
                n_f = taps * l.cin_p * l.cout_p
            l.w_off = off
            off += n_f
            if not l.fold:
                l.wt_off = off
                off += 16 * l.cin_p * l.cout_p if l.d2s else n_f      # [4 taps][Cout/8][4*Cin][8] for the d2s form
            l.bn_off = bn_off
            bn_off += l.cout
            e = sp.entries[l.name]
            rows.append([e.offset, l.cout, l.cin, taps, l.w_off, l.wt_off, l.cin_p, l.cout_p,
                         3 if l.d2s else int(l.fold)])
            ent = sp.entries
            fold_rows.append([ent[l.bn + ".weight"].offset, ent[l.bn + ".bias"].offset,
                              ent[l.bn + ".running_mean"].offset, ent[l.bn + ".running_var"].offset, l.cout, l.bn_off])
        self.layers = L
        self.pack_elems = off
        self.bn_total = bn_off
        self.pack_table = torch.tensor(rows, dtype=torch.int32, device=self.device).contiguous()
        self.fold_table = torch.tensor(fold_rows, dtype=torch.int32, device=self.device).contiguous()
        self.head_w_off = sp.entries[self.head + ".weight"].offset
        self.head_b_off = sp.entries[self.head + ".bias"].offset

    # ------------------------------------------------------------------------------------------ helpers
    def _enter(self) -> None:
        """Every public entry point selects this executor's activation mode in the native library first."""
        native.set_act_dtype(self.compute_dtype)

    def pack_weights(self, W: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        self._enter()
        K = W.shape[0]
        if out is None:
            out = torch.zeros(K, self.pack_elems, dtype=self.adt, device=self.device)
        native.check(self.lib.dba_pack_weights(native.ptr(W), W.stride(0), native.ptr(self.pack_table),
                                               len(self.layers), K, native.ptr(out), out.stride(0), native.stream()))
        return out

    def _wt(self, pack: torch.Tensor, l: ConvL, transposed: bool = False) -> torch.Tensor:
        V = self.V
        if l.fold:
            return pack[:, l.w_off:l.w_off + l.cin_p * l.cout_p].unflatten(1, (1, l.cin_p // V, l.cout_p, V))
        taps = l.ks * l.ks
        n = taps * l.cin_p * l.cout_p
        if transposed:
            return pack[:, l.wt_off:l.wt_off + n].unflatten(1, (taps, l.cout_p // V, l.cin_p, V))
        return pack[:, l.w_off:l.w_off + n].unflatten(1, (taps, l.cin_p // V, l.cout_p, V))

    def _conv(self, x, pack, l: ConvL, K, B, out, res=None, scale=None, shift=None, relu=False, stats=None,
              x_shared=False, w_shared=False, in_s2d=False, out_s2d=False):
        """Forward convolution of layer l.  pack rows are strided views -> pass raw pointers.
        w_shared: all K input sets are convolved with the SAME model (row 0 of pack / scale / shift).
        in_s2d: x is the space-to-depth form of the stride-2 layer's input (computed on the output grid only);
        out_s2d: write the output in that form for the next stage's stride-2 consumers."""
        wt = self._wt(pack, l)
        ks = 1 if l.fold else l.ks
        cin = l.cin_p
        H, sub = (l.H // l.sub, 1) if in_s2d else (l.H, l.sub)
        native.check(self.lib.dba_conv_tc_ex(
            native.ptr(x), native.ptr(wt), 0 if w_shared else pack.stride(0), native.ptr(out), native.ptr(res),
            native.ptr(scale), native.ptr(shift),
            (0 if w_shared else scale.stride(0)) if scale is not None else 0, native.ptr(stats),
            stats.stride(0) if stats is not None else 0, K, B, H, H, cin, l.cout_p, ks, sub, int(relu),
            int(x_shared), int(in_s2d), int(out_s2d), native.stream()))
        return out

    # ------------------------------------------------------------------------------------------ evaluation
    def _s2d_block(self, bi: int) -> bool:
        """Block bi opens with stride-2 convolutions whose input is written by the previous block's last
        convolution: that producer stores the space-to-depth form and the stride-2 layers compute output pixels only."""
        if not self.use_s2d or bi <= 0 or bi >= len(self.blocks):
            return False
        c1, _, sc = self.blocks[bi]
        return c1.sub == 2 and c1.ks == 3 and sc is not None and sc.sub == 2 and c1.H % 2 == 0

    def _chunk_for(self, n: int) -> int:
        """Images per evaluation pass: n split into equal passes of at most ``native_eval_chunk`` images (a rank that
        holds 1250 test images runs 2 x 640, not 2 x 1000 or 3 x 500 with a mostly empty last pass)."""
        cap = max(32, self.eval_chunk)
        if n <= 0:
            return min(cap, 256)
        passes = (n + cap - 1) // cap
        per = (n + passes - 1) // passes
        return min(cap, (per + 31) // 32 * 32)

    def _ebufs(self, M: int, Bt: int, per_job_input: bool = False) -> dict:
        """Activation buffers of an evaluation pass of M models x Bt images.  One set per (Bt, per_job_input), sized
        for the largest M seen so far; smaller M use the leading slices (the per-model stride does not depend on M)."""
        key = (Bt, per_job_input)
        full = self._eval_bufs.get(key)
        if full is None or full["cap"] < M:
            if full is not None:
                del self._eval_bufs[key]
                full = None
                torch.cuda.empty_cache()
            # (model, input set) pairs of a round: the clients' clean tests plus up to four extra tests per attacker -
            # sized once so that the first attack round does not re-allocate inside its critical path
            cap = max(M, int(self.params["no_models"]) + 4) if per_job_input and M > 1 else M
            dev = self.device
            z = lambda k, c, r: torch.zeros(k, c // self.V, Geom(Bt, r, r).rows, self.V, dtype=self.adt, device=dev)  # noqa: E731
            kin = cap if per_job_input else 1
            stem = self.layers[0]
            full = {"cap": cap, "stem_in": z(kin, stem.cin_p, stem.H), "lv": {}}
            for w, r in zip(self.widths, self.ress):
                full["lv"][r] = [z(cap, w, r) for _ in range(4)]
            # space-to-depth form of a stage's output when the next stage opens with stride-2 convolutions
            full["s2d"] = {}
            for bi, (c1, _, _) in enumerate(self.blocks):
                if self._s2d_block(bi):
                    full["s2d"][bi] = z(cap, 4 * c1.cin_p, c1.H // 2)
            if self.pool:
                full["stem_out"] = z(cap, self.widths[0], stem.H)
            full["labels"] = torch.zeros(kin, Bt, dtype=torch.int32, device=dev)
            full["idx"] = torch.full((1, Bt), -1, dtype=torch.long, device=dev)
            self._eval_bufs[key] = full
        if full["cap"] == M:
            return full
        kin = M if per_job_input else 1
        view = {"cap": M, "stem_in": full["stem_in"][:kin], "lv": {r: [b[:M] for b in bs] for r, bs in full["lv"].items()},
                "s2d": {bi: b[:M] for bi, b in full["s2d"].items()}, "labels": full["labels"][:kin], "idx": full["idx"]}
        if self.pool:
            view["stem_out"] = full["stem_out"][:M]
        return view

    @staticmethod
    def _shard(ids: torch.Tensor, shard) -> torch.Tensor:
        if shard is None:
            return ids
        r, w = shard
        n = ids.numel()
        per = (n + w - 1) // w
        return ids[r * per:min(n, (r + 1) * per)]

    @torch.no_grad()
    def evaluate_jobs(self, model: torch.Tensor, jobs, shard=None, defer: bool = False):
        """ONE model on several input sets at once: jobs = [(split, trig_row), ...] (e.g. the round's global clean /
        combined-trigger / per-trigger evaluations, main.py:198-231).  The J input sets take the place of the client
        dimension and all of them are convolved with the same packed weights (client stride 0).
        Returns [(loss_sum, correct, n)] per job (sums over this rank's shard of each split)."""
        return self._eval_general(model.reshape(1, -1), jobs, True, shard, defer)

    @torch.no_grad()
    def evaluate_pairs(self, models: torch.Tensor, jobs, shard=None, defer: bool = False):
        """P (model, input set) pairs in one batched pass: models [P, Df], jobs = [(split, trig_row)] * P.  This is
        how all per-client evaluations that follow local training (image_train.py:148-164,268-295: clean test of
        every client, the attackers' pre-/post-scaling backdoor tests and local-trigger tests) run together."""
        return self._eval_general(models, jobs, False, shard, defer)

    def _h2d(self, values, dtype) -> torch.Tensor:
        """Small host list -> device tensor without blocking the host on the current stream (torch.tensor(list,
        device=...) is a synchronous copy: a deferred evaluation on the side stream would wait for everything queued
        there before the next round could be enqueued)."""
        t = torch.tensor(values, dtype=dtype)
        if self.device.type != "cuda":
            return t.to(self.device)
        return t.pin_memory().to(self.device, non_blocking=True)

    def _eval_general(self, models: torch.Tensor, jobs, w_shared: bool, shard=None, defer: bool = False):
        """defer=True: no host read - returns a handle for finalize_eval() (the kernels are only enqueued on the current
        stream; the simulator uses this to overlap a round's global evaluations with the next round's training)."""
        self._enter()
        d, dev, lib = self.data, self.device, self.lib
        J = len(jobs)
        models = models if models.stride(-1) == 1 else models.contiguous()
        Mw = models.shape[0]
        id_sets = [self._shard(d.all_test_idx if sp == "clean" else d.poison_test_idx, shard) for sp, _ in jobs]
        ns = [int(t.numel()) for t in id_sets]
        # longest input sets first: the later passes then run on the leading jobs only (the backdoor test set is 10 %
        # shorter than the clean one - its last pass would be all padding)
        order = sorted(range(J), key=lambda j: -ns[j])
        if order != list(range(J)):
            jobs = [jobs[j] for j in order]
            id_sets = [id_sets[j] for j in order]
            ns = [ns[j] for j in order]
            if not w_shared:
                models = models[self._h2d(order, torch.long)]
        pack = self.pack_weights(models)
        scale = torch.empty(Mw, self.bn_total, device=dev)
        shift = torch.empty(Mw, self.bn_total, device=dev)
        native.check(lib.dba_bn_fold(native.ptr(models), models.stride(0), native.ptr(self.fold_table),
                                     len(self.layers), Mw, native.ptr(scale), native.ptr(shift), scale.stride(0),
                                     BN_EPS, native.stream()))
        nmax = max(ns) if ns else 0
        Bt = self._chunk_for(nmax)
        bufs = self._ebufs(J, Bt, per_job_input=True)
        mf = torch.zeros(J, 1, device=dev)
        mi = torch.zeros(J, 1, 2, dtype=torch.int32, device=dev)
        trow = self._h2d([max(tr, 0) for _, tr in jobs], torch.long)
        pn = self._h2d([Bt if sp == "poison" else 0 for sp, _ in jobs], torch.long)
        idx_all = torch.full((J, (nmax + Bt - 1) // Bt * Bt if nmax else Bt), -1, dtype=torch.long, device=dev)
        for j, t in enumerate(id_sets):
            idx_all[j, :t.numel()] = t
        labels = bufs["labels"]
        for s in range(0, nmax, Bt):
            Js = sum(1 for n in ns if n > s)           # jobs that still have images in this pass (a prefix)
            idx = idx_all[:, s:s + Bt]
            self._forward_eval(Js, Bt, idx, trow, pn, pack, scale, shift, bufs, labels, per_job=True,
                               ld_idx=idx_all.stride(0), w_shared=w_shared)
            a = bufs["out"]
            native.check(lib.dba_head(native.ptr(a), native.ptr(models), 0 if w_shared else models.stride(0),
                                      self.head_w_off, self.head_b_off, None, 0, None, native.ptr(labels), 0, None,
                                      None, native.ptr(mf), native.ptr(mi), 1, None, Js, Bt, self.head_C, self.head_H,
                                      self.head_H, self.ncls, Geom(Bt, self.head_H, self.head_H).rows,
                                      native.stream()))
        handle = (mf, mi, order, ns, (models, pack, scale, shift, idx_all, trow, pn))   # inputs kept alive until read
        return handle if defer else self.finalize_eval(handle)

    @staticmethod
    def finalize_eval(handle):
        mf, mi, order, ns, _keep = handle
        mf_c, mi_c = mf.cpu(), mi.cpu()
        out = [None] * len(order)
        for pos, j in enumerate(order):
            out[j] = (float(mf_c[pos, 0]), int(mi_c[pos, 0, 0]), ns[pos])
        return out

    @torch.no_grad()
    def evaluate(self, models: torch.Tensor, split: str, trig_row: int, shard=None):
        self._enter()
        d, dev, lib = self.data, self.device, self.lib
        M = models.shape[0]
        ids_all = self._shard(d.all_test_idx if split == "clean" else d.poison_test_idx, shard)
        n = int(ids_all.numel())
        Bt = self._chunk_for(n)
        models = models if models.stride(-1) == 1 else models.contiguous()
        pack = self.pack_weights(models)
        scale = torch.empty(M, self.bn_total, device=dev)
        shift = torch.empty(M, self.bn_total, device=dev)
        native.check(lib.dba_bn_fold(native.ptr(models), models.stride(0), native.ptr(self.fold_table),
                                     len(self.layers), M, native.ptr(scale), native.ptr(shift), scale.stride(0),
                                     BN_EPS, native.stream()))
        bufs = self._ebufs(M, Bt)
        mf = torch.zeros(M, 1, device=dev)
        mi = torch.zeros(M, 1, 2, dtype=torch.int32, device=dev)
        trow = torch.tensor([max(trig_row, 0)], dtype=torch.long, device=dev)
        pn = torch.tensor([Bt if split == "poison" else 0], dtype=torch.long, device=dev)
        idx, labels = bufs["idx"], bufs["labels"]
        for s in range(0, n, Bt):
            ids = ids_all[s:s + Bt]
            idx.fill_(-1)
            idx[0, :ids.numel()] = ids
            self._forward_eval(M, Bt, idx, trow, pn, pack, scale, shift, bufs, labels)
            a = bufs["out"]
            native.check(lib.dba_head(native.ptr(a), native.ptr(models), models.stride(0), self.head_w_off,
                                      self.head_b_off, None, 0, None, native.ptr(labels), 1, None, None,
                                      native.ptr(mf), native.ptr(mi), 1, None, M, Bt, self.head_C, self.head_H,
                                      self.head_H, self.ncls, Geom(Bt, self.head_H, self.head_H).rows,
                                      native.stream()))
        return mf[:, 0].double(), mi[:, 0, 0].long(), n

    def _forward_eval(self, M, Bt, idx, trow, pn, pack, scale, shift, bufs, labels, per_job=False,
                      ld_idx=None, w_shared=None) -> None:
        d, lib = self.data, self.lib
        stem = self.layers[0]
        xin = bufs["stem_in"]
        native.check(lib.dba_input_stage_image(
            native.ptr(d.test_x), native.ptr(d.test_y), native.ptr(idx), ld_idx if ld_idx is not None else idx.stride(0),
            native.ptr(self.trig_u8), native.ptr(trow), native.ptr(pn), None, native.ptr(xin), native.ptr(labels),
            M if per_job else 1, Bt, d.H, d.W, d.C, self.stem["ks"], self.stem["stride"], self.stem["pad"], stem.H,
            stem.H, stem.cin_p, self.label_swap, xin.shape[2], native.stream()))
        ws = per_job if w_shared is None else w_shared

        def ss(l):
            return scale[:, l.bn_off:l.bn_off + l.cout], shift[:, l.bn_off:l.bn_off + l.cout]

        lv = bufs["lv"]
        sc_, sh_ = ss(stem)
        if self.pool:
            so = bufs["stem_out"]
            self._conv(xin, pack, stem, M, Bt, so, scale=sc_, shift=sh_, relu=True, x_shared=not per_job, w_shared=ws)
            cur = lv[self.ress[0]][0]
            native.check(lib.dba_maxpool_fwd(native.ptr(so), native.ptr(cur), M, self.widths[0], Bt, stem.H, stem.H,
                                             so.shape[2], cur.shape[2], native.stream()))
        else:
            cur = lv[self.ress[0]][0]
            self._conv(xin, pack, stem, M, Bt, cur, scale=sc_, shift=sh_, relu=True, x_shared=not per_job,
                       w_shared=ws)
        res = self.ress[0]
        cur_s2d = False                           # cur holds the space-to-depth form (feeds stride-2 convolutions only)
        for bi, (c1, c2, sc) in enumerate(self.blocks):
            r_out = res // c1.sub
            pool = lv[r_out]
            free = [b for b in pool if b is not cur]
            mid, out, proj = free[0], free[1], free[2] if len(free) > 2 else None
            s1, h1 = ss(c1)
            self._conv(cur, pack, c1, M, Bt, mid, scale=s1, shift=h1, relu=True, w_shared=ws, in_s2d=cur_s2d)
            if sc is not None:
                s3, h3 = ss(sc)
                proj = free[2] if len(free) > 2 else pool[3]
                self._conv(cur, pack, sc, M, Bt, proj, scale=s3, shift=h3, relu=False, w_shared=ws, in_s2d=cur_s2d)
                skip = proj
            else:
                skip = cur
            s2, h2 = ss(c2)
            nxt_s2d = self._s2d_block(bi + 1)
            if nxt_s2d:
                out = bufs["s2d"][bi + 1]
            self._conv(mid, pack, c2, M, Bt, out, res=skip, scale=s2, shift=h2, relu=True, w_shared=ws,
                       out_s2d=nxt_s2d)
            cur, cur_s2d = out, nxt_s2d
            res = r_out
        bufs["out"] = cur

    # -------------------------------------------------------------------------------------------- training
    def _tbufs(self, K: int) -> dict:
        if K in self._train_bufs:
            return self._train_bufs[K]
        dev, B = self.device, self.B

        def z(c, r):
            return torch.zeros(K, c // self.V, Geom(B, r, r).rows, self.V, dtype=self.adt, device=dev)

        split = self.compute_dtype == "tf32"

        def zs(c, r):
            """bf16 hi/lo split copy of a z(c, r) tensor: the weight-gradient operand form in tf32 mode."""
            if not split:
                return None
            return torch.zeros(K, 2, c // 8, Geom(B, r, r).rows, 8, dtype=torch.bfloat16, device=dev)

        stem = self.layers[0]
        s2d_sc = {sc.name for bi, (_, _, sc) in enumerate(self.blocks) if sc is not None and self._s2d_block(bi)}
        t = {"xin": z(stem.cin_p, stem.H), "labels": torch.zeros(K, B, dtype=torch.int32, device=dev)}
        t[("split", "xin")] = zs(stem.cin_p, stem.H)
        for l in self.layers:
            ro = l.H // l.sub
            t[("y", l.name)] = z(l.cout_p, ro)
            t[("a", l.name)] = z(l.cout_p, ro)
            t[("split", "a", l.name)] = zs(l.cout_p, ro)
            if l.fold:
                rdy = ro
            elif l.d2s or (l.ks == 1 and l.sub == 2 and l.name in s2d_sc):
                rdy = ro                                      # stride-2 layers of a space-to-depth block: compact
            else:
                rdy = l.H                                     # on the input grid (zero-upsampled when sub == 2)
            t[("dy", l.name)] = z(l.cout_p, rdy)
            t[("split", "dy", l.name)] = zs(l.cout_p, rdy)
        # gradient scratch per resolution level: g (block output grad), dz, da1, dx
        for w, r in zip(self.widths, self.ress):
            t[("g", r)] = [z(w, r) for _ in range(4)]
        for bi, (c1, _, _) in enumerate(self.blocks):      # forward operands of the stride-2 layers (output pixels only)
            if self._s2d_block(bi):
                t[("s2d", bi)] = z(4 * c1.cin_p, c1.H // 2)
                t[("split", "s2d", bi)] = zs(4 * c1.cin_p, c1.H // 2)
        # plain tensor -> its split copy (looked up by the weight-gradient launcher)
        t["split_of"] = {}
        for key, v in list(t.items()):
            if isinstance(key, tuple) and key[0] == "split" and v is not None:
                src = t[key[1]] if len(key) == 2 else t[key[1:]]
                t["split_of"][src.data_ptr()] = v
        if self.pool:
            t["pool_out"] = z(self.widths[0], self.ress[0])
            t[("split", "pool_out")] = zs(self.widths[0], self.ress[0])
            if split:
                t["split_of"][t["pool_out"].data_ptr()] = t[("split", "pool_out")]
            t["pool_g"] = z(self.widths[0], stem.H)
        t["stats"] = torch.zeros(K, 2 * self.bn_total, device=dev)
        t["saved"] = torch.zeros(K, 2 * self.bn_total, device=dev)
        t["sums"] = torch.zeros(K, 2 * self.bn_total, device=dev)
        t["Gr"] = torch.zeros(K, (self.spec.Dp + 3) // 4 * 4, device=dev)
        t["V"] = torch.zeros_like(t["Gr"])
        t["gs"] = torch.zeros_like(t["Gr"])
        t["W"] = torch.zeros(K, (self.spec.Df + 3) // 4 * 4, device=dev)
        t["cnt"] = torch.zeros(K, self.spec.n_counters, dtype=torch.long, device=dev)
        t["mf"] = torch.zeros(K, self._metric_slots(), device=dev)
        t["mf_dist"] = torch.zeros(K, self._metric_slots(), device=dev)
        t["anchor"] = torch.zeros(K, (self.spec.Dp + 3) // 4 * 4, device=dev)
        t["dist"] = torch.zeros(K, dtype=torch.float64, device=dev)
        t["dist_scratch"] = torch.zeros(int(self.lib.dba_rows_sqdist_scratch_elems(max(K, 1))), dtype=torch.float64,
                                        device=dev)
        t["mi"] = torch.zeros(K, self._metric_slots(), 2, dtype=torch.int32, device=dev)
        t["nan"] = torch.zeros(K, dtype=torch.uint8, device=dev)
        t["st_idx"] = torch.full((K, B), -1, dtype=torch.long, device=dev)
        t["st_flip"] = torch.zeros(K, B, dtype=torch.uint8, device=dev)
        t["st_nvalid"] = torch.zeros(K, dtype=torch.int32, device=dev)
        t["st_active"] = torch.zeros(K, dtype=torch.uint8, device=dev)
        t["st_active_l"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["st_slot"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["st_lr"] = torch.zeros(K, dtype=torch.float32, device=dev)
        t["trig_row"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["poison_n"] = torch.zeros(K, dtype=torch.long, device=dev)
        t["pack"] = torch.zeros(K, self.pack_elems, dtype=self.adt, device=dev)
        self._train_bufs[K] = t
        return t

    def _bn_apply(self, l: ConvL, K, y, out, res, relu, W, t, nvalid, active, out_s2d=None):
        e = self.spec.entries
        r = l.H // l.sub
        so = t["split_of"]
        native.check(self.lib.dba_bn_apply_split(
            native.ptr(y), native.ptr(res), native.ptr(out), native.ptr(out_s2d),
            out_s2d.shape[2] if out_s2d is not None else 0, native.ptr(so.get(out.data_ptr())),
            native.ptr(so.get(out_s2d.data_ptr())) if out_s2d is not None else None,
            native.ptr(t["stats"]), t["stats"].stride(0),
            2 * l.bn_off, native.ptr(t["saved"]), t["saved"].stride(0), 2 * l.bn_off, native.ptr(W), W.stride(0),
            e[l.bn + ".weight"].offset, e[l.bn + ".bias"].offset, e[l.bn + ".running_mean"].offset,
            e[l.bn + ".running_var"].offset, native.ptr(nvalid), native.ptr(active), K, l.cout_p, self.B, r, r,
            y.shape[2], int(relu), 0.1, BN_EPS, native.stream()))

    def _bn_bwd(self, l: ConvL, K, g, y, act, dy, gmask, W, t, nvalid, up):
        e = self.spec.entries
        r = l.H // l.sub
        Gr = t["Gr"]
        ready = l.name in self._sums_ready        # accumulated by the convolution that produced g (see _dgrad)
        native.check(self.lib.dba_bn_backward_ex(
            native.ptr(g), native.ptr(y), native.ptr(act), native.ptr(dy), native.ptr(t["split_of"].get(dy.data_ptr())),
            native.ptr(gmask), native.ptr(t["sums"]),
            t["sums"].stride(0), 2 * l.bn_off, native.ptr(t["saved"]), t["saved"].stride(0), 2 * l.bn_off,
            native.ptr(W), W.stride(0), native.ptr(Gr), Gr.stride(0), e[l.bn + ".weight"].offset,
            e[l.bn + ".bias"].offset, native.ptr(nvalid), K, l.cout_p, self.B, r, r, y.shape[2], int(up),
            dy.shape[2], int(ready), native.stream()))

    def _wgrad(self, l: ConvL, K, dy, x, t, s2d=False):
        Gr = t["Gr"]
        e = self.spec.entries[l.name]
        if l.fold:
            cin, cin_real, ks, H = l.cin_p, l.ks * l.ks * l.cin, 1, l.H
        else:
            cin, cin_real, ks, H = l.cin_p, l.cin, l.ks, l.H
        if self.compute_dtype == "tf32":        # operands = the bf16 hi/lo split copies stored next to the fp32 tensors
            dy, x = t["split_of"][dy.data_ptr()], t["split_of"][x.data_ptr()]
        xcs = 0
        if s2d:             # x is the space-to-depth tensor: its first cin/8 planes are the even-position pixels
            H, xcs = l.H // 2, x.stride(0)

        def launch():
            if s2d and ks == 3:
                native.check(self.lib.dba_wgrad_tc_s2d(native.ptr(dy), native.ptr(x), native.ptr(Gr), Gr.stride(0),
                                                       e.offset, K, self.B, H, H, cin, l.cout_p, cin_real, l.cout,
                                                       native.stream()))
            else:
                native.check(self.lib.dba_wgrad_tc_ex(native.ptr(dy), native.ptr(x), xcs, native.ptr(Gr), Gr.stride(0),
                                                      e.offset, K, self.B, H, H, cin, l.cout_p, cin_real, l.cout, ks,
                                                      native.stream()))

        if not self.overlap_wgrad:
            launch()
            return
        # weight gradients are off the critical path (only the optimiser step needs them): they run on a side stream
        # (a parallel branch of the captured graph) next to the input-gradient / BatchNorm-backward chain
        main = torch.cuda.current_stream()
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            launch()

    def _dgrad(self, l: ConvL, K, dy, pack, out, res=None, bn=None, t=None):
        """Input gradient of layer l.  bn = (layer, y, act): the BatchNorm whose backward consumes ``out``; its
        reduction sums are then accumulated in this convolution's epilogue (no bn_bwd_reduce launch for that layer)."""
        wt = self._wt(pack, l, transposed=True)
        if bn is not None and self.fuse_bn_sums:
            bl, by, ba = bn
            sums, saved = t["sums"], t["saved"]
            native.check(self.lib.dba_conv_tc_dgrad_bn(
                native.ptr(dy), native.ptr(wt), pack.stride(0), native.ptr(out), native.ptr(res), native.ptr(by),
                native.ptr(ba), native.ptr(saved[:, 2 * bl.bn_off:]), saved.stride(0),
                native.ptr(sums[:, 2 * bl.bn_off:]), sums.stride(0), K, self.B, l.H, l.H, l.cout_p, l.cin_p, l.ks,
                native.stream()))
            self._sums_ready.add(bl.name)
            return out
        native.check(self.lib.dba_conv_tc_strided(
            native.ptr(dy), native.ptr(wt), pack.stride(0), native.ptr(out), native.ptr(res), None, None, 0, None, 0,
            K, self.B, l.H, l.H, l.cout_p, l.cin_p, l.ks, 1, 0, 0, native.stream()))
        return out

    def _step_body(self, K: int, t: dict, want_grad_sum: bool, with_flip: bool, use_dist: bool = False) -> None:
        """One lock-step SGD step of K clients: ~125 kernel launches, no host decisions (CUDA-graph capturable).
        All inputs come from the persistent staging tensors st_* filled by train()."""
        from ..ops import flat
        self._enter()
        lib, d, B, Dp = self.lib, self.data, self.B, self.spec.Dp
        W, Gr, V = t["W"], t["Gr"], t["V"]
        nvalid, active, slot = t["st_nvalid"], t["st_active"], t["st_slot"]
        mf, mi, nan = t["mf"], t["mi"], t["nan"]
        stem = self.layers[0]
        lv_g = {r: t[("g", r)] for r in self.ress}
        pack = self.pack_weights(W, t["pack"])
        t["stats"].zero_(); Gr.zero_()
        if not lib.dba_get_bn_fused() or self.fuse_bn_sums:     # scratch of the BatchNorm-backward reduction sums
            t["sums"].zero_()
        self._sums_ready = set()
        native.check(lib.dba_input_stage_image_split(
            native.ptr(d.train_x), native.ptr(d.train_y), native.ptr(t["st_idx"]), t["st_idx"].stride(0),
            native.ptr(self.trig_u8), native.ptr(t["trig_row"]), native.ptr(t["poison_n"]),
            native.ptr(t["st_flip"]) if with_flip else None, native.ptr(t["xin"]), native.ptr(t[("split", "xin")]),
            native.ptr(t["labels"]), K, B, d.H, d.W, d.C, self.stem["ks"], self.stem["stride"], self.stem["pad"],
            stem.H, stem.H, stem.cin_p, self.label_swap, t["xin"].shape[2], native.stream()))
        st = t["stats"]

        def fconv(l, x, y):
            self._conv(x, pack, l, K, B, y, stats=st[:, 2 * l.bn_off:], x_shared=False)

        # ---------------------------------------------------------------------- forward
        y0, a0 = t[("y", stem.name)], t[("a", stem.name)]
        fconv(stem, t["xin"], y0)
        self._bn_apply(stem, K, y0, a0, None, True, W, t, nvalid, active)
        x = a0
        if self.pool:
            x = t["pool_out"]
            native.check(lib.dba_maxpool_fwd_split(native.ptr(a0), native.ptr(x), native.ptr(t[("split", "pool_out")]), K,
                                                   self.widths[0], B, stem.H, stem.H, a0.shape[2], x.shape[2],
                                                   native.stream()))
        inputs = []
        x_s2d = None           # space-to-depth copy of x when the block about to run opens with stride-2 convolutions
        for bi, (c1, c2, sc) in enumerate(self.blocks):
            inputs.append(x)
            y1, a1 = t[("y", c1.name)], t[("a", c1.name)]
            if x_s2d is not None:
                self._conv(x_s2d, pack, c1, K, B, y1, stats=st[:, 2 * c1.bn_off:], in_s2d=True)
            else:
                fconv(c1, x, y1)
            self._bn_apply(c1, K, y1, a1, None, True, W, t, nvalid, active)
            y2, out = t[("y", c2.name)], t[("a", c2.name)]
            fconv(c2, a1, y2)
            skip = x
            if sc is not None:
                ys, as_ = t[("y", sc.name)], t[("a", sc.name)]
                if x_s2d is not None:
                    self._conv(x_s2d, pack, sc, K, B, ys, stats=st[:, 2 * sc.bn_off:], in_s2d=True)
                else:
                    fconv(sc, x, ys)
                self._bn_apply(sc, K, ys, as_, None, False, W, t, nvalid, active)
                skip = as_
            nxt = t.get(("s2d", bi + 1))
            self._bn_apply(c2, K, y2, out, skip, True, W, t, nvalid, active, out_s2d=nxt)
            x, x_s2d = out, nxt
        # ---------------------------------------------------------------------- head (fwd + bwd)
        g = lv_g[self.head_H][0]
        native.check(lib.dba_head(native.ptr(x), native.ptr(W), W.stride(0), self.head_w_off, self.head_b_off,
                                  native.ptr(Gr), Gr.stride(0), native.ptr(g), native.ptr(t["labels"]), 0,
                                  native.ptr(nvalid), native.ptr(slot), native.ptr(mf), native.ptr(mi), mf.shape[1],
                                  native.ptr(nan), K, B, self.head_C, self.head_H, self.head_H, self.ncls, x.shape[2],
                                  native.stream()))
        # ---------------------------------------------------------------------- backward
        for bi in range(len(self.blocks) - 1, -1, -1):
            c1, c2, sc = self.blocks[bi]
            xin_b = inputs[bi]
            r_out, r_in = c2.H, c1.H
            pool_o = [b for b in lv_g[r_out] if b is not g]
            dz, da1 = pool_o[0], pool_o[1]
            dy2, dy1 = t[("dy", c2.name)], t[("dy", c1.name)]
            # the BatchNorm that will consume this block's input gradient: the previous block's bn2 (or the stem's)
            if bi > 0:
                c2p = self.blocks[bi - 1][1]
                cons = (c2p, t[("y", c2p.name)], t[("a", c2p.name)])
            else:
                cons = None if self.pool else (stem, y0, a0)
            self._bn_bwd(c2, K, g, t[("y", c2.name)], t[("a", c2.name)], dy2, dz, W, t, nvalid, up=False)
            self._wgrad(c2, K, dy2, t[("a", c1.name)], t)
            self._dgrad(c2, K, dy2, pack, da1, bn=(c1, t[("y", c1.name)], t[("a", c1.name)]), t=t)
            xs = t.get(("s2d", bi)) if c1.d2s else None
            self._bn_bwd(c1, K, da1, t[("y", c1.name)], t[("a", c1.name)], dy1, None, W, t, nvalid,
                         up=(c1.sub == 2 and xs is None))
            if xs is not None:
                self._wgrad(c1, K, dy1, xs, t, s2d=True)
            else:
                self._wgrad(c1, K, dy1, xin_b, t)
            pool_i = [b for b in lv_g[r_in] if b is not g and b is not dz and b is not da1]
            dx = pool_i[0]
            if sc is None:
                self._dgrad(c1, K, dy1, pack, dx, res=dz, bn=cons, t=t)
            else:
                dx2 = pool_i[1]
                dys = t[("dy", sc.name)]
                if xs is not None:
                    # Space-to-depth block.  3x3 stride-2 input gradient from the compact dy: 2x2 taps, depth-to-space
                    # store (every pixel of the full-resolution gradient is written exactly once) ...
                    ro = c1.H // 2
                    wd = pack[:, c1.wt_off:]
                    native.check(self.lib.dba_conv_tc_ex(
                        native.ptr(dy1), native.ptr(wd), pack.stride(0), native.ptr(dx2), None, None, None, 0, None, 0,
                        K, self.B, ro, ro, c1.cout_p, 4 * c1.cin_p, 2, 1, 0, 0, 0, 3, native.stream()))
                    # ... and the 1x1 stride-2 shortcut on its own (half-resolution) grid: compact dys, weight gradient
                    # against the parity-0 planes of the input, input gradient added in place at the even positions
                    self._bn_bwd(sc, K, dz, t[("y", sc.name)], None, dys, None, W, t, nvalid, up=False)
                    self._wgrad(sc, K, dys, xs, t, s2d=True)
                    wt = self._wt(pack, sc, transposed=True)
                    native.check(self.lib.dba_conv_tc_ex(
                        native.ptr(dys), native.ptr(wt), pack.stride(0), native.ptr(dx2), native.ptr(dx2), None, None, 0,
                        None, 0, K, self.B, ro, ro, sc.cout_p, sc.cin_p, 1, 1, 0, 0, 0, 2, native.stream()))
                    dx = dx2
                else:
                    self._dgrad(c1, K, dy1, pack, dx2)
                    self._bn_bwd(sc, K, dz, t[("y", sc.name)], None, dys, None, W, t, nvalid, up=True)
                    self._wgrad(sc, K, dys, xin_b, t)
                    self._dgrad(sc, K, dys, pack, dx, res=dx2, bn=cons, t=t)
            g = dx
        if self.pool:
            native.check(lib.dba_maxpool_bwd(native.ptr(a0), native.ptr(g), native.ptr(t["pool_g"]), K, self.widths[0], B,
                                             stem.H, stem.H, a0.shape[2], g.shape[2], native.stream()))
            g = t["pool_g"]
        dy0 = t[("dy", stem.name)]
        self._bn_bwd(stem, K, g, y0, a0, dy0, None, W, t, nvalid, up=False)
        self._wgrad(stem, K, dy0, t["xin"], t)
        if use_dist:   # K11: distance-regularised poison loss, gradient term added before the optimiser step
            sc = t["dist_scratch"]
            native.check(lib.dba_rows_sqdist(native.ptr(W), W.stride(0), native.ptr(t["anchor"]), t["anchor"].stride(0),
                                             None, K, Dp, native.ptr(t["dist"]), 1, native.ptr(sc), sc.numel(),
                                             native.stream()))
            native.check(lib.dba_dist_grad(native.ptr(Gr), Gr.stride(0), native.ptr(W), W.stride(0),
                                           native.ptr(t["anchor"]), t["anchor"].stride(0), native.ptr(t["dist"]),
                                           native.ptr(t["poison_n"]), native.ptr(active), native.ptr(slot),
                                           native.ptr(t["mf_dist"]), t["mf_dist"].shape[1],
                                           float(self.params["alpha_loss"]), K, Dp, native.stream()))
        if self.overlap_wgrad:
            torch.cuda.current_stream().wait_stream(self._side)
        # ---------------------------------------------------------------------- optimiser
        flat.sgd_step_(W, V, Gr, t["st_lr"], active, Dp, self.mu, self.wd, t["gs"] if want_grad_sum else None)
        if t["cnt"].numel():
            t["cnt"].add_(t["st_active_l"].copy_(active)[:, None])
