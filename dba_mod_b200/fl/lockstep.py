"""Lock-step local training of K client replicas on persistent slabs (shared by the native executors).

A subclass provides ``_tbufs(K)`` (persistent per-width buffers: W, V, gs, Gr, cnt, anchor, mf, mi, mf_dist, nan,
trig_row, poison_n and the st_* staging tensors) and ``_step_body(K, t, want_grad_sum, with_flip, use_dist)`` (one SGD
step of all K clients, no host decisions, CUDA-graph capturable).  This class owns the round-level logic: schedule
segmentation (finished clients leave the batch), staging, graph capture / replay, start-up pre-warming and the
TrainResult (reference: the per-client loops of image_train.py:21-315, run for all clients at once).
"""
from __future__ import annotations

from typing import List

import torch

from ..ops import native


class LockStepTrainer:
    # cost model of one lock-step step over k clients: a + b*k (ms).  The class value is only the fallback (CIFAR
    # ResNet-18, tf32, B = 64, graph replay with programmatic dependent launch: 1.70 ms at k = 1 ... 5.95 ms at k = 10,
    # tools/gpu_step_cost.py); ``prewarm`` replaces it per executor by timing the captured step graphs of the model
    # actually being trained.  _CUT_COST: moving the carried state to another slab.
    _STEP_COST = (1.23, 0.472)
    _CUT_COST = 0.25

    # Slab widths: every distinct width owns a full set of activation buffers and a captured graph, so segments run on
    # the next width of a geometric ladder (at most ~33 % idle rows) instead of on their exact client count; with 80
    # clients per round that is 13 slab sets instead of up to 80.
    _LADDER = (1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512)

    compute_dtype = "bf16"

    def _metric_slots(self) -> int:
        """Per-internal-epoch metric slots of a slab: the longest local schedule the configuration can produce (the
        reference logs one line per internal epoch, image_train.py:121-133), at least 16."""
        p = self.params
        return max(16, int(p.get("internal_epochs", 1)), int(p.get("internal_poison_epochs", 1)))

    def _enter(self) -> None:
        """Selects this executor's activation element mode in the native library (library-wide switch)."""
        native.set_act_dtype(self.compute_dtype)

    @classmethod
    def _width(cls, k: int, K: int) -> int:
        for w in cls._LADDER:
            if w >= k:
                return min(w, K)
        return K

    @classmethod
    def _segments(cls, steps: List[int], max_cuts: int = 4, cost=None) -> List[tuple]:
        """Split the lock-step schedule into [(rows, lo, hi)] so that clients that finished stop being computed.

        Clients hold different numbers of samples (Dirichlet partition) and attackers train
        ``internal_poison_epochs`` x longer, so T = max steps while the average client is done much earlier: without
        this a round of 10 clients runs at ~57% utilisation.  Cut points are chosen by dynamic programming over the
        distinct step counts (cost model above); each segment runs on a slab sized for the clients still active."""
        a, b = cost if cost is not None else cls._STEP_COST
        K = len(steps)
        T = max(steps) if steps else 0
        if K == 0 or T == 0:
            return [(list(range(K)), 0, T)]
        cuts = sorted({s for s in steps if 0 < s < T})          # candidate boundaries
        pts = [0] + cuts + [T]

        def active(lo):
            return [r for r in range(K) if steps[r] > lo]

        n = len(pts)
        INF = float("inf")
        # best[j][c] = minimal cost of covering [pts[j], T) with at most c further cuts
        best = [[INF] * (max_cuts + 1) for _ in range(n)]
        nxt = [[-1] * (max_cuts + 1) for _ in range(n)]
        for c in range(max_cuts + 1):
            best[n - 1][c] = 0.0
        for j in range(n - 2, -1, -1):
            k = cls._width(len(active(pts[j])), K)
            for c in range(max_cuts + 1):
                for m in range(j + 1, n):
                    if m < n - 1 and c == 0:
                        continue
                    tail = best[m][c - 1] + cls._CUT_COST if m < n - 1 else 0.0
                    cost = (pts[m] - pts[j]) * (a + b * k) + tail
                    if cost < best[j][c] - 1e-12:
                        best[j][c], nxt[j][c] = cost, m
        segs, j, c = [], 0, max_cuts
        while j < n - 1:
            m = nxt[j][c]
            segs.append((active(pts[j]), pts[j], pts[m]))
            j, c = m, c - 1
        return segs

    def train(self, sch, W: torch.Tensor, cnt: torch.Tensor, anchor: torch.Tensor, want_grad_sum: bool):
        from .torch_backend import TrainResult
        p, dev = self.params, self.device
        K, Dp, B, Df = W.shape[0], self.spec.Dp, self.B, self.spec.Df
        alpha = float(p["alpha_loss"])
        use_dist = alpha != 1.0 and bool((sch.poison_n > 0).any())
        E = max(sch.E_max, 1)
        t = self._tbufs(K)
        if E > t["mf"].shape[1]:
            raise ValueError(f"schedule with {E} internal epochs, slabs were sized for {t['mf'].shape[1]} "
                             "(internal_epochs / internal_poison_epochs changed after the executor was built?)")
        # persistent slabs (their addresses are baked into the captured step graph)
        t["W"][:, :Df].copy_(W[:, :Df])
        if cnt.numel():
            t["cnt"].copy_(cnt)
        for key in ("V", "mf", "mi", "nan", "gs", "mf_dist"):
            t[key].zero_()
        if use_dist:
            t["anchor"][:, :Dp].copy_(anchor[:, :Dp])
        t["trig_row"].copy_(sch.trig_row)
        t["poison_n"].copy_(sch.poison_n)
        mask_all = sch.idx >= 0                                              # [K, T, B]
        ar = torch.arange(B, device=dev)
        sel = (ar[None, None, :] < sch.poison_n[:, None, None]) & mask_all
        pcount = torch.zeros(K, E, dtype=torch.long, device=dev)
        pcount.scatter_add_(1, sch.epoch_of, sel.sum(2) * sch.active)
        per_step = {"idx": sch.idx, "nvalid": mask_all.sum(2).to(torch.int32), "active": sch.active.to(torch.uint8),
                    "slot": sch.epoch_of, "lr": sch.lr,
                    "flip": sch.flip.to(torch.uint8) if sch.flip is not None else None}
        flags = (bool(want_grad_sum), sch.flip is not None, use_dist)
        steps = list(sch.steps_host) if len(sch.steps_host) == K else [sch.T] * K
        segs = self._segments(steps, cost=self._STEP_COST) if bool(p.get("compact_tail", True)) \
            else [(list(range(K)), 0, sch.T)]
        carry = ("V", "gs", "anchor") if use_dist else ("V", "gs")
        for rows, lo, hi in segs:
            if self._width(len(rows), K) == K:       # (nearly) everybody still active: stay on the home slab
                self._run_steps(K, t, per_step, None, lo, hi, flags)
                continue
            r = torch.tensor(rows, dtype=torch.long, device=dev)
            n = len(rows)
            w = self._width(n, K)
            t2 = self._tbufs(w)
            for key in carry + ("W", "cnt", "trig_row", "poison_n"):
                t2[key][:n].copy_(t[key][r])
            for key in ("mf", "mi", "nan", "mf_dist"):
                t2[key].zero_()
            if w > n:           # spare rows of the slab: idle clients (no samples, no update)
                t2["st_idx"][n:].fill_(-1)
                for key in ("st_nvalid", "st_active", "st_lr", "poison_n"):
                    t2[key][n:].zero_()
            self._run_steps(w, t2, per_step, r, lo, hi, flags)
            for key in ("V", "gs", "W", "cnt"):
                t[key][r] = t2[key][:n]
            for key in ("mf", "mi", "mf_dist"):
                t[key][r] += t2[key][:n]
            t["nan"][r] |= t2["nan"][:n]
        W[:, :Df].copy_(t["W"][:, :Df])
        if cnt.numel():
            cnt.copy_(t["cnt"])
        mf, mi = t["mf"][:, :E].clone(), t["mi"][:, :E].clone()
        if use_dist:       # logged loss of poisoning clients = alpha*CE + (1-alpha)*distance (image_train.py:88-89,103)
            pz = (sch.poison_n > 0)[:, None]
            mf = torch.where(pz, alpha * mf + (1.0 - alpha) * t["mf_dist"][:, :E], mf)
        res = TrainResult(t["gs"][:, :Dp].clone() if want_grad_sum else None, mf, mi[:, :, 0].long(),
                          mi[:, :, 1].long(), pcount, t["nan"].bool().clone())
        return res

    def _run_steps(self, K: int, t: dict, per_step: dict, rows, lo: int, hi: int, flags: tuple) -> None:
        """Steps [lo, hi) of the clients ``rows`` (None = all) on the K-client slab ``t``: stage the step's index /
        lr / mask rows, then replay the captured step graph (first use of a slab: one eager step, then capture)."""
        with_flip = flags[1]
        use_graph = bool(self.params.get("cuda_graphs", True))

        def pick(x, step):
            return x[:, step] if rows is None else x[rows, step]

        n = K if rows is None else int(rows.numel())
        ps = per_step
        native_stage = ps["idx"].is_cuda and all(ps[k].is_contiguous() for k in ("idx", "nvalid", "active", "slot", "lr")) \
            and (not with_flip or ps["flip"].is_contiguous())
        T, B = int(ps["idx"].shape[1]), int(ps["idx"].shape[2])
        for step in range(lo, hi):
            if native_stage:          # one launch instead of six indexed copies
                native.check(self.lib.dba_stage_step(
                    native.ptr(ps["idx"]), native.ptr(ps["flip"]) if with_flip else None, native.ptr(ps["nvalid"]),
                    native.ptr(ps["active"]), native.ptr(ps["slot"]), native.ptr(ps["lr"]), native.ptr(rows), n, T, B,
                    step, native.ptr(t["st_idx"]), native.ptr(t["st_flip"]) if with_flip else None,
                    native.ptr(t["st_nvalid"]), native.ptr(t["st_active"]), native.ptr(t["st_slot"]),
                    native.ptr(t["st_lr"]), native.stream()))
            else:
                t["st_idx"][:n].copy_(pick(ps["idx"], step))
                t["st_nvalid"][:n].copy_(pick(ps["nvalid"], step))
                t["st_active"][:n].copy_(pick(ps["active"], step))
                t["st_slot"][:n].copy_(pick(ps["slot"], step))
                t["st_lr"][:n].copy_(pick(ps["lr"], step))
                if with_flip:
                    t["st_flip"][:n].copy_(pick(ps["flip"], step))
            self._replay(K, t, flags, use_graph)

    def _replay(self, K: int, t: dict, flags: tuple, use_graph: bool = True) -> None:
        """One step on slab t: eager the first time (lazy kernel attributes), then capture, then graph replays."""
        want_grad_sum, with_flip, use_dist = flags
        self._enter()
        gkey = ("graph",) + flags
        g = t.get(gkey)
        if not use_graph:
            self._step_body(K, t, want_grad_sum, with_flip, use_dist)
        elif g is None:
            if not t.get(("warm", gkey)):
                self._step_body(K, t, want_grad_sum, with_flip, use_dist)
                t[("warm", gkey)] = True
                return
            l0 = native.launch_count()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._step_body(K, t, want_grad_sum, with_flip, use_dist)
            t[gkey] = graph
            t[("nl", gkey)] = native.launch_count() - l0
            graph.replay()
        else:
            g.replay()
            self.lib.dba_add_launch_count(t[("nl", gkey)])

    def prewarm(self, Kmax: int, want_grad_sum: bool) -> None:
        """Allocate the slabs and capture the step graphs of every batch width 1..Kmax up front (all-idle dummy steps:
        no client is active, nothing is updated), so that no round pays for a capture in its critical path."""
        if not bool(self.params.get("cuda_graphs", True)):
            return
        p = self.params
        with_flip = bool(getattr(self.data, "train_flip", False))
        dists = [False] + ([True] if (p["is_poison"] and float(p["alpha_loss"]) != 1.0) else [])
        if Kmax > 16:            # many clients per GPU: slabs are created on first use instead
            return
        for K in range(1, Kmax + 1):       # every K can be a round's home slab; the ladder widths are among them
            t = self._tbufs(K)
            t["st_idx"].fill_(-1); t["st_nvalid"].zero_(); t["st_active"].zero_(); t["st_lr"].zero_()
            for use_dist in dists:
                flags = (bool(want_grad_sum), with_flip, use_dist)
                for _ in range(2):
                    self._replay(K, t, flags)
            for key in ("V", "mf", "mi", "nan", "gs", "mf_dist", "W"):
                t[key].zero_()
            if t["cnt"].numel():
                t["cnt"].zero_()
        self._calibrate_step_cost(Kmax, (bool(want_grad_sum), with_flip, False))

    def _calibrate_step_cost(self, Kmax: int, flags: tuple) -> None:
        """a + b*k from the captured step graphs of THIS model / mode / batch size: replay k = 1 and k = Kmax a few
        times on idle slabs (the kernels do the same work for padding samples) and time them with CUDA events."""
        if Kmax < 2 or self.device.type != "cuda":
            return
        gkey = ("graph",) + flags
        ms = {}
        for K in (1, Kmax):
            g = self._tbufs(K).get(gkey)
            if g is None:
                return
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g.replay()
            e0.record()
            for _ in range(3):
                g.replay()
            e1.record()
            torch.cuda.synchronize(self.device)
            ms[K] = e0.elapsed_time(e1) / 3
            t = self._tbufs(K)
            for key in ("V", "mf", "mi", "nan", "gs", "mf_dist", "W"):
                t[key].zero_()
            if t["cnt"].numel():
                t["cnt"].zero_()
        b = max((ms[Kmax] - ms[1]) / (Kmax - 1), 1e-3)
        a = max(ms[1] - b, 1e-3)
        self._STEP_COST = (a, b)
