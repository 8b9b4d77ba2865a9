"""Peer-memory communicator: the product path for the aggregation collectives on multi-GPU B200 boxes.

``torch.distributed._symmetric_memory`` is used ONLY as plumbing (allocate buffers that every process of the node can
map, exchange their addresses, one host-side rendezvous at start-up); the data movement itself happens inside the
kernels of csrc/p2p_collectives.cu, which load from / store to the peers' buffers over NVLink and synchronise with
release / acquire tickets - no NCCL call and no host synchronisation on these paths:

  fedavg_            F1: local K-way delta reduce -> cross-GPU reduce -> update (+DP noise) -> broadcast, one launch
  weiszfeld          F2: one launch per Weiszfeld iteration (local distances / weights / weighted partial sum, scalars
                     in the ticket mailbox, normalised median into every rank's buffer, device-side convergence test)
  weighted_reduce_   F3b: FoolsGold's weighted gradient reduce fused with the server SGD step
  feature_row_ptrs   F3a: feature rows stay where they were produced; the Gram kernel walks peer pointers
  gather_rows        evaluation exchange: peers' model rows are pulled over NVLink by a gather kernel
  device_barrier     stream-ordered cross-GPU barrier (replaces cudaDeviceSynchronize + NCCL barrier)

``TorchDistComm`` (NCCL) stays available as the measured baseline (``fused_collectives: false``); ``nccl_calls`` counts
what still goes through it after start-up (tests assert it stays at zero inside a round).

Buffers (per rank, all symmetric):  ``stage[D]`` partial sums, ``replica[D]`` the global model, ``aux[D]`` auxiliary
vector (Weiszfeld median / all-reduce results), ``signal`` tickets, ``mail`` scalar payloads, ``table`` per-client
results, ``rows`` exchange area for feature / model rows.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import native

TABLE_CAP = 2048         # clients per round the result tables can hold


class PeerComm:
    """Wraps a torch.distributed communicator and adds the fused peer-memory operations."""

    def __init__(self, base_comm, D: int, device: torch.device, use_nvls="auto") -> None:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        self.base = base_comm
        self.rank, self.world = base_comm.rank, base_comm.world
        self.D = int(D)
        self.device = device
        self.lib = native.load()
        self._symm, self._group = symm, dist.group.WORLD
        self.nccl_calls = 0
        self.stage = symm.empty(self.D, dtype=torch.float32, device=device)
        self.replica = symm.empty(self.D, dtype=torch.float32, device=device)
        self.aux = symm.empty(self.D, dtype=torch.float32, device=device)
        self.signal = symm.empty(64, dtype=torch.int32, device=device)
        self.mail = symm.empty(128, dtype=torch.float64, device=device)
        self.table = symm.empty(2 * TABLE_CAP, dtype=torch.float32, device=device)
        for t in (self.stage, self.replica, self.aux, self.signal, self.mail, self.table):
            t.zero_()
        hs = {n: symm.rendezvous(getattr(self, n), self._group) for n in ("stage", "replica", "aux", "signal", "mail",
                                                                         "table")}
        self._handles = hs
        try:
            self._mc = {n: int(hs[n].multicast_ptr) for n in ("stage", "replica", "aux")}
            self.has_multicast = all(v != 0 for v in self._mc.values())
        except Exception:  # noqa: BLE001
            self._mc, self.has_multicast = {"stage": 0, "replica": 0, "aux": 0}, False
        # NVLS: owner phase of the fused kernels through the multicast mappings (reduction inside the NVSwitch).
        # "auto": from 4 ranks on - measured at 2 ranks the two peer loads / stores per element are faster, while the
        # switch path moves world-times fewer bytes per GPU.
        if isinstance(use_nvls, str):
            use_nvls = {"auto": self.world >= 4, "true": True, "false": False}.get(use_nvls.lower(), self.world >= 4)
        self.use_nvls = bool(use_nvls) and self.has_multicast
        PtrArr = ctypes.c_void_p * self.world

        def arr(name):
            return PtrArr(*[int(p) for p in hs[name].buffer_ptrs])

        self._counters = torch.zeros(8, dtype=torch.int32, device=device)
        self._err = torch.zeros(1, dtype=torch.int32, device=device)
        self._table_dev = torch.zeros(8, dtype=torch.int64, device=device)
        native.check(self.lib.dba_peer_init(
            self.rank, self.world, arr("stage"), arr("replica"), arr("aux"), arr("signal"), arr("mail"), arr("table"),
            TABLE_CAP, self._mc["stage"] if self.has_multicast else None,
            self._mc["replica"] if self.has_multicast else None, self._mc["aux"] if self.has_multicast else None,
            native.ptr(self._counters), native.ptr(self._err), native.ptr(self._table_dev)))
        self._rows = None            # symmetric exchange area for feature / model rows (grown on demand)
        self._rows_ptrs: List[int] = []
        torch.cuda.synchronize(device)
        self.base.barrier()

    # -- plain collectives are delegated (start-up votes, CSV row gathers) ---------------------------------------
    def all_reduce_sum_(self, t):
        """Small tensors (metric sums): one device-side all-reduce through the staging buffer when they are fp32 / fp64
        CUDA tensors that fit, else the base communicator."""
        if t.is_cuda and t.numel() <= self.D and t.dtype in (torch.float32, torch.float64):
            out = self.allreduce_axpby_(t.reshape(-1).float(), 1.0, 0.0, target="aux")
            t.copy_(out.reshape(t.shape).to(t.dtype))
            return t
        self.nccl_calls += 1
        return self.base.all_reduce_sum_(t)

    def all_gather_rows(self, t, counts):
        self.nccl_calls += 1
        return self.base.all_gather_rows(t, counts)

    def broadcast_(self, t, src=0):
        self.nccl_calls += 1
        return self.base.broadcast_(t, src)

    def barrier(self):
        self.nccl_calls += 1
        return self.base.barrier()

    def check(self) -> None:
        """Raises when a cross-GPU barrier timed out on this GPU (a peer died or never launched its kernel)."""
        if int(self.lib.dba_peer_error()) != 0:
            raise RuntimeError(f"rank {self.rank}: a peer-memory barrier timed out (a peer rank failed?)")

    # -- fused paths --------------------------------------------------------------------------------------------
    def set_global(self, G: torch.Tensor) -> torch.Tensor:
        """Moves the global model into the symmetric replica; returns the replica view to use as ``G`` from now on."""
        self.replica[: G.numel()].copy_(G)
        return self.replica[: G.numel()]

    def device_barrier(self) -> None:
        native.check(self.lib.dba_peer_barrier(native.stream()))

    def fedavg_(self, W: torch.Tensor, coef: float, total_clients: int = 0, sigma: float = 0.0, seed: int = 0) -> None:
        """replica += coef * sum over ALL ranks' client rows (W[k] - replica)  (+ noise), one launch per rank."""
        K = W.shape[0]
        ldw = W.stride(0) if K else self.D
        native.check(self.lib.dba_peer_fedavg(native.ptr(W) if K else None, ldw, K, self.D, coef, sigma, seed,
                                              int(self.use_nvls), native.stream()))

    def allreduce_axpby_(self, part: torch.Tensor, alpha: float, beta: float, n: Optional[int] = None,
                         target: str = "replica") -> torch.Tensor:
        """target[:n] = beta*target[:n] + alpha * sum_r part_r on every rank (``part`` goes through the staging
        buffer); target is the global-model replica or the auxiliary vector."""
        n = n or part.numel()
        self.stage[:n].copy_(part.reshape(-1)[:n])
        native.check(self.lib.dba_peer_axpby(None, 0, 0, None, n, alpha, beta, 0 if target == "replica" else 1,
                                             int(self.use_nvls), native.stream()))
        return (self.replica if target == "replica" else self.aux)[:n]

    def weighted_reduce_(self, S: torch.Tensor, wv: torch.Tensor, n: int, alpha: float, beta: float) -> None:
        """replica[:n] = beta*replica[:n] + alpha * sum over ALL ranks' rows of wv[k] * S[k, :n]: the local weighted
        K-way reduce is phase 1 of the same launch (FoolsGold's aggregated gradient + server SGD step)."""
        K = S.shape[0]
        wv64 = wv.double().contiguous() if K else None
        if K == 0:          # a rank without clients this round contributes a zero vector (phase 1 is skipped)
            self.stage[:n].zero_()
        native.check(self.lib.dba_peer_axpby(native.ptr(S) if K else None, S.stride(0) if K else 0, K,
                                             native.ptr(wv64), n, alpha, beta, 0, int(self.use_nvls), native.stream()))
        self._keep_wv = wv64

    def weiszfeld(self, W: torch.Tensor, G: torch.Tensor, alphas_local: torch.Tensor, slot0: int, n_total: int,
                  maxiter: int, eps: float = 1e-5, ftol: float = 1e-6):
        """RFA over the deltas (W[k] - G) of all ranks: maxiter + 2 launches, no host sync, no NCCL.  Returns
        (median [D] view of the auxiliary vector, table [2, n_total] = (wv, distance) of every client, state, obj)."""
        K = W.shape[0]
        dev = self.device
        kmax = max(K, 1)
        dist2 = torch.empty(2 * kmax, dtype=torch.float64, device=dev)
        wts = torch.zeros(kmax, dtype=torch.float32, device=dev)
        wv = torch.zeros(kmax, dtype=torch.float32, device=dev)
        obj = torch.empty(2, dtype=torch.float64, device=dev)
        state = torch.empty(4, dtype=torch.int32, device=dev)
        a = alphas_local.float().contiguous()
        native.check(self.lib.dba_peer_weiszfeld(native.ptr(W) if K else None, W.stride(0) if K else self.D, K, self.D,
                                                 native.ptr(G), native.ptr(a) if K else None, int(maxiter), eps, ftol,
                                                 int(slot0), int(n_total), native.ptr(dist2), kmax, native.ptr(wts),
                                                 native.ptr(wv), native.ptr(obj), native.ptr(state), native.stream()))
        self._keep = (dist2, wts, wv, a)       # alive until the stream has run the launches
        return self.aux[: self.D], self.table[: 2 * n_total].view(2, n_total), state, obj

    # -- row exchange over peer pointers ------------------------------------------------------------------------
    def _rows_area(self, need: int) -> None:
        """(Re)allocates the symmetric row-exchange area.  All ranks take the same decision (``need`` is computed from
        global quantities), so the rendezvous inside is collective."""
        if self._rows is None or self._rows.numel() < need:
            cap = max(int(need), 1)
            self._rows = self._symm.empty(cap, dtype=torch.float32, device=self.device)
            h = self._symm.rendezvous(self._rows, self._group)
            self._rows_ptrs = [int(p) for p in h.buffer_ptrs]
            self._h_rows = h

    def reserve_rows(self, n_floats: int) -> None:
        """Sizes the exchange area up front (collective) so that no round pays for a rendezvous."""
        self._rows_area(int(n_floats))

    def publish_rows(self, rows_local: torch.Tensor, counts: List[int], L: int) -> torch.Tensor:
        """Copies this rank's rows [counts[rank], L] into its exchange area at the offset of its first global row and
        returns, for the rank-concatenated row order, a device array of pointers to where each row lives (own memory or
        a peer's, mapped over NVLink).  A device-side barrier orders the peers' copies before anybody reads."""
        n = sum(counts)
        self._rows_area(n * L)
        lo = sum(counts[: self.rank])
        if rows_local.shape[0]:
            self._rows[lo * L:(lo + rows_local.shape[0]) * L].copy_(rows_local.reshape(-1))
        self.device_barrier()
        ptrs, row = [], 0
        for r, c in enumerate(counts):
            for _ in range(c):
                ptrs.append(self._rows_ptrs[r] + row * L * 4)
                row += 1
        return torch.tensor(ptrs, dtype=torch.int64, device=self.device)

    def feature_row_ptrs(self, feats_local: torch.Tensor, counts) -> torch.Tensor:
        """FoolsGold: publishes this rank's feature rows; kernels that walk the returned pointers perform the gather
        themselves (csrc/flat_ops.cu fg_history_ptrs_kernel)."""
        return self.publish_rows(feats_local.float().contiguous(), list(counts), feats_local.shape[1])

    def gather_rows(self, rows_local: torch.Tensor, counts: List[int]) -> torch.Tensor:
        """All ranks' rows [sum(counts), L] in this GPU's memory, pulled over NVLink by a gather kernel (no NCCL)."""
        L = rows_local.shape[1]
        n = sum(counts)
        ptrs = self.publish_rows(rows_local.contiguous(), list(counts), L)
        out = torch.empty(n, L, dtype=torch.float32, device=self.device)
        if n:
            native.check(self.lib.dba_gather_rows_ptrs(native.ptr(ptrs), n, L, native.ptr(out), out.stride(0),
                                                       native.stream()))
        self._keep_ptrs = ptrs
        # the exchange area may be rewritten by its owner only after every reader is done
        self.device_barrier()
        return out
