"""Peer-memory communicator: the product path for the aggregation collectives on multi-GPU B200 boxes.

``torch.distributed._symmetric_memory`` is used ONLY as plumbing (allocate buffers that every process of the node can
map, exchange their addresses, one host-side rendezvous); the data movement itself happens inside the kernels of
csrc/p2p_collectives.cu, which load from / store to the peers' buffers over NVLink and synchronise with release /
acquire flags - no NCCL call on these paths.  ``TorchDistComm`` (NCCL) stays available as the measured baseline
(``fused_collectives: false``).

Buffers (per rank, all symmetric):  ``stage[D]`` partial sums, ``replica[D]`` the global model, ``signal`` tickets.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import native


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
        group = dist.group.WORLD
        self.stage = symm.empty(self.D, dtype=torch.float32, device=device)
        self.replica = symm.empty(self.D, dtype=torch.float32, device=device)
        self.aux = symm.empty(self.D, dtype=torch.float32, device=device)
        self.signal = symm.empty(64, dtype=torch.int32, device=device)
        self.stage.zero_(); self.replica.zero_(); self.aux.zero_(); self.signal.zero_()
        self._h_stage = symm.rendezvous(self.stage, group)
        self._h_replica = symm.rendezvous(self.replica, group)
        self._h_aux = symm.rendezvous(self.aux, group)
        self._h_signal = symm.rendezvous(self.signal, group)
        try:
            self._mc = {"stage": int(self._h_stage.multicast_ptr), "replica": int(self._h_replica.multicast_ptr),
                        "aux": int(self._h_aux.multicast_ptr)}
            self.has_multicast = all(v != 0 for v in self._mc.values())
        except Exception:  # noqa: BLE001
            self._mc, self.has_multicast = {}, False
        # NVLS: phase 2 of the fused kernels through the multicast mappings (reduction inside the NVSwitch).  "auto":
        # from 4 ranks on - measured at 2 ranks the two peer loads / stores per element are faster (77 vs 87 us for
        # 2.8 M floats), while the switch path moves world-times fewer bytes per GPU.
        if isinstance(use_nvls, str):
            use_nvls = {"auto": self.world >= 4, "true": True, "false": False}.get(use_nvls.lower(), self.world >= 4)
        self.use_nvls = bool(use_nvls) and self.has_multicast
        PtrArr = ctypes.c_void_p * self.world
        self._p_stage = PtrArr(*[int(p) for p in self._h_stage.buffer_ptrs])
        self._p_replica = PtrArr(*[int(p) for p in self._h_replica.buffer_ptrs])
        self._p_aux = PtrArr(*[int(p) for p in self._h_aux.buffer_ptrs])
        self._p_signal = PtrArr(*[int(p) for p in self._h_signal.buffer_ptrs])
        self.counter = torch.zeros(1, dtype=torch.int32, device=device)
        self.call_id = 0
        torch.cuda.synchronize(device)
        self.base.barrier()

    # -- plain collectives are delegated (metrics, ragged gathers of scalars) -----------------------------------
    def all_reduce_sum_(self, t):
        """Large fp32 vectors (a model-sized partial sum, e.g. each Weiszfeld iteration's weighted average) go through
        the fused peer-memory kernel; scalars and small / non-fp32 tensors use the base communicator."""
        if t.is_cuda and t.dtype == torch.float32 and 4096 <= t.numel() <= self.D and t.is_contiguous():
            out = self.allreduce_axpby_(t, 1.0, 0.0, target="aux")
            t.copy_(out.reshape(t.shape))
            return t
        return self.base.all_reduce_sum_(t)

    def all_gather_rows(self, t, counts):
        return self.base.all_gather_rows(t, counts)

    def broadcast_(self, t, src=0):
        return self.base.broadcast_(t, src)

    def barrier(self):
        return self.base.barrier()

    # -- fused paths --------------------------------------------------------------------------------------------
    def set_global(self, G: torch.Tensor) -> torch.Tensor:
        """Moves the global model into the symmetric replica; returns the replica view to use as ``G`` from now on."""
        self.replica[: G.numel()].copy_(G)
        return self.replica[: G.numel()]

    def fedavg_(self, W: torch.Tensor, coef: float, total_clients: int, sigma: float = 0.0, seed: int = 0) -> None:
        """replica += coef * sum over ALL ranks' client rows (W[k] - replica)  (+ noise), one launch per rank."""
        K = W.shape[0]
        ldw = W.stride(0) if K else self.D
        if self.use_nvls:
            native.check(self.lib.dba_fedavg_nvls(native.ptr(W) if K else None, ldw, K, self.D, coef, sigma, seed,
                                                  self._p_stage, self._p_replica, self._p_signal, self._mc["stage"],
                                                  self._mc["replica"], self.rank, self.world, total_clients,
                                                  native.ptr(self.counter), self.call_id, native.stream()))
        else:
            native.check(self.lib.dba_fedavg_p2p(native.ptr(W) if K else None, ldw, K, self.D, coef, sigma, seed,
                                                 self._p_stage, self._p_replica, self._p_signal, self.rank, self.world,
                                                 total_clients, native.ptr(self.counter), self.call_id,
                                                 native.stream()))
        self.call_id += 1

    def allreduce_axpby_(self, part: torch.Tensor, alpha: float, beta: float, n: Optional[int] = None,
                         target: str = "replica") -> torch.Tensor:
        """target[:n] = beta*target[:n] + alpha * sum_r part_r on every rank.  ``part`` is copied into the staging
        buffer; target is the global-model replica or the auxiliary vector."""
        n = n or part.numel()
        self.stage[:n].copy_(part.reshape(-1)[:n])
        tgt_ptrs, tgt = (self._p_replica, self.replica) if target == "replica" else (self._p_aux, self.aux)
        if self.use_nvls:
            native.check(self.lib.dba_allreduce_axpby_nvls(n, alpha, beta, self._p_stage, tgt_ptrs, self._p_signal,
                                                           self._mc["stage"], self._mc[target], self.rank, self.world,
                                                           native.ptr(self.counter), self.call_id, native.stream()))
        else:
            native.check(self.lib.dba_allreduce_axpby_p2p(n, alpha, beta, self._p_stage, tgt_ptrs, self._p_signal,
                                                          self.rank, self.world, native.ptr(self.counter),
                                                          self.call_id, native.stream()))
        self.call_id += 1
        return tgt[:n]

    def feature_row_ptrs(self, feats_local: torch.Tensor, counts) -> torch.Tensor:
        """Publishes this rank's FoolsGold feature rows in a symmetric buffer and returns, for the concatenated
        global row order, a device array of pointers to where each row lives (own memory or a peer's, mapped over
        NVLink).  Kernels that walk these pointers perform the gather themselves."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        n, L = sum(counts), feats_local.shape[1]
        need = n * L
        if getattr(self, "_feat", None) is None or self._feat.numel() < need:
            self._feat = symm.empty(max(need, 1), dtype=torch.float32, device=self.device)
            self._h_feat = symm.rendezvous(self._feat, dist.group.WORLD)
        lo = sum(counts[: self.rank])
        if feats_local.shape[0]:
            self._feat[lo * L:(lo + feats_local.shape[0]) * L].copy_(feats_local.reshape(-1))
        torch.cuda.synchronize(self.device)      # rows must be in memory before any peer starts reading them
        self.base.barrier()
        ptrs, row = [], 0
        for r, c in enumerate(counts):
            base = int(self._h_feat.buffer_ptrs[r])
            for _ in range(c):
                ptrs.append(base + row * L * 4)
                row += 1
        return torch.tensor(ptrs, dtype=torch.int64, device=self.device)
