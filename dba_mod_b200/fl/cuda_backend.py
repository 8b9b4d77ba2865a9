"""CUDA backend: the engine's compute interface on hand-written sm_100a kernels (csrc/, loaded through ctypes).

Fails loudly when the native library is missing - there is no silent fallback to library kernels on a GPU box.
What runs where:
  * flat-slab work (SGD + FoolsGold grad-sum, poison scaling, norms, FedAvg, Weiszfeld loop, FoolsGold weights and
    server step): csrc/flat_ops.cu
  * ResNet-18 (CIFAR / Tiny-ImageNet) evaluation and training: native executor over the tcgen05 implicit-GEMM
    convolution (csrc/conv_tc.cu) + fused BatchNorm / head kernels (``NativeResNet``), when available for the model
  * anything the native executor does not cover yet falls back to the batched torch formulation *explicitly*
    (``self.native_train`` / ``self.native_eval`` say which path is active; bench.py reports them).
"""
from __future__ import annotations

import logging
from typing import List, Optional

import torch

from .. import config as C
from ..ops import flat, native
from . import aggregation as agg
from .torch_backend import TorchBackend

logger = logging.getLogger("logger")


class CudaBackend(TorchBackend):
    name = "cuda"

    def __init__(self, params, spec, data, device) -> None:
        if device.type != "cuda":
            raise RuntimeError("CudaBackend needs a CUDA device")
        native.load()                       # raises NativeError when libdba_b200.so is absent
        super().__init__(params, spec, data, device)
        self.native_train = False
        self.native_eval = False
        self.exec = None
        if params["type"] in (C.TYPE_CIFAR, C.TYPE_TINYIMAGENET) and params.get("native_net", True):
            try:
                from .native_resnet import NativeResNet
                self.exec = NativeResNet(params, spec, data, device)
                self.native_eval = self.exec.can_eval
                self.native_train = self.exec.can_train
            except ImportError:
                self.exec = None
        if params["type"] == C.TYPE_MNIST and params.get("native_net", True):
            from .native_mnist import NativeMnist
            self.exec = NativeMnist(params, spec, data, device)
            self.native_eval, self.native_train = self.exec.can_eval, self.exec.can_train
        if params["type"] == C.TYPE_LOAN and params.get("native_net", True):
            from .native_loan import NativeLoan
            self.exec = NativeLoan(params, spec, data, device)
            self.native_eval = self.native_train = True

    # ---------------------------------------------------------------------------------------- flat ops
    def _sgd_update(self, W, V, g, lr_t, act, grad_sum, mu, wd) -> None:
        g = g.contiguous()
        flat.sgd_step_(W, V, g, lr_t.contiguous(), act.to(torch.uint8), self.spec.Dp, mu, wd, grad_sum)

    @staticmethod
    def scale_(W, anchor, scale) -> None:
        flat.scale_rows_(W, anchor, scale.float().contiguous())

    def dist_norm(self, W, anchor):
        return flat.rows_dist(W, anchor, n=self.spec.Dp).float()

    def global_norm(self, W):
        return flat.rows_dist(W, None, n=self.spec.Dp).float()

    # ------------------------------------------------------------------------------------- aggregation
    def fedavg_from_states_(self, G, G_cnt, W, cnt, eta, no_models, comm, sigma, seed) -> bool:
        """G += eta/no_models * sum_k (W[k] - G): delta extraction + reduce + apply (+ DP noise) in one launch.
        Returns False when the caller must use the collective path (several ranks)."""
        if comm.world != 1:
            return False
        flat.fedavg_update_(G, W, eta / no_models, sigma or 0.0, seed, bcast=False)
        if cnt is not None and cnt.numel() and G_cnt.numel():
            d = (cnt - G_cnt[None, :]).sum(0).double()
            G_cnt.add_((d * (eta / no_models)).trunc().long())
        return True

    def geometric_median_states(self, W, G, num_samples: List[float], maxiter: int):
        """Device-resident Weiszfeld loop over the deltas W[c] - G (single rank)."""
        ns = torch.tensor(num_samples, dtype=torch.float64, device=W.device)
        alphas = (ns / ns.sum()).float()
        med, dist, wv, obj, state = flat.weiszfeld(W, G, alphas, maxiter)
        st = state.cpu()
        return agg.RFAResult(int(st[2]), [float(v) for v in wv.cpu().tolist()],
                             [float(v) for v in dist.cpu().tolist()], med, float(obj.cpu()[0]), int(st[1]))

    def foolsgold_native(self, names_in_order, feats_local, counts, comm, use_memory: bool):
        """FoolsGold weights on the device (helper.py:527-607): per-client fp64 history, Gram, cosine, pardoning, logit.
        With several ranks the clients' feature rows stay where they were produced: every rank's kernel pulls the
        remote rows through peer-mapped pointers (NVLink loads inside the history/Gram kernels, no all-gather)."""
        from ..ops import flat
        dev = self.device
        L = feats_local.shape[1]
        n = len(names_in_order)
        if not hasattr(self, "_fg_hist"):
            cap = max(int(self.params["number_of_total_participants"]), n) + 64
            self._fg_hist = torch.zeros(cap, L, dtype=torch.float64, device=dev)
            self._fg_ids = {}
        for nm in names_in_order:
            if nm not in self._fg_ids:
                if len(self._fg_ids) >= self._fg_hist.shape[0]:
                    self._fg_hist = torch.cat([self._fg_hist, torch.zeros_like(self._fg_hist)], 0)
                self._fg_ids[nm] = len(self._fg_ids)
        ids = torch.tensor([self._fg_ids[nm] for nm in names_in_order], dtype=torch.int32, device=dev)
        if comm.world == 1:
            return flat.foolsgold_weights(self._fg_hist, ids, feats_local.float().contiguous(), use_memory)
        if not hasattr(comm, "feature_row_ptrs"):
            return None
        ptrs = comm.feature_row_ptrs(feats_local.float().contiguous(), counts)      # device int64 [n]
        work = torch.empty(n, L, dtype=torch.float64, device=dev)
        gram = torch.empty(n, n, dtype=torch.float64, device=dev)
        cs = torch.empty(n, n, dtype=torch.float64, device=dev)
        wv = torch.empty(n, dtype=torch.float64, device=dev)
        alpha = torch.empty(n, dtype=torch.float64, device=dev)
        lib = native.load()
        native.check(lib.dba_foolsgold_weights_ptrs(native.ptr(self._fg_hist), self._fg_hist.stride(0),
                                                    native.ptr(ids), native.ptr(ptrs), n, L, int(use_memory),
                                                    native.ptr(work), native.ptr(gram), native.ptr(cs),
                                                    native.ptr(wv), native.ptr(alpha), native.stream()))
        return wv, alpha

    def foolsgold_apply_native(self, G, grad_sum, wv_local, n_clients, eta, lr, decay, comm) -> bool:
        """G[:Dp] -= lr * (eta/n * sum_c wv_c * gradsum_c + decay * G[:Dp]); one fused launch (per rank)."""
        from ..ops import flat
        Dp = self.spec.Dp
        if comm.world == 1:
            S = grad_sum if grad_sum is not None else torch.zeros(0, Dp, device=self.device)
            flat.foolsgold_apply_(G, S.contiguous(), wv_local.double().contiguous(), Dp, eta / n_clients, lr, decay)
            return True
        if hasattr(comm, "weighted_reduce_"):
            # F3b: local weighted K-way reduce + cross-GPU reduce + server SGD step + broadcast in one launch per rank
            S = grad_sum if grad_sum is not None else torch.zeros(0, Dp, device=self.device)
            comm.weighted_reduce_(S, wv_local, Dp, -lr * eta / n_clients, 1.0 - lr * decay)
            return True
        return False

    # -------------------------------------------------------------------------------- train / evaluate
    def _empty_result(self, sch, want_grad_sum):
        from .torch_backend import TrainResult
        dev, E = self.device, max(sch.E_max, 1)
        z = lambda dt: torch.zeros(0, E, dtype=dt, device=dev)  # noqa: E731
        return TrainResult(torch.zeros(0, self.spec.Dp, device=dev) if want_grad_sum else None, z(torch.float32),
                           z(torch.long), z(torch.long), z(torch.long), torch.zeros(0, dtype=torch.bool, device=dev))

    def train(self, sch, W, cnt, anchor, want_grad_sum, generator=None):
        if W.shape[0] == 0:               # a rank without clients this round (more ranks than selected clients)
            return self._empty_result(sch, want_grad_sum)
        needs_dist = float(self.params["alpha_loss"]) != 1.0 and bool((sch.poison_n > 0).any())
        per_batch = bool(self.params["vis_train_batch_loss"] or self.params["batch_track_distance"])
        if self.native_train and not per_batch and (not needs_dist or getattr(self.exec, "supports_dist_loss", False)):
            # (the distance-regularised poison loss and the per-batch loss / distance traces run on the torch path)
            return self.exec.train(sch, W, cnt, anchor, want_grad_sum)
        return super().train(sch, W, cnt, anchor, want_grad_sum, generator)

    def evaluate(self, models, split, trig_row, shard=None):
        if self.native_eval:
            return self.exec.evaluate(models, split, trig_row, shard=shard)
        return super().evaluate(models, split, trig_row, shard=shard)

    def evaluate_pairs(self, models, jobs, shard=None):
        if self.native_eval and hasattr(self.exec, "evaluate_pairs"):
            return self.exec.evaluate_pairs(models, jobs, shard=shard)
        return None

    def evaluate_jobs(self, model, jobs, shard=None):
        if self.native_eval and hasattr(self.exec, "evaluate_jobs"):
            return self.exec.evaluate_jobs(model, jobs, shard=shard)
        return super().evaluate_jobs(model, jobs, shard=shard)

    @property
    def can_defer_eval(self) -> bool:
        """The executor can enqueue an evaluation without reading its result (NativeResNet.finalize_eval)."""
        return bool(self.native_eval and hasattr(self.exec, "finalize_eval"))

    def evaluate_jobs_deferred(self, model, jobs, shard=None):
        return self.exec.evaluate_jobs(model, jobs, shard=shard, defer=True)

    def finalize_eval(self, handle):
        return self.exec.finalize_eval(handle)
