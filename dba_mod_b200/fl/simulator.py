"""The experiment driver: round loop, client training, aggregation, evaluation, recording, checkpoints.

Public API (what ``main.py --params X.yaml`` and ``bench.py`` call):

    sim = Simulator(Params.from_yaml(path))      # loads data, builds / resumes the global model
    sim.run()                                     # all rounds  (reference main.py:135-235)
    sim.run_round(epoch)                          # one round, returns a RoundReport

Reference flow reproduced per round (main.py:135-235, image_train.py:12-315, loan_train.py:11-261, test.py):
select clients -> local training of every selected client from the current global model (benign or poisoned, with
scaling) -> per-client evaluations (Q16) -> aggregation (mean / geom_median / foolsgold) -> global evaluations
(clean, combined trigger, per-trigger) -> checkpoint -> CSV dump.  What differs is *how*: the clients of a rank
train as one batched group over flat slabs, all evaluations of a round that share inputs are batched over models,
metrics stay on the device until the end of the phase, and with several ranks (one process per GPU) the clients are
sharded over ranks and aggregation runs as collectives (fused peer-memory kernels on B200, ``torch.distributed`` as
the baseline / CPU path).
"""
from __future__ import annotations

import logging
import os
import random
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import config as C
from ..config import Params
from ..data import ImageData, LoanData
from ..models import display_name, initial_state
from ..utils import checkpoint as ckpt
from ..utils.csv_record import Recorder
from ..utils.logging_utils import close_logger, make_run_folder, now_tag, open_logger
from ..utils.metrics import MetricsStream, dict_html
from ..utils.timing import PhaseTimer
from . import aggregation as agg
from .schedule import ClientPlan, build_schedule, plan_clients
from .selection import select_clients

logger = logging.getLogger("logger")


@dataclass
class RoundReport:
    epoch: int
    names: List
    adversaries: List
    global_acc: float = 0.0
    global_loss: float = 0.0
    poison_acc: Optional[float] = None
    poison_loss: Optional[float] = None
    trigger_acc: Dict[str, float] = field(default_factory=dict)
    agg_weights: Optional[List[float]] = None
    agg_alphas: Optional[List[float]] = None
    steps: int = 0
    nan_clients: List = field(default_factory=list)
    seconds: float = 0.0


def pick_device(params) -> torch.device:
    if params["backend"] == "torch" and not torch.cuda.is_available():
        return torch.device("cpu")
    if torch.cuda.is_available():
        return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    return torch.device("cpu")


class Simulator:
    def __init__(self, params: Params, device: Optional[torch.device] = None, comm=None, out_root: str = "saved_models",
                 quiet: bool = False, data=None, write_files: bool = True) -> None:
        self.params = params
        self.device = device if device is not None else pick_device(params)
        self.comm = comm if comm is not None else agg.LocalComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.is_main = self.rank == 0
        self.write_files = write_files and self.is_main
        # per-client CSV rows of the other ranks are only gathered when somebody writes them (same decision on all ranks)
        self.gather_rows = bool(write_files)
        seed = int(params["seed"])
        # main.py:36-38,86 — one python / numpy stream drives partition shuffles, Dirichlet draws and selection
        self.py_rng = random.Random(seed)
        self.np_rng = np.random.RandomState(seed)
        self.torch_gen = torch.Generator(device=self.device).manual_seed(seed + 7919 * self.rank)
        self.noise_gen = torch.Generator(device=self.device).manual_seed(seed + 104729)
        self.sched_gen = torch.Generator(device=self.device)

        name = params.get("name") or C.DEFAULT_NAME[params["type"]]
        self.name = name
        self.current_time = now_tag()
        if self.write_files:
            self.folder_path = make_run_folder(name, self.current_time, out_root)
            self._log_handlers = open_logger(self.folder_path, stream=not quiet)
        else:
            self.folder_path = None
            self._log_handlers = []
            logging.getLogger("logger").setLevel(logging.WARNING if quiet else logging.INFO)
        if not params.get("environment_name"):
            params["environment_name"] = name
        params["current_time"] = self.current_time
        params["folder_path"] = self.folder_path
        logger.info(f"current path: {self.folder_path}")

        self.recorder = Recorder()
        self.metrics = MetricsStream(self.folder_path, enabled=self.write_files)
        self.ckpt_writer = ckpt.CheckpointWriter(params)
        self.timer = PhaseTimer(self.device)

        # ---- data (image_helper.load_data / loan_helper.load_data)
        if data is not None:
            self.data = data
        elif params.is_image:
            self.data = ImageData(params, self.device, self.py_rng, self.np_rng)
        else:
            self.data = LoanData(params, self.device)
        logger.info("load data done")

        # ---- model (create_model + resume; image_helper.py:32-70, loan_helper.py:85-109)
        self.spec, flat, cnt = initial_state(params["type"], seed)
        self.start_epoch = 1
        if params["resumed_model"]:
            path = ckpt.resume_path(params, out_root if os.path.isdir(out_root) else "saved_models")
            if os.path.exists(path) or params.get("resume_required", False):
                sd, ep, lr = ckpt.load_checkpoint(path)
                flat, cnt = self.spec.flatten(sd)
                self.start_epoch = ep + 1
                if lr is not None:
                    params["lr"] = lr
                logger.info(f"Loaded parameters from saved model: LR is {params['lr']} and current epoch is "
                            f"{self.start_epoch}")
            else:
                # the shipped YAMLs resume from a pre-trained checkpoint that is not part of any repository
                # (the reference dies in torch.load here); start from the random initialisation instead
                logger.warning(f"resumed_model is set but {path} does not exist: starting from random "
                               f"initialisation at epoch 1 (set resume_required: true to make this an error)")
        self.G = flat.to(self.device)
        self.G_cnt = cnt[: max(self.spec.n_counters, 0)].to(self.device) if self.spec.n_counters else \
            torch.zeros(0, dtype=torch.long, device=self.device)
        logger.info("create model done")
        self.local_name = display_name(params["type"], "Local")
        self.target_name = display_name(params["type"], "Target")

        self.backend = self._make_backend()
        ex = getattr(self.backend, "exec", None)
        if (ex is not None and hasattr(ex, "prewarm") and getattr(self.backend, "native_train", False)
                and params.get("compact_tail", True) and params.get("prewarm", True)):
            # every batch width a round can shrink to gets its slab and captured step graph now, not mid-round
            # the cost-model placement can give one rank anything between 1 and no_models - (world - 1) clients
            kmax = max(1, int(params["no_models"]) - (self.world - 1))
            ex.prewarm(kmax, params["aggregation_methods"] == C.AGGR_FOOLSGOLD)
            if self.world > 1:
                # the measured cost model drives the client placement, which every rank must compute identically:
                # everybody adopts rank 0's measurement (start-up plumbing, before the first round)
                c = torch.tensor(ex._STEP_COST, dtype=torch.float64, device=self.device)
                self.comm.broadcast_(c, 0)
                ex._STEP_COST = (float(c[0]), float(c[1]))
        self.collective_path = "single process" if self.world == 1 else "torch.distributed baseline"
        if (self.world > 1 and self.device.type == "cuda" and params["fused_collectives"]
                and not hasattr(self.comm, "fedavg_")):
            # all ranks must agree on the path: try to set the peer-memory communicator up, then vote
            peer, err = None, None
            try:
                from ..parallel.symm import PeerComm
                peer = PeerComm(self.comm, self.spec.Df, self.device, use_nvls=params.get("nvls", "auto"))
            except Exception as e:  # noqa: BLE001
                err = e
            ok = torch.tensor([1.0 if peer is not None else 0.0], device=self.device)
            self.comm.all_reduce_sum_(ok)
            if int(ok.item()) == self.world:
                self.comm = peer
                self.G = self.comm.set_global(self.G)
                # evaluation exchange area: every client's model plus the attackers' extra tests of a round
                n_adv = len(params["adversary_list"]) if params["is_poison"] else 0
                peer.reserve_rows((int(params["no_models"]) + 4 * n_adv + 4) * self.spec.Df)
                self.collective_path = ("fused peer-memory kernels (NVLS multimem)" if getattr(peer, "use_nvls", False)
                                        else "fused peer-memory kernels (NVLink P2P)")
            else:
                logger.warning(f"peer-memory communicator unavailable ({err}); using the NCCL baseline collectives")
                self.collective_path = "nccl baseline"
        self.fg = agg.FoolsGoldState(use_memory=bool(params["fg_use_memory"]))
        self._loan_local_epoch = self.start_epoch - 1
        self.stream_inputs = False          # bench end-to-end mode: re-upload the round's inputs from pinned host
        self._pinned = None
        self._test_rows = None
        self._host_idx = {}
        # overlap_global_eval: a round's global evaluations (main.py:195-231) are only ENQUEUED - on a side stream - and
        # read back after the next round's local training, which they overlap; records / CSV rows keep the reference's
        # order, log lines of the global tests appear one training phase late.  Pays when the training phase leaves SMs
        # idle (one client per GPU: launch-bound K = 1 steps).
        self.overlap_global_eval = (bool(self.params.get("overlap_global_eval", False)) and self.device.type == "cuda"
                                    and bool(getattr(self.backend, "can_defer_eval", False)))
        self._eval_stream = torch.cuda.Stream(device=self.device) if self.overlap_global_eval else None
        self._pending_eval = None
        self._pending_client = None         # deferred per-client evaluations of the same round (image data only)
        self._pending_rows = None
        self.h2d_bytes_last = 0
        self.d2h_bytes_last = 0

        if params["is_poison"]:
            logger.info(f"Poisoned following participants: {params['adversary_list']}")
        if self.write_files:
            params.dump_yaml(os.path.join(self.folder_path, "params.yaml"))
            with open(os.path.join(self.folder_path, "params.html"), "w") as f:
                f.write(dict_html(params, self.current_time))
        unknown = params.unknown_keys()
        if unknown:
            logger.info(f"ignoring unknown parameters: {unknown}")

    # ------------------------------------------------------------------------------------------ set-up
    def _make_backend(self):
        want = self.params["backend"]
        if want == "auto":
            want = "cuda" if self.device.type == "cuda" else "torch"
        if want == "cuda":
            from .cuda_backend import CudaBackend
            return CudaBackend(self.params, self.spec, self.data, self.device)
        from .torch_backend import TorchBackend
        return TorchBackend(self.params, self.spec, self.data, self.device)

    def state_dict(self):
        return self.spec.unflatten(self.G.detach().cpu(), self.G_cnt.detach().cpu() if self.G_cnt.numel() else None)

    def load_state_dict(self, sd) -> None:
        flat, cnt = self.spec.flatten(sd)
        self.G.copy_(flat.to(self.device))
        if self.spec.n_counters:
            self.G_cnt.copy_(cnt[: self.spec.n_counters].to(self.device))

    def close(self) -> None:
        self.flush_pending()
        self.metrics.close()
        close_logger(self._log_handlers)
        self._log_handlers = []

    # ------------------------------------------------------------------------------------- client sharding
    def _placement(self, names: List, epoch: Optional[int] = None) -> List[int]:
        """owner[slot] = rank.  Greedy makespan placement under the lock-step cost model: a rank's local phase costs
        ``a * T + b * sum(steps)`` (T = its longest client; every step pays the launch floor ``a`` once, however many
        clients share it), so a client goes to the rank whose cost grows least - long clients end up sharing their
        floor with short ones instead of being balanced by step count alone."""
        p = self.params
        if epoch is not None:
            plans = plan_clients(p, self.data, names, epoch)
            steps = [pl.n_epochs * pl.nb for pl in plans]
        else:
            steps = []
            for nm in names:
                adv = p.adversary_index(nm) if (p["is_poison"] and nm in p["adversary_list"]) else -1
                ne = p["internal_poison_epochs"] if adv >= 0 else p["internal_epochs"]
                steps.append(self.data.batches_per_epoch(nm) * ne)
        a, b = getattr(getattr(self.backend, "exec", None), "_STEP_COST", (1.0, 0.3))
        T = [0] * self.world
        S = [0] * self.world
        owner = [0] * len(names)
        for st, slot in sorted(((st, s) for s, st in enumerate(steps)), key=lambda t: (-t[0], t[1])):
            r = min(range(self.world), key=lambda i: (a * max(T[i], st) + b * (S[i] + st), i))
            T[r], S[r] = max(T[r], st), S[r] + st
            owner[slot] = r
        return owner

    def _shard(self, names: List, epoch: Optional[int] = None) -> List[int]:
        """Slots of this rank."""
        if self.world == 1:
            return list(range(len(names)))
        owner = self._placement(names, epoch)
        return [s for s in range(len(names)) if owner[s] == self.rank]

    def _owner_counts(self, names: List, epoch: Optional[int] = None) -> Tuple[List[int], List[int]]:
        """(rows per rank, global-slot order of the concatenated rank shards)."""
        if self.world == 1:
            return [len(names)], list(range(len(names)))
        owner = self._placement(names, epoch)
        counts, order = [], []
        for r in range(self.world):
            sl = [s for s in range(len(names)) if owner[s] == r]
            counts.append(len(sl))
            order += sl
        return counts, order

    # ------------------------------------------------------------------------------------------ the loop
    def run(self, max_rounds: Optional[int] = None) -> List[RoundReport]:
        p = self.params
        reports = []
        for i, epoch in enumerate(range(self.start_epoch, p["epochs"] + 1, p["aggr_epoch_interval"])):
            if max_rounds is not None and i >= max_rounds:
                break
            reports.append(self.run_round(epoch))
        self.flush_pending()
        logger.info("Saving all the graphs.")
        logger.info(f"This run has a label: {p['current_time']}. Metrics stream: {self.metrics.path}")
        return reports

    def prepare_streaming(self) -> None:
        """Pins the host copies of the data set once (end-to-end mode); afterwards every round only gathers and
        uploads that round's inputs."""
        d = self.data
        if self._pinned is None:
            pin = (lambda t: t.pin_memory()) if self.device.type == "cuda" else (lambda t: t)
            self._pinned = {"train": pin(d.train_x_host), "test": pin(d.test_x_host)}
        self.stream_inputs = True

    def _stage_inputs(self, names: List) -> None:
        """End-to-end mode: the round's inputs (selected clients' training samples + the test split) come from
        pinned host memory every round, like the reference's DataLoader -> .to(device) path."""
        d = self.data
        if self._pinned is None:
            pin = (lambda t: t.pin_memory()) if self.device.type == "cuda" else (lambda t: t)
            self._pinned = {"train": pin(d.train_x_host), "test": pin(d.test_x_host)}
        idx = torch.cat([d.client_index_tensor(n) for n in names]) if names else None
        nbytes = 0
        if idx is not None and idx.numel():
            # host copies of the clients' index lists are kept (a .cpu() per round would stall the host on the stream)
            cache = self._host_idx
            for n in names:
                if n not in cache:
                    cache[n] = d.client_index_tensor(n).cpu()
            host_idx = torch.cat([cache[n] for n in names])
            n = int(host_idx.numel())
            stage = self._pinned.get("stage")
            if stage is None or stage.shape[0] < n:
                stage = torch.empty((max(n, 8192),) + tuple(self._pinned["train"].shape[1:]),
                                    dtype=self._pinned["train"].dtype)
                if self.device.type == "cuda":
                    stage = stage.pin_memory()
                self._pinned["stage"] = stage
            rows = stage[:n]
            torch.index_select(self._pinned["train"], 0, host_idx, out=rows)      # gather into pinned memory
            dev_rows = rows.to(self.device, non_blocking=True)
            d.train_x.index_copy_(0, idx, dev_rows)
            nbytes += rows.numel() * rows.element_size()
        self._h2d_train = nbytes

    def _stage_test_inputs(self) -> None:
        """End-to-end mode, second half: the test split, uploaded right before the round's first evaluation (after a
        deferred global evaluation of the previous round has stopped reading it)."""
        d = self.data
        if self.world > 1 and self.params["type"] in C.IMAGE_TYPES:
            # several ranks: a rank evaluates its slice of the clean and of the backdoor test set (the executors' _shard:
            # contiguous slices of ceil(n / world) indices) - those rows are this rank's inputs
            if self._test_rows is None:
                def part(ids):
                    n = int(ids.numel())
                    per = (n + self.world - 1) // self.world
                    return ids[self.rank * per:min(n, (self.rank + 1) * per)]
                rows = torch.unique(torch.cat([part(d.all_test_idx), part(d.poison_test_idx)]))
                self._test_rows = (rows, rows.cpu())
                stage = torch.empty((int(rows.numel()),) + tuple(self._pinned["test"].shape[1:]),
                                    dtype=self._pinned["test"].dtype)
                self._pinned["test_stage"] = stage.pin_memory() if self.device.type == "cuda" else stage
                # the test split never changes: its shard is gathered into the pinned staging buffer once, every round
                # then pays the host->device copy only
                torch.index_select(self._pinned["test"], 0, self._test_rows[1], out=self._pinned["test_stage"])
            rows_dev, rows_host = self._test_rows
            stage = self._pinned["test_stage"]
            d.test_x.index_copy_(0, rows_dev, stage.to(self.device, non_blocking=True))
            self.h2d_bytes_last = self._h2d_train + stage.numel() * stage.element_size()
            return
        d.test_x.copy_(self._pinned["test"], non_blocking=True)
        self.h2d_bytes_last = self._h2d_train + d.test_x.numel() * d.test_x.element_size()

    def run_round(self, epoch: int) -> RoundReport:
        p, be, dev = self.params, self.backend, self.device
        t0 = time.time()
        names, adversaries = select_clients(p, epoch, self.data.participants_list, self.data.benign_namelist,
                                            self.py_rng)
        logger.info(f"Server Epoch:{epoch} choose agents : {names}.")
        rep = RoundReport(epoch, list(names), list(adversaries))
        self._round_epoch, self._loan_tle, self._round_names = epoch, {}, list(names)
        my_slots = self._shard(names, epoch)
        my_names = [names[s] for s in my_slots]
        counts, order = self._owner_counts(names, epoch)
        if self.stream_inputs:
            self._stage_inputs(my_names)
        K, Df = len(my_names), self.spec.Df
        use_fg = p["aggregation_methods"] == C.AGGR_FOOLSGOLD

        # R1: global -> client broadcast (image_train.py:26-32)
        ld = (Df + 31) // 32 * 32               # 128-byte aligned rows: the flat kernels vectorise over them
        W = torch.empty(K, ld, device=dev)[:, :Df]
        W.copy_(self.G.unsqueeze(0).expand(K, Df))
        cnt = self.G_cnt.unsqueeze(0).repeat(K, 1) if self.spec.n_counters else torch.zeros(K, 0, dtype=torch.long,
                                                                                             device=dev)
        grad_sum = None
        num_samples = [0] * K
        rows_local: List[Tuple[int, str, list]] = []     # (slot, table, row) emitted in reference order later

        with self.timer.phase("train"):
            for seg_epoch in range(epoch, epoch + p["aggr_epoch_interval"]):
                anchor, anchor_cnt = W.clone(), cnt.clone()
                lr_override = self._loan_adaptive_lr(my_names, seg_epoch, W)
                with self.timer.phase("train/schedule"):
                    plans = plan_clients(p, self.data, my_names, seg_epoch, lr_override)
                    sch = build_schedule(p, self.data, plans, seg_epoch, dev, self.sched_gen)
                with self.timer.phase("train/local_steps"):
                    res = be.train(sch, W, cnt, anchor, use_fg, generator=self.torch_gen)
                self.flush_pending()            # previous round's global evaluations (overlapped the training above)
                if self.stream_inputs and seg_epoch == epoch:
                    self._stage_test_inputs()
                rep.steps += sch.T
                if use_fg:
                    grad_sum = res.grad_sum     # "only works for aggr_epoch_interval=1" (image_train.py:24)
                with self.timer.phase("train/client_eval"):
                    self._after_training(seg_epoch, plans, my_slots, sch, res, W, cnt, anchor, anchor_cnt,
                                         rows_local, num_samples, rep)
        # emit per-client rows in the reference's order (client by client)
        if self._pending_client is not None:
            self._pending_rows = (rows_local, counts)          # completed and emitted by flush_pending()
        else:
            self._emit_rows(rows_local, counts)

        # ---------------- aggregation (main.py:172-191)
        with self.timer.phase("aggregate"):
            self._aggregate(epoch, names, adversaries, my_slots, counts, order, W, cnt, grad_sum, num_samples, rep)

        # ---------------- global evaluation (main.py:195-231)
        with self.timer.phase("global_eval"):
            if self.overlap_global_eval:
                self._enqueue_client_evals()
                self._launch_global_eval(epoch, rep, t0)
                return rep                # completed (in place) by flush_pending() during the next round / at the end
            self._global_eval(epoch, rep)
        self._end_of_round(epoch, rep, t0)
        return rep

    def _end_of_round(self, epoch: int, rep: RoundReport, t0: float) -> None:
        p = self.params
        if hasattr(self.comm, "check"):
            self.comm.check()            # a timed-out cross-GPU barrier (dead peer) surfaces here, once per round
        if self.write_files:
            self.ckpt_writer.save_model(self.state_dict(), epoch, rep.poison_loss if p["is_poison"] else rep.global_loss)
            self.recorder.save_result_csv(epoch, p["is_poison"], self.folder_path)
            self.metrics.flush()
        else:
            self.recorder.flush_scale_row()
        rep.seconds = time.time() - t0
        logger.info(f"Done in {rep.seconds} sec.")

    def _launch_global_eval(self, epoch: int, rep: RoundReport, t0: float) -> None:
        """Enqueue the round's global evaluations on the side stream (they read self.G and the test split, both left
        untouched until flush_pending())."""
        jobs, tests = self._global_jobs()
        shard = (self.rank, self.world) if self.world > 1 else None
        side = self._eval_stream
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            handle = self.backend.evaluate_jobs_deferred(self.G, jobs, shard=shard)
        self._pending_eval = (epoch, rep, t0, tests, handle)

    def flush_pending(self) -> None:
        """Read back and record a deferred global evaluation (no-op when none is pending)."""
        pend, self._pending_eval = self._pending_eval, None
        if pend is None:
            return
        epoch, rep, t0, tests, handle = pend
        torch.cuda.current_stream().wait_stream(self._eval_stream)     # evaluation buffers are shared with the main stream
        pc, self._pending_client = self._pending_client, None
        if pc is not None:              # the round's per-client tests first: their rows precede the global rows
            c_epoch, plans, my_slots, rows, chandle = pc
            self._record_client_evals(c_epoch, plans, my_slots, rows, self._finish_eval_jobs(chandle))
            (rows_local, counts), self._pending_rows = self._pending_rows, None
            self._emit_rows(rows_local, counts)
        res = self.backend.finalize_eval(handle)
        self._record_global_eval(epoch, rep, tests, self._reduce_eval_rows(res))
        self._end_of_round(epoch, rep, t0)

    # -------------------------------------------------------------------------------------- local phase
    def _loan_adaptive_lr(self, my_names, seg_epoch, W) -> Optional[dict]:
        """Q18 (loan_train.py:65-75): a LOAN attacker first measures the backdoor accuracy of the model it starts
        from and shrinks ``poison_lr`` (/5 above 20 %, a further /10 above 60 %) unless ``baseline``."""
        p = self.params
        if p["type"] != C.TYPE_LOAN or not p["is_poison"]:
            return None
        ks = [k for k, nm in enumerate(my_names)
              if nm in p["adversary_list"] and seg_epoch in p.poison_epochs_of(p.adversary_index(nm))]
        if not ks:
            return None
        loss, correct, n = self.backend.evaluate(W[ks], "poison", 0)
        out = {}
        for j, k in enumerate(ks):
            acc_p = 100.0 * float(correct[j]) / n if n else 0.0
            logger.info(acc_p)
            lr = float(p["poison_lr"])
            if not p["baseline"]:
                if acc_p > 20:
                    lr /= 5
                if acc_p > 60:
                    lr /= 10
            out[my_names[k]] = lr
        return out

    def _after_training(self, epoch, plans: List[ClientPlan], my_slots, sch, res, W, cnt, anchor, anchor_cnt,
                        rows, num_samples, rep) -> None:
        p, be = self.params, self.backend
        is_loan = p["type"] == C.TYPE_LOAN
        K = len(plans)
        # ---- one D2H of the phase's metric block
        loss_sum = res.loss_sum.cpu()
        correct = res.correct.cpu()
        count = res.count.cpu()
        pcount = res.poison_count.cpu()
        nan = res.nan_flag.cpu()
        self.d2h_bytes_last = sum(t.numel() * t.element_size() for t in (loss_sum, correct, count, pcount, nan))
        for k, pl in enumerate(plans):
            if bool(nan[k]):
                rep.nan_clients.append(pl.name)
                if is_loan:
                    raise ValueError(f"NaN in the outputs of client {pl.name}")   # loan_model.py:25-26
        # ---- train rows / logs
        for k, pl in enumerate(plans):
            slot = my_slots[k]
            for e in range(pl.n_epochs):
                n = int(count[k, e])
                acc = 100.0 * float(correct[k, e]) / n if n else 0.0
                tl = float(loss_sum[k, e]) / n if n else 0.0
                if is_loan:   # loan_train.py:32: starts at (round start epoch - 1), +1 per internal epoch
                    tle = self._loan_tle.get(pl.name, self._round_epoch - 1) + 1
                    self._loan_tle[pl.name] = tle
                else:
                    tle = (epoch - 1) * pl.n_epochs + e + 1
                tag = "___PoisonTrain" if pl.poisoning else "___Train"
                extra = f", train_poison_data_count: {int(pcount[k, e])}" if pl.poisoning else ""
                logger.info("{} {},  epoch {:3d}, local model {}, internal_epoch {:3d},  Average loss: {:.4f}, "
                            "Accuracy: {}/{} ({:.4f}%){}".format(tag, self.local_name, epoch, pl.name, e + 1, tl,
                                                                 int(correct[k, e]), n, acc, extra))
                rows.append((slot, "train", [pl.name, tle, epoch, e + 1, tl, acc, int(correct[k, e]), n]))
                if p["vis_train"]:
                    self.metrics.train_vis(tle, acc, tl, str(pl.name), pl.poisoning)
            if pl.n_epochs:
                num_samples[k] = int(count[k, pl.n_epochs - 1])

        # ---- per-batch traces (vis_train_batch_loss / batch_track_distance windows of the reference)
        for trace, fn in ((getattr(res, "step_loss", None), "loss"), (getattr(res, "step_dist", None), "dist")):
            if trace is None:
                continue
            tr = trace.cpu()
            for k, pl in enumerate(plans):
                for step in range(pl.n_epochs * pl.nb):
                    e, b = divmod(step, pl.nb)
                    tle = (epoch - 1) * pl.n_epochs + e + 1
                    if fn == "loss":
                        self.metrics.train_batch_vis(tle, pl.nb, b, float(tr[k, step]), str(pl.name), pl.poisoning)
                    else:
                        self.metrics.track_distance_batch_vis(tle, pl.nb, b, float(tr[k, step]), str(pl.name),
                                                              pl.poisoning)

        # ---- evaluations + scaling; jobs are batched over models that share the same inputs
        scaling = [k for k, pl in enumerate(plans) if pl.poisoning and not p["baseline"]]
        jobs: List[Tuple[str, int, torch.Tensor, int, str]] = []   # (split, trig_row, model_row, k, tag)
        if scaling:
            logger.info(f"Global model norm: {float(be.global_norm(self.G[None])[0])}.")
            gn, dn = be.global_norm(W[scaling]).cpu(), be.dist_norm(W[scaling], anchor[scaling]).cpu()
            for j, k in enumerate(scaling):
                logger.info(f"Norm before scaling: {float(gn[j])}. Distance: {float(dn[j])}")
        pre = W[scaling].clone() if scaling else None
        pre_of = {k: j for j, k in enumerate(scaling)}
        if scaling:
            sc = torch.tensor([plans[k].scale for k in scaling], device=self.device)
            sub = W[scaling]
            be.scale_(sub, anchor[scaling], sc)
            W[scaling] = sub
            if cnt.numel():
                for k in scaling:       # counters scale too (integer arithmetic under torch<=1.4)
                    cnt[k] = anchor_cnt[k] + ((cnt[k] - anchor_cnt[k]).double() * plans[k].scale).trunc().long()
            gn, dn = be.global_norm(W[scaling]).cpu(), be.dist_norm(W[scaling], anchor[scaling]).cpu()
            for j, k in enumerate(scaling):
                logger.info(f"Scaling by  {plans[k].scale}")
                logger.info(f"Scaled Norm after poisoning: {float(gn[j])}, distance: {float(dn[j])}")
                if not is_loan:
                    rows.append((my_slots[k], "scale", [epoch, float(dn[j])]))
        for split, trig, k, tag in self._eval_job_meta(plans):
            row = pre[pre_of[k]] if tag.startswith("pre_") else W[k]
            jobs.append((split, trig, row, k, tag))
        if self._defer_client_eval():
            # enqueue only: the rows of these tests are recorded by flush_pending(), after the next round's local phase
            self._pending_client = (epoch, plans, list(my_slots), rows, self._launch_eval_jobs(jobs, epoch))
            return
        self._record_client_evals(epoch, plans, my_slots, rows, self._run_eval_jobs(jobs, epoch))

    def _record_client_evals(self, epoch, plans, my_slots, rows, results) -> None:
        p = self.params
        order_tags = ["pre_clean", "pre_poison", "clean", "poison", "agent"]
        for k, pl in enumerate(plans):
            slot = my_slots[k]
            for tag in order_tags:
                r = results.get((k, tag))
                if r is None:
                    continue
                loss, acc, corr, n = r
                poisoned = tag in ("pre_poison", "poison")
                if tag != "agent":
                    logger.info("___Test {} poisoned: {}, epoch: {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.4f}%)"
                                .format(self.local_name, poisoned, epoch, loss, corr, n, acc))
                if tag in ("pre_clean", "clean"):
                    rows.append((slot, "test", [pl.name, epoch, loss, acc, corr, n]))
                    if tag == "clean":
                        self.metrics.test_vis(epoch, acc, None, str(pl.name))
                elif poisoned:
                    rows.append((slot, "poison", [pl.name, epoch, loss, acc, corr, n]))
                    if tag == "poison":
                        self.metrics.poison_test_vis(epoch, acc, None, str(pl.name))
                else:
                    rows.append((slot, "trigger", [pl.name, f"{pl.name}_trigger", "", epoch, loss, acc, corr, n]))
                    if p["vis_trigger_split_test"]:
                        self.metrics.trigger_agent_test_vis(epoch, acc, f"{pl.name}_trigger")

    def _eval_job_meta(self, plans: List[ClientPlan]) -> List[Tuple[str, int, int, str]]:
        """(split, trigger row, client index, tag) of every per-client evaluation that follows local training, in the
        order they are run.  Depends on the round plan only, so every rank can also derive the other ranks' lists."""
        p = self.params
        is_loan = p["type"] == C.TYPE_LOAN
        out: List[Tuple[str, int, int, str]] = []
        if not p["client_eval"]:
            return out
        scaling = [k for k, pl in enumerate(plans) if pl.poisoning and not p["baseline"]]
        if not is_loan:
            for k in scaling:                       # image_train.py:148-164: tests of the unscaled poisoned model
                out.append(("clean", -1, k, "pre_clean"))
                out.append(("poison", 0, k, "pre_poison"))
        for k, pl in enumerate(plans):
            if is_loan or not pl.poisoning:                       # loan_train.py:198-201 / image_train.py:268-271
                out.append(("clean", -1, k, "clean"))
            if p["is_poison"] and pl.poisoning:                   # image_train.py:274-282
                out.append(("poison", 0, k, "poison"))
            if p["is_poison"] and pl.adv_index >= 0:              # image_train.py:285-295
                out.append(("poison", pl.adv_index + 1, k, "agent"))
        return out

    def _run_eval_jobs(self, jobs, epoch=None) -> Dict[Tuple[int, str], Tuple[float, float, int, int]]:
        pairs = getattr(self.backend, "evaluate_pairs", None)
        if pairs is not None and self.world > 1:
            out = self._run_eval_jobs_balanced(jobs, pairs, epoch)
            if out is not None:
                return out
        if pairs is not None and jobs:
            res = pairs(torch.stack([j[2] for j in jobs]), [(j[0], max(j[1], 0)) for j in jobs])
            if res is not None:
                out = {}
                for (split, trig, _, k, tag), (loss, corr, n) in zip(jobs, res):
                    out[(k, tag)] = (loss / n if n else 0, 100.0 * corr / n if n else 0, corr, n)
                return out
        groups: Dict[Tuple[str, int], list] = {}
        for split, trig, row, k, tag in jobs:
            groups.setdefault((split, trig), []).append((row, k, tag))
        out = {}
        for (split, trig), items in groups.items():
            models = torch.stack([it[0] for it in items])
            loss, correct, n = self.backend.evaluate(models, split, max(trig, 0))
            loss, correct = loss.cpu(), correct.cpu()
            for j, (_, k, tag) in enumerate(items):
                acc = 100.0 * float(correct[j]) / n if n else 0
                out[(k, tag)] = (float(loss[j]) / n if n else 0, acc, int(correct[j]), n)
        return out

    def _defer_client_eval(self) -> bool:
        """overlap_client_eval (off by default, needs overlap_global_eval): the per-client tests of a round are deferred
        the same way.  Measured at 8 GPUs: 18.7 ms of tests disappear from the round, the overlapped training phase gets
        7.4 ms longer, host-side scheduling 2.7 ms - 82.5 -> 81.4 ms per round device-timed, but end to end 89.9 -> 92.6:
        not worth it by default (the evaluation passes keep every SM busy for 50-100 us at a time and the launch-bound
        training kernels queue behind them)."""
        return (self.overlap_global_eval and bool(self.params.get("overlap_client_eval", False))
                and self.params["type"] in C.IMAGE_TYPES and int(self.params["aggr_epoch_interval"]) == 1)

    def _launch_eval_jobs(self, jobs, epoch):
        """Deferred form of _run_eval_jobs, first half: the inputs of the evaluation - a copy of the model rows, after the
        exchange between ranks (a collective, so it happens now, on the main stream).  The kernels are enqueued by
        _enqueue_client_evals() once the aggregation kernels are in the queue (an evaluation pass keeps every SM busy
        for milliseconds; the cooperative aggregation kernels would wait behind it)."""
        if self.world > 1:
            prep = self._balanced_inputs(jobs, epoch)
            if prep is None:
                return ["none", None, jobs, None]
            models_all, jobs_all, counts = prep
            return ["balanced", (models_all, jobs_all), jobs, counts]
        if not jobs:
            return ["none", None, jobs, None]
        return ["single", (torch.stack([j[2] for j in jobs]), [(j[0], max(j[1], 0)) for j in jobs]), jobs, None]

    def _enqueue_client_evals(self) -> None:
        if self._pending_client is None:
            return
        handle = self._pending_client[4]
        if handle[0] == "none":
            return
        models, meta = handle[1]
        side = self._eval_stream
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            shard = (self.rank, self.world) if handle[0] == "balanced" else None
            handle[1] = self.backend.exec.evaluate_pairs(models, meta, shard=shard, defer=True)

    def _finish_eval_jobs(self, handle):
        kind, h, jobs, counts = handle
        if kind == "none":
            return {}
        res = self.backend.finalize_eval(h)
        if kind == "balanced":
            return self._balanced_reduce(res, jobs, counts)
        out = {}
        for (split, trig, _, k, tag), (loss, corr, n) in zip(jobs, res):
            out[(k, tag)] = (loss / n if n else 0, 100.0 * corr / n if n else 0, corr, n)
        return out

    def _run_eval_jobs_balanced(self, jobs, pairs, epoch=None):
        """Several ranks: the client -> rank placement is uneven (10 clients on 8 GPUs, attackers with extra tests),
        so the models to evaluate are exchanged (~11 MB each over NVLink) and EVERY rank evaluates ALL
        (model, input set) pairs on its own slice of the test split; the (loss, correct, count) sums are then
        all-reduced and each rank keeps the rows of its own jobs.  Every rank derives all ranks' job lists from the
        round plan (no metadata exchange); with the peer-memory communicator the models are pulled over NVLink by a
        gather kernel and the sums go through the fused all-reduce - no NCCL call, no pickling."""
        prep = self._balanced_inputs(jobs, epoch)
        if prep is None:
            return {}
        models_all, jobs_all, counts = prep
        res = pairs(models_all, jobs_all, shard=(self.rank, self.world))
        if res is None:
            return None
        return self._balanced_reduce(res, jobs, counts)

    def _balanced_inputs(self, jobs, epoch=None):
        names = self._round_names
        ep = self._round_epoch if epoch is None else epoch
        owner = self._placement(names, self._round_epoch)
        gathered = []
        for r in range(self.world):
            nm_r = [names[s] for s in range(len(names)) if owner[s] == r]
            meta = self._eval_job_meta(plan_clients(self.params, self.data, nm_r, ep))
            gathered.append([(m[0], max(m[1], 0)) for m in meta])
        assert gathered[self.rank] == [(j[0], max(j[1], 0)) for j in jobs], "evaluation job lists out of step"
        counts = [len(g) for g in gathered]
        total = sum(counts)
        if total == 0:
            return None
        Df = self.spec.Df
        mine = torch.stack([j[2][:Df] for j in jobs]) if jobs else torch.zeros(0, Df, device=self.device)
        if hasattr(self.comm, "gather_rows"):
            models_all = self.comm.gather_rows(mine.contiguous(), counts)
        else:
            models_all = self.comm.all_gather_rows(mine.contiguous(), counts)
        jobs_all = [m for g in gathered for m in g]
        return models_all, jobs_all, counts

    def _balanced_reduce(self, res, jobs, counts):
        t = torch.tensor([[r[0], float(r[1]), float(r[2])] for r in res], dtype=torch.float64, device=self.device)
        self.comm.all_reduce_sum_(t)
        lo = sum(counts[: self.rank])
        rows = t[lo: lo + counts[self.rank]].cpu().tolist()
        out = {}
        for (split, trig, _, k, tag), (loss, corr, n) in zip(jobs, rows):
            corr, n = int(round(corr)), int(round(n))
            out[(k, tag)] = (loss / n if n else 0, 100.0 * corr / n if n else 0, corr, n)
        return out

    def _emit_rows(self, rows_local, counts) -> None:
        rows = rows_local
        if self.world > 1 and self.gather_rows:
            import torch.distributed as dist
            gathered = [None] * self.world
            dist.all_gather_object(gathered, rows_local)
            rows = [r for part in gathered for r in part]
        if not self.is_main:
            return
        rows = sorted(rows, key=lambda r: r[0])     # stable: keeps each client's internal order
        rec = self.recorder
        for _, table, row in rows:
            if table == "train":
                rec.train_result.append(row)
            elif table == "test":
                rec.test_result.append(row)
            elif table == "poison":
                rec.posiontest_result.append(row)
            elif table == "trigger":
                rec.poisontriggertest_result.append(row)
            elif table == "scale":
                rec.add_scale(row[0], row[1])

    # ---------------------------------------------------------------------------------------- aggregation
    def _aggregate(self, epoch, names, adversaries, my_slots, counts, order, W, cnt, grad_sum, num_samples,
                   rep) -> None:
        p, be, comm = self.params, self.backend, self.comm
        method = p["aggregation_methods"]
        sigma = float(p["sigma"]) if p["diff_privacy"] else None
        all_ns = self._gather_num_samples(num_samples, counts, order, len(names))
        names_in_order = [names[s] for s in order]
        if method == C.AGGR_MEAN:
            fused = getattr(be, "fedavg_from_states_", None)
            if fused is not None and fused(self.G, self.G_cnt, W, cnt if cnt.numel() else None, float(p["eta"]),
                                           int(p["no_models"]), comm, sigma, int(p["seed"]) * 1000003 + epoch):
                return
            if hasattr(comm, "fedavg_"):      # fused peer-memory kernel: reduce + update + broadcast, no NCCL
                comm.fedavg_(W, float(p["eta"]) / int(p["no_models"]), len(names), sigma or 0.0,
                             int(p["seed"]) * 1000003 + epoch)
                if cnt.numel():
                    c = (cnt - self.G_cnt[None, :]).sum(0).double()
                    comm.all_reduce_sum_(c)
                    self.G_cnt.add_((c * (float(p["eta"]) / int(p["no_models"]))).trunc().long())
                return
            delta = W - self.G[None, :]
            dcnt = (cnt - self.G_cnt[None, :]) if cnt.numel() else None
            agg.fedavg_update_(self.G, self.G_cnt, delta, dcnt, float(p["eta"]), int(p["no_models"]), comm, sigma,
                               self.noise_gen)
            return
        self._log_poison_ratio("rfa" if method == C.AGGR_GEO_MED else "foolsgold", names, all_ns)
        if method == C.AGGR_GEO_MED:
            fused = getattr(be, "geometric_median_states", None)
            if fused is not None and comm.world == 1:
                res = fused(W, self.G, [all_ns[s] for s in order], int(p["geom_median_maxiter"]))
            elif hasattr(comm, "weiszfeld") and self.device.type == "cuda":
                # F2: one fused launch per Weiszfeld iteration and rank, convergence decided on the device
                tot = float(sum(all_ns.values())) or 1.0
                a_loc = torch.tensor([all_ns[s] / tot for s in my_slots], dtype=torch.float32, device=self.device)
                n_tot = len(names)
                med, table, state, obj = comm.weiszfeld(W, self.G, a_loc, sum(counts[: self.rank]), n_tot,
                                                        int(p["geom_median_maxiter"]))
                st, tb, ob = state.cpu(), table.cpu(), obj.cpu()           # the round's only RFA read-back
                comm.check()
                iters = int(st[1])
                res = agg.RFAResult(int(st[2]), [float(v) for v in tb[0].tolist()], [float(v) for v in tb[1].tolist()],
                                    med, float(ob[int(st[3])]), iters)
            else:
                delta = W - self.G[None, :]
                ns = torch.tensor(num_samples, dtype=torch.float64, device=self.device)
                res = agg.geometric_median(delta, ns, [all_ns[s] for s in order], counts,
                                           maxiter=int(p["geom_median_maxiter"]), comm=comm)
            if hasattr(comm, "weiszfeld") and comm.world > 1 and self.device.type == "cuda":
                # every rank holds the same median and draws the same DP noise (shared seed): no broadcast
                self.G.add_(res.median[: self.G.numel()], alpha=float(p["eta"]))
                if sigma is not None:
                    self.G.add_(agg.dp_noise_like(self.G, sigma, self.noise_gen))
            else:
                agg.rfa_update_(self.G, res, float(p["eta"]), sigma, comm, self.noise_gen)
            logger.info(f"[rfa agg] iters: {res.iters}, objective: {res.objective}, weight: {res.wv}")
            wv, alphas = res.wv, res.distances
        else:
            feats_local = grad_sum[:, self.spec.fg_slice] if grad_sum is not None and grad_sum.shape[0] else \
                torch.zeros(0, self.spec.fg_len, device=self.device)
            res = None
            if hasattr(be, "foolsgold_native"):
                res = be.foolsgold_native(names_in_order, feats_local.contiguous(), counts, comm,
                                          bool(p["fg_use_memory"]))
            if res is None:       # torch / NCCL path (CPU, or peer memory unavailable)
                feats = comm.all_gather_rows(feats_local.contiguous(), counts)
                fmat = self.fg.features(names_in_order, feats)
                res = agg.foolsgold_weights(fmat)
            wv_t, alpha_t = res
            logger.info(f"[foolsgold agg] wv: {wv_t.tolist()}")
            self.fg.wv_history.append(wv_t.tolist())
            lo = sum(counts[: self.rank])
            wv_local = wv_t[lo: lo + counts[self.rank]].to(self.device)
            gs = grad_sum if grad_sum is not None else torch.zeros(0, self.spec.Dp, device=self.device)
            done = hasattr(be, "foolsgold_apply_native") and be.foolsgold_apply_native(
                self.G, gs, wv_local, len(names), float(p["eta"]), float(p["lr"]), float(p["decay"]), comm)
            if not done:
                agg.foolsgold_update_(self.G, self.spec.Dp, gs, wv_local, len(names), float(p["eta"]),
                                      float(p["lr"]), float(p["decay"]), comm)
            wv, alphas = [float(v) for v in wv_t.tolist()], [float(v) for v in alpha_t.tolist()]
        rep.agg_weights, rep.agg_alphas = wv, alphas
        if self.is_main:
            self.recorder.add_weight_result(names_in_order, wv, alphas)
            for nm, w_, a_ in zip(names_in_order, wv, alphas):
                self.metrics.weight_vis(epoch, w_, str(nm), nm in adversaries)
                self.metrics.alpha_vis(epoch, a_, str(nm), nm in adversaries)

    def _gather_num_samples(self, num_samples, counts, order, n) -> Dict[int, float]:
        """Samples seen by every selected client in its last internal epoch = its partition size (every rank knows
        the partition, so nothing is exchanged)."""
        names = self._round_names
        return {s: float(self.data.client_size(names[s])) for s in range(n)}

    def _log_poison_ratio(self, tag, names, all_ns) -> None:
        p = self.params
        tot = sum(all_ns.values()) or 1.0
        adv = sum(all_ns[s] for s, nm in enumerate(names) if nm in p["adversary_list"]) / tot
        logger.info(f"[{tag} agg] training data poison_ratio: {adv}  data num: {[all_ns[s] for s in range(len(names))]}")
        logger.info(f"[{tag} agg] considering poison per batch poison_fraction: "
                    f"{adv * p['poisoning_per_batch'] / p['batch_size']}")

    # ---------------------------------------------------------------------------------------- global eval
    def _eval_global_jobs(self, jobs) -> List[Tuple[float, float, int, int]]:
        """All global evaluations of a round in one batched pass; with several ranks every rank evaluates its shard
        of the test split and the three sums are all-reduced (R5)."""
        shard = (self.rank, self.world) if self.world > 1 else None
        return self._reduce_eval_rows(self.backend.evaluate_jobs(self.G, jobs, shard=shard))

    def _reduce_eval_rows(self, res) -> List[Tuple[float, float, int, int]]:
        t = torch.tensor([[r[0], float(r[1]), float(r[2])] for r in res], dtype=torch.float64, device=self.device)
        if self.world > 1:
            self.comm.all_reduce_sum_(t)
        out = []
        for row in t.cpu().tolist():
            loss, c, n = row[0], int(round(row[1])), int(round(row[2]))
            out.append((loss / n if n else 0, 100.0 * c / n if n else 0, c, n))
        return out

    def _global_eval(self, epoch: int, rep: RoundReport) -> None:
        jobs, tests = self._global_jobs()
        self._record_global_eval(epoch, rep, tests, self._eval_global_jobs(jobs))

    def _global_jobs(self):
        p = self.params
        jobs = [("clean", 0)]
        tests = []
        if p["is_poison"]:
            jobs.append(("poison", 0))
            if p.is_centralized_attack:
                if p["centralized_test_trigger"]:
                    tests = [(f"global_in_index_{j}_trigger", j + 1) for j in range(int(p["trigger_num"]))]
            else:
                tests = [(f"global_in_{a}_trigger", p.adversary_index(a) + 1) for a in p["adversary_list"]]
            jobs += [("poison", row) for _, row in tests]
        return jobs, tests

    def _record_global_eval(self, epoch: int, rep: RoundReport, tests, results) -> None:
        p, rec = self.params, self.recorder
        tge = epoch + p["aggr_epoch_interval"] - 1
        loss, acc, corr, n = results[0]
        logger.info("___Test {} poisoned: {}, epoch: {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.4f}%)".format(
            self.target_name, False, tge, loss, corr, n, acc))
        rep.global_loss, rep.global_acc = loss, acc
        if self.is_main:
            rec.add_test("global", tge, loss, acc, corr, n)
            rec.close_scale_row(acc)
            self.metrics.test_vis(tge, acc, None, "global")
        if not p["is_poison"]:
            return
        loss, acc_p, corr, n = results[1]
        logger.info("___Test {} poisoned: {}, epoch: {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.4f}%)".format(
            self.target_name, True, tge, loss, corr, n, acc_p))
        rep.poison_loss, rep.poison_acc = loss, acc_p
        if self.is_main:
            rec.add_poison_test("global", tge, loss, acc_p, corr, n)
            rec.add_trigger_test("global", "combine", "", tge, loss, acc_p, corr, n)
            if p["vis_trigger_split_test"]:
                self.metrics.trigger_agent_test_vis(epoch, acc_p, "global_combine")
        for (tname, _), (l_, a_, c_, n_) in zip(tests, results[2:]):
            rep.trigger_acc[tname] = a_
            if self.is_main:
                rec.add_trigger_test("global", tname, "", epoch, l_, a_, c_, n_)
                if p["vis_trigger_split_test"]:
                    self.metrics.trigger_agent_test_vis(epoch, a_, tname)
