"""Headline benchmark: FL rounds/sec, CIFAR-10 ResNet-18, 100 clients / 10 per round, DBA + FedAvg (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N --steps K ...  # unmodified reference (baseline/_ref) + compat shim

A "step" is one full FL round exactly as the reference's main loop performs it (main.py:135-235): client selection,
local training of the 10 selected clients (attackers train poisoned + scale), every client's full-test-set
evaluations (Q16), FedAvg aggregation, global clean / backdoor / per-trigger evaluations.  Synthetic CIFAR-shaped data
(50 000 train / 10 000 test, 32x32x3 uint8) and random-init weights - there is no network for datasets/checkpoints.
Both arms use the same YAML (utils/cifar_params.yaml with ``resumed_model: false`` and the four adversaries'
single-shot attack rounds moved into the timed window) and the same synthetic data generator.

Timing: W untimed warm-up rounds, then K rounds bracketed by barrier + torch.cuda.synchronize(), timed with CUDA
events on the launching stream, MAX over ranks.  ``value`` runs with the dataset resident on the device; ``e2e`` is a
second timed window through the same public API (``Simulator.run_round``) in which every round's inputs (the selected
clients' training images and the test images) are first copied host->device from pinned memory and the round's
metric block is read back device->host.  Strong scaling: the round's work is fixed and sharded over the N ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fl_rounds_per_sec_cifar10_resnet18"


def metric_name(config: str) -> str:
    return METRIC if config == "cifar_fedavg" else f"fl_rounds_per_sec_{config}"


def data_label(p) -> str:
    t = p["type"]
    if t == "loan":
        return f"synthetic LOAN-shaped table ({p.get('synthetic_train_size') or 60000} rows, 91 features, 51 states), random-init weights"
    shape = {"cifar": "CIFAR-10-shaped uint8 images (50000 train / 10000 test)",
             "mnist": "MNIST-shaped uint8 images (60000 train / 10000 test)",
             "tiny-imagenet-200": "Tiny-ImageNet-shaped uint8 images (100000 train / 10000 test, 64x64, 200 classes)"}[t]
    return f"synthetic {shape}, random-init weights"


def dtype_label(p, be, cuda: bool) -> str:
    """Compute dtype of the native executor that actually ran (LOAN is an fp32 CUDA-core MLP, the image networks
    follow compute_dtype)."""
    if not cuda:
        return "fp32"
    ex = getattr(be, "exec", None)
    if p["type"] == "loan":
        return "fp32"
    return getattr(ex, "compute_dtype", "fp32") if ex is not None else "fp32"


# --config: the BASELINE.json configurations (the headline is the default).  name -> (yaml, overrides, description)
CONFIGS = {
    "cifar_fedavg": ("cifar_params.yaml", {}, "CIFAR ResNet-18 (narrow 32/64/128/256, 2.8M params), 100 clients / 10 per "
                     "round, DBA 4 local triggers scale 100, FedAvg"),
    "cifar_rfa": ("cifar_params.yaml", {"aggregation_methods": "geom_median"}, "CIFAR ResNet-18, 100 / 10, DBA, RFA"),
    "cifar_foolsgold": ("cifar_params.yaml", {"aggregation_methods": "foolsgold"},
                        "CIFAR ResNet-18, 100 / 10, DBA, FoolsGold"),
    "cifar_rfa_1000": ("cifar_params.yaml", {"aggregation_methods": "geom_median", "number_of_total_participants": 1000,
                                             "no_models": 80, "native_eval_chunk": 250},
                       "CIFAR ResNet-18, RFA geometric median, 1000 clients / 80 per round"),
    "tiny_foolsgold": ("tiny_params.yaml", {"aggregation_methods": "foolsgold", "native_eval_chunk": 250,
                                            "dirichlet_alpha": 0.5},
                       "Tiny-ImageNet ResNet-18 (64x64, 200 classes), 100 / 10, DBA, FoolsGold"),
    "loan_fedavg": ("loan_params.yaml", {"synthetic_train_size": 200000, "number_of_total_participants": 51},
                    "LOAN MLP 91-46-23-9, 51 state clients / 10 per round, DBA feature triggers, FedAvg"),
    "mnist_fedavg": ("mnist_params.yaml", {}, "MnistNet, 100 clients / 10 per round, DBA 4-part trigger, FedAvg"),
}


def bench_params(warmup: int, steps: int, windows: int = 2, config: str = "cifar_fedavg") -> dict:
    import yaml
    fname, over, _ = CONFIGS[config]
    with open(os.path.join(ROOT, "utils", fname)) as f:
        p = yaml.safe_load(f)
    p["resumed_model"] = False
    p["save_model"] = False
    p["synthetic_data"] = True
    p["epochs"] = warmup + windows * steps
    p.update(over)
    # shipped schedules resume from a checkpoint and strike at rounds r, r+2, r+4, r+6 -> the same offsets are placed
    # inside every timed window (single-shot attacks of the adversaries, 3 / 5 / 7 / 9 rounds into the window)
    for i in range(int(p.get("trigger_num", 0))):
        if f"{i}_poison_epochs" in p:
            p[f"{i}_poison_epochs"] = [warmup + w * steps + 3 + 2 * i for w in range(windows)
                                       if 3 + 2 * i <= steps]
    # one attack inside the warm-up as well, so that the attacker-only paths (poisoned schedule, scaling, backdoor
    # tests, their buffers) exist before the first timed round - in both arms, they share this schedule
    if warmup >= 2 and "0_poison_epochs" in p:
        p["0_poison_epochs"] = [2] + list(p["0_poison_epochs"])
    return p


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0) -> None:
        super().__init__(daemon=True)
        self.idx, self.samples, self._stop_evt, self.proc = gpu_index, [], threading.Event(), None

    def run(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                if self._stop_evt.is_set():
                    break
                self.samples.append(line.strip())
        except Exception:  # noqa: BLE001
            pass

    def stop(self) -> dict:
        self._stop_evt.set()
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [t.strip() for t in s.split(",")]
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
                for nm, v in zip(names, f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _dist_setup(n_gpus: int):
    import torch
    from dba_mod_b200.parallel.launcher import make_comm
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    return make_comm()


def _max_over_ranks(ms: float, comm) -> float:
    import torch
    if comm.world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device="cuda" if torch.cuda.is_available() else "cpu")
    import torch.distributed as dist
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_own(args) -> dict:
    import torch
    from dba_mod_b200.config import Params
    from dba_mod_b200.fl.simulator import Simulator
    from dba_mod_b200.ops import native

    comm = _dist_setup(args.gpus)
    W, K = args.warmup, args.steps
    raw = bench_params(W, K, config=args.config)
    raw["compute_dtype"] = args.dtype
    if comm.world > 1:
        # several GPUs: the training phase is launch-bound (one or two clients per GPU), so a round's global evaluations
        # run on a side stream under the next round's training (Simulator.overlap_global_eval)
        raw["overlap_global_eval"] = True
    if args.small:
        raw.update(synthetic_train_size=5000, synthetic_test_size=1000)
    params = Params(raw)
    cuda = torch.cuda.is_available()
    sim = Simulator(params, comm=comm, quiet=True, write_files=False)
    dev = sim.device
    be = sim.backend

    def sync():
        comm.barrier()
        if cuda:
            torch.cuda.synchronize()

    epoch = 1
    for _ in range(W):
        sim.run_round(epoch); epoch += 1
    sim.flush_pending()
    sim.timer.resolve()
    sim.timer.reset()                   # phase_ms covers the timed windows only

    def timed_window(stream_inputs: bool):
        nonlocal epoch
        sim.stream_inputs = stream_inputs
        sampler = ClockSampler(dev.index or 0) if (cuda and comm.rank == 0) else None
        if sampler:
            sampler.start()
        sync()
        l0 = native.launch_count() if native.available() else 0
        if cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = time.perf_counter()
        reports = []
        for _ in range(K):
            reports.append(sim.run_round(epoch)); epoch += 1
        sim.flush_pending()             # a deferred global evaluation of the last round belongs to the timed region
        if cuda:
            e1.record()
        sync()
        wall_ms = (time.perf_counter() - t0) * 1e3
        ms = e0.elapsed_time(e1) if cuda else wall_ms
        clocks = sampler.stop() if sampler else None
        launches = (native.launch_count() - l0) if native.available() else 0
        return _max_over_ranks(ms, comm), _max_over_ranks(wall_ms, comm), clocks, launches, reports

    ms, wall_ms, clocks, launches, reps = timed_window(False)
    sim.prepare_streaming()            # one-time pinning of the host data set (not part of any round)
    ms_e2e, wall_e2e, _, _, reps2 = timed_window(True)
    out = {
        "metric": metric_name(args.config), "value": K / (ms / 1e3), "unit": "rounds/s", "n_gpus": comm.world,
        "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": dtype_label(params, be, cuda),
        "data": data_label(params),
        "impl": "dba_mod_b200",
        "config": {"name": args.config, "model": CONFIGS[args.config][2],
                   "global_batch": int(params["no_models"]) * int(params["batch_size"]),
                   "seq_len": None, "clients_total": int(params["number_of_total_participants"]),
                   "clients_per_round": int(params["no_models"]), "batch_size": int(params["batch_size"]),
                   "aggregation": params["aggregation_methods"], "attack": "DBA, single-shot, scaled",
                   "parallelism": f"client-sharded x{comm.world}", "timed_l2": "inputs larger than L2 (no flush)",
                   "per_client_eval": True, "backend": be.name,
                   "native_train": bool(getattr(be, "native_train", False)),
                   "native_eval": bool(getattr(be, "native_eval", False)),
                   "collectives": sim.collective_path, "overlap_global_eval": bool(sim.overlap_global_eval)},
        "e2e": {"value": K / (ms_e2e / 1e3), "unit": "rounds/s", "h2d_bytes_per_step": sim.h2d_bytes_last,
                "d2h_bytes_per_step": sim.d2h_bytes_last, "ms_per_step": ms_e2e / K,
                "wall_ms_per_step": wall_e2e / K},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "wall_ms_per_step": wall_ms / K,
        "phase_ms": {k: v / (2 * K) for k, v in sim.timer.resolve().items()},
        "final": {"global_acc": reps2[-1].global_acc, "poison_acc": reps2[-1].poison_acc},
    }
    sim.close()
    return out


def run_reference(args) -> dict:
    W, K = args.warmup, args.steps
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_harness
    if not ref_harness.reference_dir():
        return {"impl": "reference", "unavailable": "neither baseline/_ref nor /root/reference holds the reference"}
    if int(os.environ.get("RANK", "0")) != 0:
        return {}          # the reference is a single-process, single-GPU program (config.py:2); other ranks idle
    try:
        import torch
        _run = ref_harness.run_reference
        raw = bench_params(W, K, windows=1, config=args.config)
        for k in ("synthetic_data",):
            raw.pop(k, None)
        sampler = ClockSampler(0) if torch.cuda.is_available() else None
        if sampler:
            sampler.start()
        small = (5000, 1000) if args.small else (raw.get("synthetic_train_size"), raw.get("synthetic_test_size"))
        for k in ("synthetic_train_size", "synthetic_test_size", "native_eval_chunk"):
            raw.pop(k, None)
        res = _run(raw, W, K, train_size=small[0], test_size=small[1],
                   log_path=os.path.join(ROOT, "gpurun_out", "reference_run.log") if os.path.isdir(
                       os.path.join(ROOT, "gpurun_out")) else None)
        clocks = sampler.stop() if sampler else None
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference run failed: {type(e).__name__}: {str(e)[:200]}"}
    ms = res.get("ms_per_round")
    if ms is None:
        return {"impl": "reference", "unavailable": f"reference finished only {res.get('rounds')} rounds"}
    return {
        "metric": metric_name(args.config), "value": 1e3 / ms, "unit": "rounds/s", "n_gpus": 1, "steps": K,
        "warmup": W,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp32" if raw["type"] == "loan" else "tf32",      # fp32 tensors, cuDNN TF32 convolutions (torch default)
        "data": data_label(raw),
        "impl": "reference",
        "config": {"name": args.config, "model": CONFIGS[args.config][2],
                   "global_batch": int(raw["no_models"]) * int(raw["batch_size"]),
                   "clients_total": int(raw["number_of_total_participants"]), "clients_per_round": int(raw["no_models"]),
                   "batch_size": int(raw["batch_size"]), "aggregation": raw["aggregation_methods"],
                   "attack": "DBA, single-shot, scaled", "parallelism": "single process, single GPU",
                   "note": "wall-clock between round ends after cuda synchronize; H2D per batch is part of its stock path"},
        "e2e": {"value": 1e3 / ms, "unit": "rounds/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None},
        "gpu_launches": None, "clocks": clocks,
        "final": {"global_acc": res.get("global_acc"), "poison_acc": res.get("poison_acc")},
    }


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--config", default="cifar_fedavg", choices=sorted(CONFIGS),
                    help="which BASELINE.json configuration to time (default: the headline)")
    ap.add_argument("--dtype", default="tf32", choices=["tf32", "bf16"],
                    help="compute mode of the native executor: tf32 = fp32 activations + TF32 tensor-core math (the "
                         "reference's precision class, default), bf16 = fast mode")
    ap.add_argument("--small", action="store_true", help="5000/1000 images (CPU smoke test only; not a valid number)")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: whatever libraries print while the benchmark runs (NCCL's version banner,
    # the reference's progress prints) is diverted to stderr at the file-descriptor level
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        out = run_reference(args) if args.impl == "reference" else run_own(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if int(os.environ.get("RANK", "0")) == 0 and out:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
