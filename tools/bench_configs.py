"""Times the other BASELINE.json configurations through the public API (own arm only), device-timed rounds:

    python tools/bench_configs.py loan|rfa1000|tiny_fg|cifar_fg|cifar_rfa|mnist [--rounds K] [--warmup W]

Prints one JSON line per config.  Multi-GPU: launch under torchrun like bench.py.
"""
import argparse
import json
import os
import sys
import time

import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dba_mod_b200.config import Params  # noqa: E402
from dba_mod_b200.fl.simulator import Simulator  # noqa: E402
from dba_mod_b200.parallel.launcher import make_comm  # noqa: E402


def cfg(name, W, K):
    def load(f):
        raw = yaml.safe_load(open(os.path.join(ROOT, "utils", f)))
        raw.update(resumed_model=False, synthetic_data=True, epochs=W + K)
        return raw
    if name == "loan":          # config 2: LOAN MLP, 51 state clients, DBA feature trigger, FedAvg
        raw = load("loan_params.yaml")
        raw.update(synthetic_train_size=200000, number_of_total_participants=51)
        for i in range(3):
            raw[f"{i}_poison_epochs"] = [W + 2 + 2 * i]
    elif name == "rfa1000":     # config 4: CIFAR RFA, 1000 clients / 80 per round
        raw = load("cifar_params.yaml")
        raw.update(aggregation_methods="geom_median", number_of_total_participants=1000, no_models=80,
                   native_eval_chunk=250)
        for i in range(4):
            raw[f"{i}_poison_epochs"] = [W + 2 + i]
    elif name == "cifar_rfa":
        raw = load("cifar_params.yaml")
        raw.update(aggregation_methods="geom_median")
        for i in range(4):
            raw[f"{i}_poison_epochs"] = [W + 2 + i]
    elif name == "cifar_fg":
        raw = load("cifar_params.yaml")
        raw.update(aggregation_methods="foolsgold")
        for i in range(4):
            raw[f"{i}_poison_epochs"] = [W + 2 + i]
    elif name == "tiny_fg":     # config 5: Tiny-ImageNet ResNet-18, FoolsGold, 100 clients / 10 per round
        raw = load("tiny_params.yaml")
        raw.update(aggregation_methods="foolsgold", native_eval_chunk=250, dirichlet_alpha=0.5)
        for i in range(4):
            raw[f"{i}_poison_epochs"] = [W + 2 + i]
    elif name == "mnist":       # config 1: MNIST MnistNet, 100 clients / 10 per round, DBA, FedAvg
        raw = load("mnist_params.yaml")
        for i in range(4):
            raw[f"{i}_poison_epochs"] = [W + 2 + (i % 2)]
    else:
        raise SystemExit(f"unknown config {name}")
    return raw


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    a = ap.parse_args()
    if a.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        from ref_harness import run_reference
        raw = cfg(a.config, a.warmup, a.rounds)
        n_train = raw.get("synthetic_train_size")
        for k in ("synthetic_data", "synthetic_train_size", "synthetic_test_size", "native_eval_chunk"):
            raw.pop(k, None)
        res = run_reference(raw, a.warmup, a.rounds, train_size=n_train)
        print(json.dumps({"config": a.config, "impl": "reference", "ms_per_round": res.get("ms_per_round"),
                          "rounds_per_s": 1e3 / res["ms_per_round"] if res.get("ms_per_round") else None,
                          "global_acc": res.get("global_acc"), "poison_acc": res.get("poison_acc")}), flush=True)
        return
    comm = make_comm()
    sim = Simulator(Params(cfg(a.config, a.warmup, a.rounds)), comm=comm, quiet=True, write_files=False)
    ep = 1
    for _ in range(a.warmup):
        sim.run_round(ep); ep += 1
    comm.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    reps = []
    for _ in range(a.rounds):
        reps.append(sim.run_round(ep)); ep += 1
    e1.record(); comm.barrier(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.rounds
    if comm.rank == 0:
        be = sim.backend
        print(json.dumps({"config": a.config, "n_gpus": comm.world, "ms_per_round": ms, "rounds_per_s": 1e3 / ms,
                          "wall_ms_per_round": (time.perf_counter() - t0) * 1e3 / a.rounds,
                          "steps_last_round": reps[-1].steps, "global_acc": reps[-1].global_acc,
                          "poison_acc": reps[-1].poison_acc, "agg_weights": reps[-1].agg_weights,
                          "backend": be.name, "native_train": bool(getattr(be, "native_train", False)),
                          "native_eval": bool(getattr(be, "native_eval", False)),
                          "collectives": sim.collective_path,
                          "phase_ms": {k: v / (a.rounds + a.warmup) for k, v in sim.timer.resolve().items()}}), flush=True)
    sim.close()


if __name__ == "__main__":
    main()
