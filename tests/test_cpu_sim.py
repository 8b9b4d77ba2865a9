"""CPU integration tests: full rounds through the public API on synthetic data (BASELINE config 1 plumbing), the
three aggregators, LOAN, checkpoint resume, CSV output, and world-size invariance with a 2-process gloo group."""
import csv
import os
import subprocess
import sys

import pytest
import torch
import yaml

from dba_mod_b200.config import Params
from dba_mod_b200.fl.simulator import Simulator

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _raw(name, **over):
    raw = yaml.safe_load(open(os.path.join(ROOT, "utils", name)))
    raw.update(resumed_model=False, synthetic_data=True)
    raw.update(over)
    return raw


def _mnist(**over):
    base = dict(synthetic_train_size=3000, synthetic_test_size=400, epochs=3, internal_poison_epochs=2,
                **{"0_poison_epochs": [2], "1_poison_epochs": [3], "2_poison_epochs": [9], "3_poison_epochs": [9]})
    base.update(over)
    return _raw("mnist_params.yaml", **base)


def test_mnist_fedavg_dba_round_trip(tmp_path):
    """BASELINE config 1: MNIST MnistNet, 100 clients / 10 per round, DBA 4-part trigger, FedAvg, CPU."""
    sim = Simulator(Params(_mnist(save_model=True, save_on_epochs=[2])), out_root=str(tmp_path), quiet=True)
    reps = sim.run()
    sim.close()
    assert [r.epoch for r in reps] == [1, 2, 3]
    assert reps[1].adversaries == [41] and reps[1].names[0] == 41 and reps[2].adversaries == [73]
    # single-shot model replacement: eta/no_models * scale = 1 -> the scaled attacker flips the global ASR
    assert reps[1].poison_acc > 90.0 and reps[0].poison_acc < 30.0
    folder = sim.folder_path
    files = set(os.listdir(folder))
    assert {"log.txt", "params.yaml", "train_result.csv", "test_result.csv", "posiontest_result.csv",
            "poisontriggertest_result.csv", "scale_result.csv", "metrics.jsonl", "model_last.pt.tar",
            "model_last.pt.tar.epoch_2", "model_last.pt.tar.best"} <= files
    rows = list(csv.reader(open(os.path.join(folder, "train_result.csv"))))
    assert rows[0][0] == "local_model" and len(rows) == 1 + 3 * 10 + 2 * 1     # attackers: 2 internal epochs
    trig = list(csv.reader(open(os.path.join(folder, "poisontriggertest_result.csv"))))
    names = {r[1] for r in trig[1:]}
    assert {"combine", "global_in_41_trigger", "41_trigger"} <= names
    test_rows = list(csv.reader(open(os.path.join(folder, "test_result.csv"))))
    assert sum(1 for r in test_rows if r[0] == "global") == 3
    # resume from the epoch-2 checkpoint (reference format) continues at epoch 3 with the saved weights
    ck = torch.load(os.path.join(folder, "model_last.pt.tar.epoch_2"), weights_only=False)
    assert ck["epoch"] == 2 and list(ck["state_dict"].keys())[0] == "conv1.weight"
    rel = os.path.relpath(os.path.join(folder, "model_last.pt.tar.epoch_2"), str(tmp_path))
    sim2 = Simulator(Params(_mnist(resumed_model=True, resumed_model_name=rel)), out_root=str(tmp_path), quiet=True)
    assert sim2.start_epoch == 3
    for k, v in ck["state_dict"].items():
        assert torch.equal(sim2.state_dict()[k], v)
    sim2.close()


@pytest.mark.parametrize("method", ["geom_median", "foolsgold"])
def test_mnist_robust_aggregators(tmp_path, method):
    sim = Simulator(Params(_mnist(aggregation_methods=method)), out_root=str(tmp_path), quiet=True)
    reps = sim.run()
    sim.close()
    r = reps[1]
    assert len(r.agg_weights) == 10 and len(r.agg_alphas) == 10
    if method == "geom_median":
        assert r.agg_weights[0] < 0.02 and r.poison_acc < 50.0       # scaled attacker is down-weighted, attack fails
    rows = list(csv.reader(open(os.path.join(sim.folder_path, "weight_result.csv"))))
    assert len(rows) == 9 and rows[3][0] == "41"


def test_centralized_attack_uses_global_trigger(tmp_path):
    raw = _mnist(adversary_list=[45], epochs=2, **{"0_poison_epochs": [2]})
    sim = Simulator(Params(raw), out_root=str(tmp_path), quiet=True)
    reps = sim.run()
    sim.close()
    assert reps[1].adversaries == [45]
    assert set(reps[1].trigger_acc) == {f"global_in_index_{j}_trigger" for j in range(4)}


def test_loan_round(tmp_path):
    raw = _raw("loan_params.yaml", synthetic_train_size=6000, epochs=2, internal_poison_epochs=3,
               **{"0_poison_epochs": [2], "1_poison_epochs": [9], "2_poison_epochs": [9]})
    sim = Simulator(Params(raw), out_root=str(tmp_path), quiet=True)
    assert len(sim.data.state_names) == 51 and sim.data.n_features == 91
    reps = sim.run()
    sim.close()
    assert reps[1].adversaries == ["CT"] and reps[1].names[0] == "CT"
    assert reps[1].poison_acc is not None and "global_in_CT_trigger" in reps[1].trigger_acc
    rows = list(csv.reader(open(os.path.join(sim.folder_path, "train_result.csv"))))
    ct = [r for r in rows if r[0] == "CT" and r[2] == "2"]
    assert [int(r[3]) for r in ct] == [1, 2, 3] and [int(r[1]) for r in ct] == [2, 3, 4]   # temp_local_epoch counter


def test_aggr_epoch_interval_two(tmp_path):
    raw = _mnist(aggr_epoch_interval=2, epochs=4, **{"0_poison_epochs": [2]})
    sim = Simulator(Params(raw), out_root=str(tmp_path), quiet=True)
    reps = sim.run()
    sim.close()
    assert [r.epoch for r in reps] == [1, 3] and reps[0].adversaries == [41]


_WORKER = r"""
import os, sys, json, torch, yaml
sys.path.insert(0, {root!r})
import torch.distributed as dist
from dba_mod_b200.config import Params
from dba_mod_b200.fl.simulator import Simulator
from dba_mod_b200.parallel.launcher import make_comm
raw = yaml.safe_load(open(os.path.join({root!r}, "utils", "mnist_params.yaml")))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=2000, synthetic_test_size=300, epochs=2,
           internal_poison_epochs=2, aggregation_methods={method!r}, backend="torch")
raw["0_poison_epochs"] = [2]
comm = make_comm("gloo")
sim = Simulator(Params(raw), device=torch.device("cpu"), comm=comm, quiet=True, write_files=False)
reps = sim.run()
if comm.rank == 0:
    print("RESULT " + json.dumps({{"norm": float(sim.G.double().norm()), "acc": reps[-1].global_acc,
                                   "asr": reps[-1].poison_acc, "w": reps[-1].agg_weights}}))
sim.close()
if comm.world > 1:
    dist.destroy_process_group()
"""


@pytest.mark.parametrize("method", ["mean", "geom_median", "foolsgold"])
def test_world_size_invariance_gloo(tmp_path, method):
    """Aggregation results do not depend on the number of ranks / client placement (fp32 reduction-order tolerance)."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT, method=method))

    def run(world):
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29500 + os.getpid() % 400 + world),
                   WORLD_SIZE=str(world), OMP_NUM_THREADS="2")
        procs = []
        for r in range(world):
            e = dict(env, RANK=str(r), LOCAL_RANK=str(r))
            procs.append(subprocess.Popen([sys.executable, str(script)], env=e, stdout=subprocess.PIPE,
                                          stderr=subprocess.PIPE, text=True))
        outs = [p.communicate(timeout=600) for p in procs]
        for p, (o, e_) in zip(procs, outs):
            assert p.returncode == 0, e_[-2000:]
        line = [l for l in outs[0][0].splitlines() if l.startswith("RESULT ")][0]
        import json
        return json.loads(line[7:])

    a, b = run(1), run(2)
    assert abs(a["norm"] - b["norm"]) / a["norm"] < 1e-4, (a, b)
    assert abs(a["acc"] - b["acc"]) < 1.0 and abs(a["asr"] - b["asr"]) < 2.0
    if a["w"] is not None:
        assert max(abs(x - y) for x, y in zip(sorted(a["w"]), sorted(b["w"]))) < 1e-3


def test_client_placement_follows_the_lock_step_cost_model(tmp_path):
    """Placement of a round's clients on ranks: everybody placed exactly once, the modelled makespan
    (a * longest client + b * sum of steps per rank) is no worse than balancing by step count alone, and a long
    attacker shares its launch floor with short clients instead of being isolated by step count."""
    from dba_mod_b200.fl.schedule import plan_clients
    sim = Simulator(Params(_mnist(synthetic_train_size=6000, internal_poison_epochs=6)), out_root=str(tmp_path),
                    quiet=True, write_files=False)
    p = sim.params
    names = list(p["adversary_list"][:1]) + [n for n in sim.data.participants_list if n not in p["adversary_list"]][:9]
    epoch = p.poison_epochs_of(0)[0]
    steps = [pl.n_epochs * pl.nb for pl in plan_clients(p, sim.data, names, epoch)]
    assert max(steps) >= 3 * sorted(steps)[len(steps) // 2]          # the attacker really is the long one
    a, b = 1.0, 0.3

    def makespan(owner, world):
        return max(a * max([steps[s] for s in range(len(names)) if owner[s] == r], default=0)
                   + b * sum(steps[s] for s in range(len(names)) if owner[s] == r) for r in range(world))

    for world in (2, 3, 4):
        sim.world = world
        owner = sim._placement(names, epoch)
        assert len(owner) == len(names) and set(owner) <= set(range(world))
        counts, order = sim._owner_counts(names, epoch)
        assert sum(counts) == len(names) and sorted(order) == list(range(len(names)))
        load, lpt = [0] * world, [0] * len(names)               # plain longest-processing-time on step counts
        for st, s in sorted(((st, s) for s, st in enumerate(steps)), key=lambda t: (-t[0], t[1])):
            r = min(range(world), key=lambda i: (load[i], i))
            load[r] += st
            lpt[s] = r
        assert makespan(owner, world) <= makespan(lpt, world) + 1e-9
    sim.world = 1
    sim.close()
