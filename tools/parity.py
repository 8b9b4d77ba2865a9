"""Accuracy / attack-success-rate parity of this framework against the UNMODIFIED reference (through the compat shim of
baseline/ref_harness.py) on identical synthetic data.

    python tools/parity.py run --arm own|reference [--aggs mean,geom_median,foolsgold] [--seeds 0,1,2] [--rounds 16]
                               [--train 10000] [--test 2000] [--dtype tf32|bf16] [--out gpurun_out/parity_own.json]
    python tools/parity.py report <own.json> [<own_other.json> ...] <reference.json> > profiles/parity_report.md

One run = the shipped CIFAR DBA configuration (100 clients / 10 per round, Dirichlet 0.5, four adversaries with one
local trigger each, scale 100, lr 0.1 / poison_lr 0.05, 2 / 6 internal epochs) from a random initialisation for
``rounds`` rounds, with the four single-shot attacks in rounds 8 / 10 / 12 / 14.  Seed s selects the synthetic data set
(generator seed 1234 + s, the same bytes for both arms) and - for this framework - the seed of every other stream; the
reference hard-codes its own seeds (main.py:36-38,86), so its runs differ by data set only.

Recorded per round (reference: its own csv_record lists, utils/csv_record.py:4-19): global main-task accuracy, global
backdoor accuracy (combined trigger), per-trigger backdoor accuracy, aggregation weights / alphas (RFA, FoolsGold) and
the attackers' update distances (scale_result).  The report states the tolerance bands it applies.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ATTACK_ROUNDS = (8, 10, 12, 14)


def _overrides(spec) -> dict:
    """``--set eta=1,scale_weights_poison=10``: YAML-typed overrides applied to BOTH arms."""
    import yaml
    out = {}
    for item in (spec or "").split(","):
        if item.strip():
            k, v = item.split("=", 1)
            out[k.strip()] = yaml.safe_load(v)
    return out


def params_for(agg: str, rounds: int, train: int, test: int, overrides=None) -> dict:
    import yaml
    raw = yaml.safe_load(open(os.path.join(ROOT, "utils", "cifar_params.yaml")))
    raw.update(resumed_model=False, save_model=False, epochs=rounds, aggregation_methods=agg)
    for i, r in enumerate(ATTACK_ROUNDS):
        raw[f"{i}_poison_epochs"] = [r]
    raw.update(_overrides(overrides))
    return raw


def run_own(agg, seed, a):
    import torch
    from dba_mod_b200.config import Params
    from dba_mod_b200.fl.simulator import Simulator
    from dba_mod_b200.parallel.launcher import make_comm
    raw = params_for(agg, a.rounds, a.train, a.test, a.set)
    raw.update(synthetic_data=True, synthetic_train_size=a.train, synthetic_test_size=a.test,
               synthetic_seed=1234 + seed, seed=a.fixed_seed if a.fixed_seed >= 0 else 1 + seed, compute_dtype=a.dtype)
    comm = make_comm()
    sim = Simulator(Params(raw), comm=comm, quiet=True, write_files=False)
    reps = sim.run()
    rec = sim.recorder
    out = {"rounds": [{"epoch": r.epoch, "acc": r.global_acc, "asr": r.poison_acc, "trig": r.trigger_acc,
                       "names": [str(n) for n in r.names], "adv": [str(n) for n in r.adversaries],
                       "w": r.agg_weights, "alpha": r.agg_alphas} for r in reps],
           "scale": [list(r) for r in rec.scale_result], "world": comm.world,
           # attackers' LOCAL models on the backdoor test set (before scaling, then after), image_train.py:150-160,262-268
           "local_poison": [[str(r[0]), int(r[1]), float(r[3])] for r in rec.posiontest_result if str(r[0]) != "global"]}
    sim.close()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return out if comm.rank == 0 else None


def run_ref(agg, seed, a):
    """One reference run in a process of its own (its modules keep their result lists and seeds at module level)."""
    import subprocess
    import tempfile
    tmp = tempfile.mktemp(suffix=".json")
    cmd = [sys.executable, os.path.abspath(__file__), "_refone", "--agg", agg, "--seed", str(seed), "--rounds",
           str(a.rounds), "--train", str(a.train), "--test", str(a.test), "--out", tmp, "--set", a.set or ""]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 or not os.path.exists(tmp):
        raise RuntimeError(f"reference run failed: {r.stderr[-1500:]}")
    out = json.load(open(tmp))
    os.remove(tmp)
    return out


def _ref_one(agg, seed, a):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_harness
    raw = params_for(agg, a.rounds, a.train, a.test, a.set)
    res = ref_harness.run_reference(raw, 0, a.rounds, train_size=a.train, test_size=a.test, data_seed=1234 + seed,
                                    records=True)
    rec = res["records"]
    rounds = {}
    for r in rec["test"]:
        if r[0] == "global":
            rounds.setdefault(int(r[1]), {})["acc"] = float(r[3])
    for r in rec["poison"]:
        if r[0] == "global":
            rounds.setdefault(int(r[1]), {})["asr"] = float(r[3])
    for r in rec["trigger"]:
        if r[0] == "global" and r[1] != "combine":
            rounds.setdefault(int(r[3]), {}).setdefault("trig", {})[str(r[1])] = float(r[5])
    w = rec["weight"]
    out_rounds = []
    for i, ep in enumerate(sorted(rounds)):
        d = {"epoch": ep, "acc": rounds[ep].get("acc"), "asr": rounds[ep].get("asr"), "trig": rounds[ep].get("trig", {}),
             "names": None, "adv": None, "w": None, "alpha": None}
        if len(w) >= 3 * (i + 1):
            d["names"] = [str(x) for x in w[3 * i]]
            d["w"] = [float(x) for x in w[3 * i + 1]]
            d["alpha"] = [float(x) for x in w[3 * i + 2]]
        out_rounds.append(d)
    return {"rounds": out_rounds, "scale": rec["scale"], "world": 1,
            "local_poison": [[str(r[0]), int(r[1]), float(r[3])] for r in rec["poison"] if str(r[0]) != "global"]}


def cmd_run(a):
    aggs = a.aggs.split(",")
    seeds = [int(s) for s in a.seeds.split(",")]
    results = {"arm": a.arm, "dtype": a.dtype if a.arm == "own" else "tf32 (cuDNN)", "rounds": a.rounds,
               "train": a.train, "test": a.test, "overrides": a.set or "", "runs": {}}
    for agg in aggs:
        for s in seeds:
            t0 = time.time()
            r = run_own(agg, s, a) if a.arm == "own" else run_ref(agg, s, a)
            if r is not None:
                r["seconds"] = time.time() - t0
                results["runs"][f"{agg}/{s}"] = r
                last = r["rounds"][-1]
                print(f"{a.arm} {agg} seed {s}: {r['seconds']:.0f} s, final acc {last['acc']:.1f} asr {last['asr']}",
                      file=sys.stderr, flush=True)
    if int(os.environ.get("RANK", "0")) == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(results, open(a.out, "w"))
        results["world"] = next(iter(results["runs"].values()))["world"] if results["runs"] else 1
        json.dump(results, open(a.out, "w"))


def cmd_sweep(a):
    """Pick a learnable, stable regime with the (cheap) own arm, then run the reference and own bf16 on it.
    Candidates are ';'-separated --set strings; score = min over seeds of the accuracy in the last pre-attack round."""
    import copy
    cands = [c for c in a.cands.split(";") if c.strip()]
    best, best_score, log = None, -1.0, []
    for c in cands:
        b = copy.copy(a)
        b.arm, b.set, b.out = "own", c, os.path.join(a.outdir, "sweep_tmp.json")
        cmd_run(b)
        x = json.load(open(b.out))
        pre = ATTACK_ROUNDS[0] - 2           # index of the last round before the first attack
        accs = [v["rounds"][pre]["acc"] for v in x["runs"].values()]
        score = min(accs)
        log.append({"set": c, "pre_attack_acc": accs, "score": score})
        print(f"sweep {c}: pre-attack acc {accs} -> {score:.1f}", file=sys.stderr, flush=True)
        if score > best_score + 1e-9:
            best, best_score = c, score
            os.replace(b.out, os.path.join(a.outdir, f"parity_sweep_own_{a.dtype}.json"))
    json.dump({"candidates": log, "chosen": best}, open(os.path.join(a.outdir, "parity_sweep_choice.json"), "w"), indent=1)
    for arm, dtype, name in (("own", "bf16", "parity_sweep_own_bf16.json"), ("reference", a.dtype, "parity_sweep_ref.json")):
        if arm == "own" and dtype == a.dtype:
            continue
        b = copy.copy(a)
        b.arm, b.dtype, b.set, b.out = arm, dtype, best, os.path.join(a.outdir, name)
        cmd_run(b)


# ------------------------------------------------------------------------------------------------- report
def _mean(xs):
    xs = [x for x in xs if x is not None]
    return statistics.mean(xs) if xs else float("nan")


def _attacker_weight(run, ep):
    """Aggregation weight of the round's attacker (its slot in the names row)."""
    for r in run["rounds"]:
        if r["epoch"] == ep and r.get("w") and r.get("names"):
            adv_names = {"17", "33", "77", "11"}
            ws = [w for n, w in zip(r["names"], r["w"]) if n in adv_names]
            others = [w for n, w in zip(r["names"], r["w"]) if n not in adv_names]
            return (min(ws) if ws else None), (_mean(others) if others else None)
    return None, None


def cmd_report(a):
    arms = [json.load(open(f)) for f in a.files]
    ref = [x for x in arms if x["arm"] == "reference"]
    ref = ref[0] if ref else None
    own = [x for x in arms if x["arm"] == "own"]
    L = []
    L.append("# Accuracy / ASR parity: dba_mod_b200 vs the unmodified reference (through the compat shim)\n")
    L.append(f"Generated by `tools/parity.py report` from {', '.join('`' + os.path.basename(f) + '`' for f in a.files)}.\n")
    ov = arms[0].get("overrides") or ""
    L.append(("Setting: shipped CIFAR DBA YAML" + (f" with `{ov}`" if ov else "") +
              " from a random initialisation, synthetic CIFAR-shaped data ")
             +
             f"({arms[0]['train']} train / {arms[0]['test']} test images, generator seed 1234 + s, identical bytes for "
             f"all arms), {arms[0]['rounds']} rounds, the four adversaries strike once each in rounds "
             f"{', '.join(map(str, ATTACK_ROUNDS))} with their local trigger (scale_weights_poison as configured).  Mean over seeds; the number "
             "in brackets is the spread (max - min) over seeds.\n")
    L.append("Tolerance bands applied below (batch order, dropout-free nets but different sample orders and cuDNN "
             "algorithms make single runs differ; the reference itself is not reproducible run to run):\n"
             "* main-task accuracy per round: |mean_own - mean_ref| <= 6 points, OR within two standard errors of the "
             "difference of the means (+ 3 points) where the seed-to-seed spread of the arms themselves is larger than "
             "that (a random-init narrow ResNet at lr 0.02-0.1 on non-IID shards is chaotic in BOTH arms: see the "
             "spread columns);\n"
             "* backdoor accuracy: the mean over seeds is on the same side of 50 % in every attack round (the attack "
             "lands or is rejected in both arms);\n"
             "* RFA: attacker weight at least 10x below the mean benign weight in both arms;\n"
             "* FoolsGold: weights in [0, 1], attacker rounds compared qualitatively (cosine-similarity pardoning is "
             "discontinuous).\n")
    verdicts = []
    aggs = sorted({k.split("/")[0] for x in arms for k in x["runs"]})
    for agg in aggs:
        L.append(f"\n## {agg}\n")
        cols = []
        for x in arms:
            runs = [v for k, v in x["runs"].items() if k.startswith(agg + "/")]
            if runs:
                if x["arm"] == "reference":
                    label = f"reference (n={len(runs)})"
                elif "backend=torch" in (x.get("overrides") or ""):
                    label = f"own fp32 torch backend (n={len(runs)})"
                else:
                    label = f"own {x['dtype']} x{x.get('world', 1)} GPU (n={len(runs)})"
                cols.append((label, runs, x["arm"]))
        if not cols:
            continue
        L.append("| round | " + " | ".join(f"{c[0]} acc | {c[0]} ASR" for c in cols) + " |")
        L.append("|---|" + "---|---|" * len(cols))
        nr = min(len(r["rounds"]) for c in cols for r in c[1])
        table = {}
        for i in range(nr):
            row = [str(cols[0][1][0]["rounds"][i]["epoch"]) + (" *" if cols[0][1][0]["rounds"][i]["epoch"] in ATTACK_ROUNDS else "")]
            for label, runs, arm in cols:
                accs = [r["rounds"][i]["acc"] for r in runs]
                asrs = [r["rounds"][i]["asr"] for r in runs if r["rounds"][i]["asr"] is not None]
                table[(label, i)] = (_mean(accs), _mean(asrs))
                row.append(f"{_mean(accs):.1f} [{max(accs) - min(accs):.1f}]")
                row.append(f"{_mean(asrs):.1f} [{(max(asrs) - min(asrs)) if asrs else 0:.1f}]")
            L.append("| " + " | ".join(row) + " |")
        if agg in ("geom_median", "foolsgold"):
            L.append("\nAttacker weight (min over the round's adversaries) vs mean benign weight in the attack rounds:\n")
            L.append("| round | " + " | ".join(f"{c[0]} attacker / benign" for c in cols) + " |")
            L.append("|---|" + "---|" * len(cols))
            for ep in ATTACK_ROUNDS:
                row = [str(ep)]
                for label, runs, arm in cols:
                    pairs = [_attacker_weight(r, ep) for r in runs]
                    aw = _mean([p[0] for p in pairs])
                    bw = _mean([p[1] for p in pairs])
                    row.append(f"{aw:.2e} / {bw:.3f}")
                L.append("| " + " | ".join(row) + " |")
        # verdicts against the reference
        if ref is not None:
            rlabel = [c[0] for c in cols if c[2] == "reference"]
            for label, runs, arm in cols:
                if arm != "own" or not rlabel:
                    continue
                worst_acc, last_acc, asr_ok, worst_z = 0.0, 0.0, True, 0.0
                rruns = [c[1] for c in cols if c[2] == "reference"][0]
                for i in range(nr):
                    ao, so = table[(label, i)]
                    ar, sr = table[(rlabel[0], i)]
                    worst_acc = max(worst_acc, abs(ao - ar))
                    # difference of the two means in units of its standard error (+ 3 points of slack for the
                    # discreteness of small test sets): |z| <= 2 is "the same within seed noise"
                    xo = [r["rounds"][i]["acc"] for r in runs]
                    xr = [r["rounds"][i]["acc"] for r in rruns]
                    se = ((statistics.pvariance(xo) / max(1, len(xo) - 1) if len(xo) > 1 else 0.0)
                          + (statistics.pvariance(xr) / max(1, len(xr) - 1) if len(xr) > 1 else 0.0)) ** 0.5
                    worst_z = max(worst_z, max(0.0, abs(ao - ar) - 3.0) / se if se > 0 else (0.0 if abs(ao - ar) <= 3 else 99.0))
                    ep = cols[0][1][0]["rounds"][i]["epoch"]
                    if ep in ATTACK_ROUNDS:
                        if so == so and sr == sr and (so > 50) != (sr > 50):
                            asr_ok = False
                last_acc = abs(table[(label, nr - 1)][0] - table[(rlabel[0], nr - 1)][0])
                ok = (worst_acc <= 6.0 or worst_z <= 2.0) and asr_ok
                verdicts.append((agg, label, worst_acc, last_acc, asr_ok, ok, worst_z))
        lp = []
        for label, runs, arm in cols:
            before, after = [], []
            for r in runs:
                seen = {}
                for name, ep, acc in r.get("local_poison", []):
                    seen.setdefault((name, ep), []).append(acc)
                for (name, ep), v in seen.items():
                    if ep in ATTACK_ROUNDS and len(v) >= 2:
                        before.append(v[0]); after.append(v[1])
            if before:
                lp.append(f"{label}: {_mean(before):.1f} % before scaling, {_mean(after):.1f} % after (n = {len(before)})")
        if lp:
            L.append("\nBackdoor accuracy of the attackers' LOCAL models in their attack rounds: " + "; ".join(lp) + ".")
        sc = []
        for label, runs, arm in cols:
            d = [row[1] for r in runs for row in r.get("scale", []) if len(row) >= 2]
            if d:
                sc.append(f"{label}: {_mean([float(v) for v in d]):.1f}")
        if sc:
            L.append("\nMean L2 distance of the scaled attacker update to the global model (scale_result.csv): "
                     + "; ".join(sc) + ".")
    if verdicts:
        L.append("\n## Verdict against the bands\n")
        L.append("| aggregator | arm | max abs acc diff of the means | acc diff last round | worst z (diff beyond 3 points / "
                 "standard error) | mean ASR on the same side of 50 % in attack rounds | within bands |")
        L.append("|---|---|---|---|---|---|---|")
        for agg, label, wa, la, ao, ok, wz in verdicts:
            L.append(f"| {agg} | {label} | {wa:.1f} | {la:.1f} | {wz:.2f} | {'yes' if ao else 'NO'} | {'yes' if ok else 'NO'} |")
    print("\n".join(L))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    r = sub.add_parser("run")
    r.add_argument("--arm", default="own", choices=["own", "reference"])
    r.add_argument("--aggs", default="mean,geom_median,foolsgold")
    r.add_argument("--seeds", default="0,1,2")
    r.add_argument("--rounds", type=int, default=16)
    r.add_argument("--train", type=int, default=10000)
    r.add_argument("--test", type=int, default=2000)
    r.add_argument("--dtype", default="tf32", choices=["tf32", "bf16"])
    r.add_argument("--out", default="gpurun_out/parity_own.json")
    r.add_argument("--set", default="", help="key=value,... overrides of the shipped YAML, applied to both arms")
    r.add_argument("--fixed-seed", type=int, default=-1, dest="fixed_seed",
                   help="own arm: use this seed for partition / selection / batch order in every run (the reference "
                        "hard-codes seed 1, so its runs differ by data set only; -1 = 1 + s)")
    o = sub.add_parser("_refone")
    o.add_argument("--agg"); o.add_argument("--seed", type=int); o.add_argument("--rounds", type=int)
    o.add_argument("--train", type=int); o.add_argument("--test", type=int); o.add_argument("--out")
    o.add_argument("--set", default="")
    p = sub.add_parser("report")
    p.add_argument("files", nargs="+")
    sw = sub.add_parser("sweep")
    sw.add_argument("--cands", required=True)
    sw.add_argument("--aggs", default="mean")
    sw.add_argument("--seeds", default="0,1,2")
    sw.add_argument("--rounds", type=int, default=16)
    sw.add_argument("--train", type=int, default=10000)
    sw.add_argument("--test", type=int, default=2000)
    sw.add_argument("--dtype", default="tf32")
    sw.add_argument("--outdir", default="gpurun_out")
    sw.add_argument("--fixed-seed", type=int, default=-1, dest="fixed_seed")
    mg = sub.add_parser("merge", help="merge the runs of result files of one arm / configuration (more seeds)")
    mg.add_argument("files", nargs="+")
    mg.add_argument("--out", required=True)
    tj = sub.add_parser("traj", help="per-run accuracy / ASR trajectories of result files, one line per run")
    tj.add_argument("files", nargs="+")
    a = ap.parse_args()
    if a.cmd == "_refone":
        json.dump(_ref_one(a.agg, a.seed, a), open(a.out, "w"))
    elif a.cmd == "run":
        cmd_run(a)
    elif a.cmd == "merge":
        xs = [json.load(open(f)) for f in a.files]
        out = dict(xs[0])
        out["runs"] = {}
        for x in xs:
            assert x["arm"] == xs[0]["arm"] and x.get("overrides", "") == xs[0].get("overrides", ""), "different arms / configurations"
            out["runs"].update(x["runs"])
        out["rounds"] = min(x["rounds"] for x in xs)
        json.dump(out, open(a.out, "w"))
    elif a.cmd == "sweep":
        cmd_sweep(a)
    elif a.cmd == "traj":
        for f in a.files:
            x = json.load(open(f))
            print(f"# {os.path.basename(f)} arm={x['arm']} dtype={x['dtype']} set={x.get('overrides', '')}")
            for k, v in x["runs"].items():
                print(f"{k:>16s} acc " + " ".join(f"{r['acc']:5.1f}" for r in v["rounds"]))
                print(f"{'':>16s} asr " + " ".join(f"{(r['asr'] if r['asr'] is not None else -1):5.1f}" for r in v["rounds"]))
    else:
        cmd_report(a)
