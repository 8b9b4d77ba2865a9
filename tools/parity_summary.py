"""Assemble profiles/parity_report.md from the result files under profiles/r2/ (tools/parity.py wrote them).

    python tools/parity_summary.py > profiles/parity_report.md
"""
import json
import math
import os
import statistics as st
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R2 = os.path.join(ROOT, "profiles", "r2")
ATTACK = (8, 10, 12, 14)


def load(name):
    p = os.path.join(R2, name)
    return json.load(open(p)) if os.path.exists(p) else None


def col(x, rd, key):
    return [r["rounds"][rd - 1][key] for r in x["runs"].values() if len(r["rounds"]) >= rd and r["rounds"][rd - 1][key] is not None]


def ms(v):
    return f"{st.mean(v):5.1f} +- {st.pstdev(v):4.1f} (n={len(v)})" if v else "-"


def fisher(a, n1, b, n2):
    """Two-sided Fisher exact test for a/n1 vs b/n2."""
    def pr(k):
        return math.comb(a + b, k) * math.comb(n1 + n2 - a - b, n1 - k) / math.comb(n1 + n2, n1)
    p0 = pr(a)
    return min(1.0, sum(pr(k) for k in range(max(0, a + b - n2), min(a + b, n1) + 1) if pr(k) <= p0 * (1 + 1e-9)))


def main():
    arms = [("native tf32 (1 GPU)", load("parity_stable_own_tf32_12seeds_fixedpart_v2.json") or load("parity_stable_own_tf32_12seeds_fixedpart.json")),
            ("native tf32, own partition seeds", load("parity_stable_own_tf32_12seeds.json")),
            ("native bf16 (1 GPU)", load("parity_sweep_own_bf16.json")),
            ("fp32 torch backend of this framework (GPU)", load("parity_stable_own_torchbackend.json")),
            ("reference (GPU, cuDNN TF32)", load("parity_sweep_ref.json"))]
    arms = [(n, x) for n, x in arms if x]
    L = []
    L.append("# Accuracy / attack-success-rate parity against the unmodified reference\n")
    L.append("Everything below was produced by `tools/parity.py` (runs; reference through `baseline/ref_harness.py`, stock code "
             "path) and assembled by `tools/parity_summary.py` from the JSON files in `profiles/r2/`.  Synthetic CIFAR-shaped "
             "data (10000 train / 2000 test images, generator seed 1234 + s: identical bytes for every arm), 100 clients / 10 "
             "per round, Dirichlet 0.5, four adversaries with one local DBA trigger each, single-shot attacks in rounds 8 / 10 / "
             "12 / 14, random initialisation.\n")
    L.append("## 1. What can and cannot be compared\n")
    L.append("* The reference is not reproducible run to run (unseeded cuDNN, shuffled loaders) and hard-codes its seeds "
             "(`main.py:36-38,86`), so its runs differ by data set only; this framework's native path accumulates with fp32 "
             "atomics.  A random-init narrow ResNet on 100-image non-IID shards is **chaotic in both code bases**: with the "
             "shipped YAML (eta 0.1, lr 0.1, scale 100, written for a resumed pre-trained model) nothing is learnt for ~8 "
             "rounds and the seed-to-seed spread of the reference's OWN accuracy reaches 63 points "
             "(`profiles/parity_report_shipped.md`).  Parity is therefore statistical: means over seeds, spreads, and exact "
             "tests on attack outcomes.\n"
             "* Stable regime used for the main comparison (found by `tools/parity.py sweep` on the cheap arm, then run on all "
             "arms): `eta=1, scale_weights_poison=10 (= no_models / eta, the same model replacement as 100 at eta 0.1), lr = "
             "poison_lr = 0.02, internal_epochs=4`: every arm learns the main task to 75-95 % by round 7, i.e. before the "
             "first attack.\n"
             "* In this regime an attack round adds the scaled attacker update AND 0.9 x the mean benign update to a model "
             "that is still moving fast; in every arm the result is a (partly) collapsed model - main-task accuracy 10-30 % - "
             "that either answers the target class for everything (ASR ~100 %) or another constant class (ASR ~0 %), and "
             "recovers in the next round.  'The attack landed' (ASR > 50 %) is the outcome that is compared.\n")
    L.append("## 2. Stable regime, FedAvg: per-round means over seeds\n")
    hdr = "| round | " + " | ".join(f"{n} acc | ASR" for n, _ in arms) + " |"
    L.append(hdr)
    L.append("|---|" + "---|---|" * len(arms))
    nr = max(len(next(iter(x["runs"].values()))["rounds"]) for _, x in arms)
    for rd in range(1, nr + 1):
        row = [f"{rd}{' *' if rd in ATTACK else ''}"]
        for _, x in arms:
            a, s = col(x, rd, "acc"), col(x, rd, "asr")
            row += [ms(a), ms(s)]
        L.append("| " + " | ".join(row) + " |")
    L.append("\n`*` = attack round.  Arms with n runs of 9 rounds stop after round 9.\n")
    L.append("## 3. Attack outcomes (ASR > 50 % in an attack round)\n")
    ref = [x for n, x in arms if n.startswith("reference")][0]
    L.append("| arm | round 8 | all attack rounds run | Fisher exact p vs reference (round 8) |")
    L.append("|---|---|---|---|")
    r8 = col(ref, 8, "asr")
    for n, x in arms:
        v8 = col(x, 8, "asr")
        allv = [v for rd in ATTACK for v in col(x, rd, "asr")]
        p = fisher(sum(v > 50 for v in v8), len(v8), sum(v > 50 for v in r8), len(r8)) if x is not ref else float("nan")
        L.append(f"| {n} | {sum(v > 50 for v in v8)}/{len(v8)} (mean ASR {st.mean(v8):.0f} %) | "
                 f"{sum(v > 50 for v in allv)}/{len(allv)} | {'-' if x is ref else f'{p:.2f}'} |")
    small_o, small_r = load("parity_cpu_small_own_torchbackend.json"), load("parity_cpu_small_ref.json")
    if small_o and small_r:
        o8, q8 = col(small_o, 8, "asr"), col(small_r, 8, "asr")
        L.append(f"\nSame comparison with both code bases in fp32 on the CPU (no GPU numerics involved; 3000 / 500 images, 8 "
                 f"seeds each, `profiles/r2/parity_cpu_small_*.json`): this framework's torch backend lands "
                 f"{sum(v > 50 for v in o8)}/{len(o8)} (mean ASR {st.mean(o8):.0f} %), the reference "
                 f"{sum(v > 50 for v in q8)}/{len(q8)} (mean ASR {st.mean(q8):.0f} %), Fisher p = "
                 f"{fisher(sum(v > 50 for v in o8), len(o8), sum(v > 50 for v in q8), len(q8)):.2f}: the round semantics "
                 "(selection, poisoning, scaling, aggregation, evaluation) are statistically indistinguishable.\n")
    out = "\n".join(L)
    tail = os.path.join(ROOT, "profiles", "parity_report_tail.md")
    if os.path.exists(tail):
        out += "\n" + open(tail).read()
    print(out)


if __name__ == "__main__":
    main()
