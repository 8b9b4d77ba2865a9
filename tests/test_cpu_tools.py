"""Benchmark / parity tooling that runs without a GPU: schedule of the benchmark's attack rounds, parity file merging and
the exact test used in the parity report."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def test_bench_places_attacks_inside_every_window_and_in_the_warm_up():
    bench = _load("bench.py", "bench_under_test")
    p = bench.bench_params(3, 5)                       # warm-up 3, 5 timed rounds, two windows (device-timed + end to end)
    assert p["epochs"] == 13 and p["resumed_model"] is False and p["synthetic_data"] is True
    assert p["0_poison_epochs"] == [2, 6, 11]          # warm-up attack + round 3 of both windows
    assert p["1_poison_epochs"] == [8, 13]             # round 5 of both windows
    assert p["2_poison_epochs"] == [] and p["3_poison_epochs"] == []        # rounds 7 / 9 do not fit into 5 rounds
    ref = bench.bench_params(3, 5, windows=1)          # the reference arm is timed over one window, same offsets
    assert ref["0_poison_epochs"] == [2, 6] and ref["1_poison_epochs"] == [8]
    for name in bench.CONFIGS:
        q = bench.bench_params(1, 2, config=name)
        assert q["epochs"] == 5 and bench.metric_name(name)
    assert bench.bench_params(3, 5, config="cifar_rfa_1000")["no_models"] == 80


def test_parity_merge_and_exact_test(tmp_path):
    parity = _load("tools/parity.py", "parity_under_test")
    summ = _load("tools/parity_summary.py", "parity_summary_under_test")
    assert parity._overrides("eta=1,scale_weights_poison=10,lr=0.02,flag=true") == {
        "eta": 1, "scale_weights_poison": 10, "lr": 0.02, "flag": True}
    raw = parity.params_for("geom_median", 9, 1000, 200, "eta=1")
    assert raw["aggregation_methods"] == "geom_median" and raw["epochs"] == 9 and raw["eta"] == 1
    assert [raw[f"{i}_poison_epochs"] for i in range(4)] == [[8], [10], [12], [14]]
    # Fisher's exact test, two-sided: textbook values
    assert abs(summ.fisher(3, 12, 3, 3) - 0.0440) < 1e-3          # 3/12 vs 3/3
    assert abs(summ.fisher(5, 8, 4, 8) - 1.0) < 1e-9
    assert summ.fisher(0, 10, 10, 10) < 1e-4
    run = {"rounds": [{"epoch": 1, "acc": 10.0, "asr": 0.0, "trig": {}, "names": None, "adv": None, "w": None,
                       "alpha": None}], "scale": [], "world": 1}
    for i, seeds in enumerate(([0, 1], [2])):
        json.dump({"arm": "own", "dtype": "tf32", "rounds": 1 + i, "train": 10, "test": 2, "overrides": "eta=1",
                   "runs": {f"mean/{s}": run for s in seeds}}, open(tmp_path / f"p{i}.json", "w"))
    argv = sys.argv
    try:
        sys.argv = ["parity.py", "merge", str(tmp_path / "p0.json"), str(tmp_path / "p1.json"), "--out",
                    str(tmp_path / "m.json")]
        import runpy
        runpy.run_path(os.path.join(ROOT, "tools", "parity.py"), run_name="__main__")
    finally:
        sys.argv = argv
    m = json.load(open(tmp_path / "m.json"))
    assert sorted(m["runs"]) == ["mean/0", "mean/1", "mean/2"] and m["rounds"] == 1
