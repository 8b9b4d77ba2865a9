"""CPU tests of the auxiliary pieces: diagnostics (reference's off-path Helper functions), data prep, metrics stream."""
import json
import os

import pytest
import torch

from dba_mod_b200.fl import diagnostics as dg
from dba_mod_b200.models import initial_state


def test_diagnostics_on_flat_vectors():
    spec, flat, _ = initial_state("mnist")
    ref = flat.clone()
    w = flat.clone()
    w[:spec.Dp] += 0.01
    mv = dg.model_max_values(spec, w, ref)
    assert len(mv) == len(spec.param_names) and all(abs(float(v) - 0.01) < 1e-6 for v in mv)
    assert float(dg.cos_sim_loss(spec, ref, ref, 100.0)) == pytest.approx(0.0, abs=2e-2)
    assert float(dg.model_cosine_similarity(spec, ref, ref)) == pytest.approx(0.0, abs=2e-2)
    assert float(dg.accum_similarity(spec, w, w)) == pytest.approx(0.0, abs=1e-5)
    assert dg.get_one_vec(spec, w).shape == (spec.Dp,)


def test_epoch_submit_json(tmp_path):
    spec, flat, _ = initial_state("loan")
    deltas = torch.stack([flat * 0 + 1, flat * 0 + 2])
    store = dg.update_epoch_submit_dict(spec, deltas, ["CA", "NY"], 3)
    path = dg.save_epoch_submit_dict(str(tmp_path), store)
    back = json.load(open(path))
    assert set(back["3"].keys()) == {"CA", "NY"} and back["3"]["NY"]["layer3.0.bias"][0] == 2.0


def test_ewc_fisher_and_penalty():
    spec, flat, _ = initial_state("loan")
    ewc = dg.EWC("loan", spec)
    torch.manual_seed(0)
    batches = [(torch.randn(8, 91), torch.randint(0, 9, (8,))) for _ in range(3)]
    fisher = ewc.estimate_fisher(flat, batches)
    assert fisher.shape == (spec.Dp,) and float(fisher.min()) >= 0 and float(fisher.sum()) > 0
    assert float(ewc.ewc_loss(flat, 10.0)) == 0.0
    ewc.consolidate(flat, fisher)
    moved = flat.clone(); moved[:spec.Dp] += 0.1
    assert float(ewc.ewc_loss(moved, 10.0)) == pytest.approx(5.0 * 0.01 * float(fisher.sum()), rel=1e-4)


def test_loan_prep_and_csv_loader(tmp_path):
    import pandas as pd
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import LoanData
    from dba_mod_b200.data import prep, synthetic
    import yaml
    # raw table with an object column, a date-like dropped column and a big-mean numeric column
    df = pd.DataFrame({"id": range(40), "addr_state": ["CA", "NY"] * 20, "grade": ["A", "B", "C", "A"] * 10,
                       "loan_amnt": [12000.0] * 40, "dti": [15.0] * 40, "small": [1.0] * 40,
                       "loan_status": ["Current", "Fully Paid"] * 20, "desc": [None] * 40})
    raw = tmp_path / "loan.csv"
    df.to_csv(raw, index=False)
    out = prep.preprocess_loan(str(raw), str(tmp_path / "loan"))
    assert sorted(os.path.basename(p) for p in out) == ["loan_CA.csv", "loan_NY.csv"]
    ca = pd.read_csv(tmp_path / "loan" / "loan_CA.csv")
    assert "id" not in ca.columns and "desc" not in ca.columns and "addr_state" not in ca.columns
    assert ca["loan_amnt"].iloc[0] == pytest.approx(1.2) and ca["dti"].iloc[0] == pytest.approx(1.5)
    assert set(ca["grade"]) <= {0, 1, 2} and set(ca["loan_status"]) == {0}
    # synthetic per-state CSVs in the reference layout are read back by the loader
    synthetic.write_loan_csvs(str(tmp_path / "data" / "loan"), 3000)
    rawp = yaml.safe_load(open(os.path.join(os.path.dirname(os.path.dirname(__file__)), "utils", "loan_params.yaml")))
    rawp.update(resumed_model=False, synthetic_data=False, data_dir=str(tmp_path / "data"))
    d = LoanData(Params(rawp), torch.device("cpu"))
    assert not d.is_synthetic and d.n_features == 91 and len(d.state_names) == 51
    assert d.feature_dict["pub_rec"] == 3 and "CT" in d.participants_list


def test_tiny_val_reformat(tmp_path):
    from dba_mod_b200.data import prep
    val = tmp_path / "val"
    (val / "images").mkdir(parents=True)
    for i in range(4):
        (val / "images" / f"val_{i}.JPEG").write_bytes(b"x")
    (val / "val_annotations.txt").write_text("".join(f"val_{i}.JPEG\tn0{i % 2}\t0\t0\t1\t1\n" for i in range(4)))
    assert prep.reformat_tiny_val(str(val)) == 4
    assert sorted(os.listdir(val)) == ["n00", "n01"] and len(os.listdir(val / "n00")) == 2


def test_metrics_stream(tmp_path):
    from dba_mod_b200.utils.metrics import MetricsStream, dict_html
    m = MetricsStream(str(tmp_path))
    m.train_vis(1, 50.0, 0.7, "3", True); m.weight_vis(2, 0.1, "17", True); m.close()
    recs = [json.loads(l) for l in open(tmp_path / "metrics.jsonl")]
    assert [r["win"] for r in recs] == ["train_acc", "train_loss", "aggregation_weight"]
    assert "<td>lr</td><td>0.1</td>" in dict_html({"lr": 0.1, "folder_path": "x"}, "t") and "folder_path" not in \
        dict_html({"lr": 0.1, "folder_path": "x"}, "t")


def test_extra_model_definitions_match_reference_layout():
    """ResNet34/50 and SimpleMnist (defined but unused by the reference) keep its state_dict keys and shapes."""
    import sys
    from dba_mod_b200.models import nn_compat as nc
    REF = "/root/reference"
    if not os.path.isdir(REF):
        pytest.skip("reference tree not mounted")
    sys.path.insert(0, REF)
    try:
        for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
            del sys.modules[k]
        from models.resnet_cifar import ResNet34 as R34, ResNet50 as R50
        from models.simple import SimpleMnist
        pairs = [(nc.ResNet34(), R34(name="x", created_time="t")), (nc.ResNet50(), R50(name="x", created_time="t")),
                 (nc.SimpleMnistCompat(), SimpleMnist(name="x", created_time="t"))]
    finally:
        sys.path.remove(REF)
        for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
            del sys.modules[k]
    for ours, theirs in pairs:
        a, b = ours.state_dict(), theirs.state_dict()
        assert list(a.keys()) == list(b.keys())
        assert all(a[k].shape == b[k].shape for k in a)
    y = nc.ResNet50()(torch.randn(2, 3, 32, 32))
    assert y.shape == (2, 10)
