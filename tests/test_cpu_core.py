"""CPU unit tests: config schema, partitioners, selection, triggers, aggregation maths, recorder, checkpoints.
Oracles are the reference's own functions imported read-only from /root/reference where they are importable."""
import copy
import os
import random
import sys
import types

import numpy as np
import pytest
import torch
import yaml

from dba_mod_b200 import config as C
from dba_mod_b200.config import Params, ParamsError, multistep_lr_factors
from dba_mod_b200.data import partition, triggers
from dba_mod_b200.fl import aggregation as agg
from dba_mod_b200.fl.selection import select_clients

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _yaml(name):
    return yaml.safe_load(open(os.path.join(ROOT, "utils", name)))


def _ref_module(name):
    """Import a reference module with its heavy / missing dependencies stubbed (nothing is copied)."""
    if not os.path.isdir(REF):
        pytest.skip("reference tree not mounted")
    for stub in ("visdom", "matplotlib", "matplotlib.pyplot", "cv2"):
        if stub not in sys.modules:
            m = types.ModuleType(stub)
            if stub == "visdom":
                m.Visdom = lambda *a, **k: types.SimpleNamespace()
            sys.modules[stub] = m
    if "matplotlib" in sys.modules and not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    saved = list(sys.path)
    sys.path.insert(0, REF)
    try:
        for k in ("config", "helper", "utils", "utils.csv_record", "models", "image_helper"):
            if k in sys.modules and not (getattr(sys.modules[k], "__file__", None) or "").startswith(REF):
                del sys.modules[k]
        return __import__(name)
    finally:
        sys.path[:] = saved


# ------------------------------------------------------------------------------------------------ config
@pytest.mark.parametrize("name", ["cifar_params.yaml", "mnist_params.yaml", "tiny_params.yaml", "loan_params.yaml"])
def test_shipped_yaml_validates(name):
    p = Params(_yaml(name))
    assert p["type"] in C.ALL_TYPES and p.unknown_keys() == []
    if p.is_image:
        assert len(p.pixel_pattern(-1)) == sum(len(p.pixel_pattern(j)) for j in range(p["trigger_num"]))
    else:
        names, values = p.feature_trigger(-1)
        assert len(names) == len(values) == 6
    assert p.poison_epochs_of(0) == p["0_poison_epochs"]


def test_yaml_matches_reference_values():
    for ours, theirs in [("cifar_params.yaml", "cifar_params.yaml"), ("mnist_params.yaml", "mnist_params.yaml"),
                         ("tiny_params.yaml", "tiny_params.yaml"), ("loan_params.yaml", "loan_params.yaml")]:
        if not os.path.exists(os.path.join(REF, "utils", theirs)):
            pytest.skip("reference tree not mounted")
        a = _yaml(ours)
        b = yaml.safe_load(open(os.path.join(REF, "utils", theirs)))
        for k, v in b.items():
            if k in C.IGNORED_KEYS and k not in a:
                continue        # present in the shipped file but never read by the reference
            assert a.get(k) == v, (ours, k, a.get(k), v)


def test_params_errors():
    raw = _yaml("mnist_params.yaml")
    bad = dict(raw); bad.pop("lr")
    with pytest.raises(ParamsError):
        Params(bad)
    bad = dict(raw, aggregation_methods="median")
    with pytest.raises(ParamsError):
        Params(bad)
    bad = dict(raw); bad.pop("2_poison_pattern")
    with pytest.raises(ParamsError):
        Params(bad)


def test_multistep_lr_quirk_q4():
    assert multistep_lr_factors(10) == pytest.approx([1, 1, .1, .1, .1, .1, .1, .1, .01, .01])
    assert multistep_lr_factors(6) == [1.0] * 6            # 1.2 / 4.8 are never hit on torch >= 2
    assert multistep_lr_factors(6, compat="bisect") == pytest.approx([1, 1, .1, .1, .1, .01])
    assert multistep_lr_factors(10, step_first=True) == pytest.approx([1, .1, .1, .1, .1, .1, .1, .01, .01, .01])
    # same thing through torch's scheduler
    opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=1.0)
    sch = torch.optim.lr_scheduler.MultiStepLR(opt, milestones=[0.2 * 10, 0.8 * 10], gamma=0.1)
    seen = []
    for _ in range(10):
        seen.append(opt.param_groups[0]["lr"]); opt.step(); sch.step()
    assert seen == pytest.approx(multistep_lr_factors(10))


# -------------------------------------------------------------------------------------------- partitions
def test_dirichlet_partition_matches_reference_bit_exact():
    ih = _ref_module("image_helper")
    labels = [i % 10 for i in range(3000)]
    random.Random(3).shuffle(labels)
    classes = partition.build_classes_dict(labels)
    ours = partition.dirichlet_partition(copy.deepcopy(classes), 20, 0.5, random.Random(1), np.random.RandomState(1))
    fake = types.SimpleNamespace(classes_dict=copy.deepcopy(classes))
    random.seed(1); np.random.seed(1)
    theirs = ih.ImageHelper.sample_dirichlet_train_data(fake, 20, alpha=0.5)
    assert {k: list(v) for k, v in ours.items()} == {k: list(v) for k, v in theirs.items()}
    sizes = [len(ours[u]) for u in range(20)]
    assert sum(sizes) <= 3000 and max(sizes) > 2 * min(sizes)        # non-IID, ragged


def test_equal_partition():
    parts = partition.equal_partition(1000, 10, random.Random(1))
    allv = sorted(i for v in parts.values() for i in v)
    assert allv == list(range(1000)) and all(len(v) == 100 for v in parts.values())


# --------------------------------------------------------------------------------------------- selection
def test_selection_modes():
    p = Params(_yaml("cifar_params.yaml"))
    participants = list(range(100))
    benign = list(set(participants) - set(p["adversary_list"]))
    names, adv = select_clients(p, 203, participants, benign, random.Random(1))
    assert adv == [17] and names[0] == 17 and len(names) == 10 and len(set(names)) == 10
    names, adv = select_clients(p, 10, participants, benign, random.Random(1))
    assert adv == [] and len(names) == 10
    # reference stream: same RNG state -> same sample
    rng = random.Random(1)
    exp = random.Random(1).sample(benign + [17, 33, 77, 11], 10)
    assert select_clients(p, 10, participants, benign, rng)[0] == exp
    p2 = Params(dict(_yaml("cifar_params.yaml"), is_random_adversary=True))
    names, adv = select_clients(p2, 203, participants, benign, random.Random(5))
    assert set(adv) == set(names) & {17, 33, 77, 11}
    p3 = Params(dict(_yaml("cifar_params.yaml"), is_random_namelist=False))
    names, adv = select_clients(p3, 1, list(p3["participants_namelist"]), benign, random.Random(1))
    assert names == list(range(10)) and adv == [17, 33, 77, 11]
    # aggr_epoch_interval > 1: attacker forced in when any on-going epoch is a poison epoch
    p4 = Params(dict(_yaml("cifar_params.yaml"), aggr_epoch_interval=2))
    assert select_clients(p4, 202, participants, benign, random.Random(1))[1] == [17]


# ---------------------------------------------------------------------------------------------- triggers
@pytest.mark.parametrize("cfg,shape", [("cifar_params.yaml", (3, 32, 32)), ("mnist_params.yaml", (1, 28, 28)),
                                       ("tiny_params.yaml", (3, 64, 64))])
def test_pixel_trigger_matches_reference(cfg, shape):
    ih = _ref_module("image_helper")
    raw = _yaml(cfg)
    p = Params(raw)
    fake = types.SimpleNamespace(params=raw)
    img = torch.rand(*shape)
    masks = triggers.all_pixel_masks(p, shape[1], shape[2])
    for adv in [-1, 0, 1, 2, 3]:
        ref = ih.ImageHelper.add_pixel_pattern(fake, img, adv)
        ours = triggers.stamp_images_(img.clone(), masks[adv + 1])
        assert torch.equal(ref, ours)


def test_feature_trigger_table():
    p = Params(_yaml("loan_params.yaml"))
    from dba_mod_b200.data.synthetic import loan_feature_names
    fd = {n: i for i, n in enumerate(loan_feature_names())}
    mask, val = triggers.all_feature_triggers(p, fd, 91)
    assert mask.shape == (4, 91) and int(mask[0].sum()) == 6 and int(mask[1].sum()) == 2
    x = torch.zeros(5, 91)
    triggers.stamp_features_(x, mask[2], val[2])
    assert x[0, fd["pub_rec_bankruptcies"]] == 20 and x[0, fd["pub_rec"]] == 100 and float(x.sum()) == 5 * 120


# ------------------------------------------------------------------------------------------- aggregation
def test_foolsgold_matches_reference():
    h = _ref_module("helper")
    fg = h.FoolsGold(use_memory=False)
    rng = np.random.RandomState(0)
    for n, L in [(10, 2560), (6, 207), (25, 500)]:
        g = rng.randn(n, L)
        g[1] = g[4] * 2 + 0.01 * rng.randn(L)
        g[2] = g[4] * 0.5 + 0.3 * rng.randn(L)
        wv_ref, alpha_ref = fg.foolsgold(g.copy())
        wv, alpha = agg.foolsgold_weights(torch.from_numpy(g))
        np.testing.assert_allclose(wv.numpy(), wv_ref, rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(alpha.numpy(), alpha_ref, rtol=1e-9, atol=1e-9)


def test_foolsgold_memory_accumulates():
    st = agg.FoolsGoldState(use_memory=True)
    f = torch.ones(2, 4)
    assert torch.equal(st.features(["a", "b"], f), f.double())
    m = st.features(["b", "c"], 2 * f)
    assert torch.equal(m[0], torch.full((4,), 3.0, dtype=torch.double)) and torch.equal(m[1], 2 * f[1].double())


def test_weiszfeld_matches_reference():
    h = _ref_module("helper")
    torch.manual_seed(0)
    n, D = 8, 3000
    delta = 0.05 * torch.randn(n, D)
    delta[0] *= 60
    ns = [500 + 13 * i for i in range(n)]
    res = agg.geometric_median(delta, torch.tensor(ns, dtype=torch.float64), ns, [n], maxiter=10)
    # reference maths on dict-of-tensors points, re-stated with its own helpers (l2dist / weighted_average_oracle)
    pts = [{"w": delta[i, :1000].clone(), "v": delta[i, 1000:].clone()} for i in range(n)]
    alphas = torch.from_numpy(np.asarray(ns, dtype=np.float64) / sum(ns)).float()
    H = h.Helper
    median = H.weighted_average_oracle(pts, alphas)
    obj = H.geometric_median_objective(median, pts, alphas)
    wv = None
    for i in range(10):
        prev = obj
        w = torch.tensor([a / max(1e-5, H.l2dist(median, q)) for a, q in zip(alphas, pts)], dtype=alphas.dtype)
        w = w / w.sum()
        median = H.weighted_average_oracle(pts, w)
        obj = H.geometric_median_objective(median, pts, alphas)
        if abs(prev - obj) < 1e-6 * obj:
            break
        wv = w.clone()
    med_ref = torch.cat([median["w"], median["v"]])
    torch.testing.assert_close(res.median, med_ref, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(res.wv, wv.numpy(), rtol=1e-4)
    np.testing.assert_allclose(res.distances, [H.l2dist(median, q) for q in pts], rtol=1e-4)
    assert res.wv[0] < 0.01


def test_fedavg_update_and_counters():
    G = torch.zeros(6)
    cnt = torch.tensor([5], dtype=torch.long)
    delta = torch.arange(12, dtype=torch.float32).reshape(2, 6)
    dcnt = torch.tensor([[16], [48]], dtype=torch.long)
    agg.fedavg_update_(G, cnt, delta, dcnt, eta=0.1, no_models=10)
    torch.testing.assert_close(G, delta.sum(0) * 0.01)
    assert int(cnt[0]) == 5          # (16+48)*0.01 = 0.64 truncates to 0 (torch<=1.4 integer add, Q3)
    agg.fedavg_update_(G, cnt, delta, dcnt * 10, eta=0.1, no_models=10)
    assert int(cnt[0]) == 11         # 6.4 -> 6
    G2 = torch.zeros(100000)
    agg.fedavg_update_(G2, torch.zeros(0, dtype=torch.long), torch.zeros(1, 100000), None, 0.1, 10, sigma=0.01,
                       generator=torch.Generator().manual_seed(0))
    assert abs(float(G2.std()) - 0.01) < 5e-4


def test_foolsgold_server_step():
    G = torch.ones(10)
    S = torch.stack([torch.full((6,), 2.0), torch.full((6,), 4.0)])
    agg.foolsgold_update_(G, 6, S, torch.tensor([1.0, 0.5]), n_clients=2, eta=0.1, lr=0.1, decay=0.5)
    exp = 1 - 0.1 * (0.1 * (2 + 2) / 2 + 0.5)
    torch.testing.assert_close(G[:6], torch.full((6,), exp))
    assert torch.equal(G[6:], torch.ones(4))                  # buffers untouched (Q9)


# ------------------------------------------------------------------------------------- recorder / checkpoint
def test_csv_recorder_schema(tmp_path):
    from dba_mod_b200.utils.csv_record import Recorder
    r = Recorder()
    r.add_train(3, 1, 1, 1, 0.5, 50.0, 32, 64)
    r.add_test("global", 1, 0.4, 60.0, 6, 10)
    r.add_poison_test("global", 1, 0.4, 60.0, 6, 10)
    r.add_trigger_test("global", "combine", "", 1, 0.4, 60.0, 6, 10)
    r.add_weight_result([1, 2], [0.5, 0.5], [0.1, 0.2])
    r.add_scale(1, 12.34567); r.close_scale_row(77.12345)
    r.save_result_csv(1, True, str(tmp_path))
    rows = lambda f: [l.strip() for l in open(tmp_path / f)]   # noqa: E731
    assert rows("train_result.csv")[0] == "local_model,round,epoch,internal_epoch,average_loss,accuracy,correct_data,total_data"
    assert rows("test_result.csv")[0] == "model,epoch,average_loss,accuracy,correct_data,total_data"
    assert rows("posiontest_result.csv")[1] == "global,1,0.4,60.0,6,10"
    assert rows("poisontriggertest_result.csv")[0].startswith("model,trigger_name,trigger_value,epoch")
    assert rows("weight_result.csv") == ["1,2", "0.5,0.5", "0.1,0.2"]
    assert rows("scale_result.csv") == ["1,12.3457,77.1235"]
    # reference module writes byte-identical files from the same rows
    if not os.path.isdir(REF):
        return
    import importlib.util
    spec_ = importlib.util.spec_from_file_location("ref_csv_record", os.path.join(REF, "utils", "csv_record.py"))
    cr = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(cr)
    for lst in (cr.train_result, cr.test_result, cr.posiontest_result, cr.poisontriggertest_result, cr.weight_result,
                cr.scale_result, cr.scale_temp_one_row):
        lst.clear()
    cr.train_result.append([3, 1, 1, 1, 0.5, 50.0, 32, 64]); cr.test_result.append(["global", 1, 0.4, 60.0, 6, 10])
    cr.posiontest_result.append(["global", 1, 0.4, 60.0, 6, 10])
    cr.poisontriggertest_result.append(["global", "combine", "", 1, 0.4, 60.0, 6, 10])
    cr.add_weight_result([1, 2], [0.5, 0.5], [0.1, 0.2])
    cr.scale_temp_one_row.extend([1, 12.3457, 77.1235])
    ref_dir = tmp_path / "ref"; ref_dir.mkdir()
    cr.save_result_csv(1, True, str(ref_dir))
    for f in ["train_result.csv", "test_result.csv", "posiontest_result.csv", "poisontriggertest_result.csv",
              "weight_result.csv", "scale_result.csv"]:
        assert open(tmp_path / f).read().replace("\r\n", "\n") == open(ref_dir / f).read().replace("\r\n", "\n"), f


@pytest.mark.parametrize("kind", ["cifar", "mnist", "loan", "tiny-imagenet-200"])
def test_checkpoint_roundtrip_with_reference_state_dict(kind, tmp_path):
    """A checkpoint written by the reference's model classes loads into the flat state and comes back identical
    (key names, order, shapes, dtypes incl. int64 num_batches_tracked)."""
    from dba_mod_b200.models import initial_state
    from dba_mod_b200.utils import checkpoint as ckpt
    _ref_module("config")
    sys.path.insert(0, REF)
    try:
        for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
            del sys.modules[k]
        if kind == "cifar":
            from models.resnet_cifar import ResNet18
            m = ResNet18(name="Target", created_time="t")
        elif kind == "mnist":
            from models.MnistNet import MnistNet
            m = MnistNet(name="Target", created_time="t")
        elif kind == "loan":
            from models.loan_model import LoanNet
            m = LoanNet(name="Target", created_time="t")
        else:
            from models.resnet_tinyimagenet import resnet18
            m = resnet18(name="Target", created_time="t")
    finally:
        sys.path.remove(REF)
        for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
            del sys.modules[k]
    sd = m.state_dict()
    for k, v in sd.items():
        if v.dtype == torch.long:
            v.fill_(7)
    path = tmp_path / "model_last.pt.tar.epoch_200"
    torch.save({"state_dict": sd, "epoch": 200, "lr": 0.05}, path)
    spec, flat, cnt = initial_state(kind)
    sd2, ep, lr = ckpt.load_checkpoint(str(path))
    f, c = spec.flatten(sd2)
    back = spec.unflatten(f, c)
    assert ep == 200 and lr == 0.05 and list(back.keys()) == list(sd.keys())
    for k in sd:
        assert back[k].dtype == sd[k].dtype and torch.equal(back[k], sd[k]), k
    # and the reference model accepts what we write
    m.load_state_dict(back)


def test_schedule_segments_drop_finished_clients():
    from dba_mod_b200.fl.native_resnet import NativeResNet
    seg = NativeResNet._segments
    # ten clients, one attacker training 3x longer: the tail runs on the attacker alone
    s = seg([16] * 4 + [48] + [14] * 5)
    assert s[0] == (list(range(10)), 0, 14) and s[-1] == ([4], 16, 48)
    # nothing to gain: one segment
    assert seg([16] * 10) == [(list(range(10)), 0, 16)]
    # every step of every client is covered exactly once, segments are contiguous, rows only ever shrink
    for steps in ([5, 40, 12, 40, 9], [6, 10, 12, 14, 16, 18, 20, 24, 28, 32], [3, 1, 2], [7], [0, 4]):
        cover = [0] * len(steps)
        segs = seg(steps)
        assert segs[0][1] == 0 and segs[-1][2] == max(steps)
        for (rows, lo, hi), nxt in zip(segs, segs[1:] + [None]):
            assert rows == [r for r in range(len(steps)) if steps[r] > lo]
            if nxt is not None:
                assert nxt[1] == hi and set(nxt[0]) <= set(rows)
            for r in rows:
                cover[r] += min(hi, steps[r]) - min(lo, steps[r])
        assert cover == steps
    assert len(seg([6, 10, 12, 14, 16, 18, 20, 24, 28, 32], max_cuts=2)) <= 3
    assert seg([]) == [([], 0, 0)]


def test_space_to_depth_layout_round_trip():
    from dba_mod_b200.ops.layout import Geom, depth_to_space, from_planar, space_to_depth, to_planar
    x = torch.randn(2, 3, 5, 8, 6)
    y = space_to_depth(x)
    assert y.shape == (2, 3, 20, 4, 3) and torch.equal(depth_to_space(y), x)
    # channel p*C + c holds parity plane p = 2*(h & 1) + (w & 1)
    assert torch.equal(y[:, :, 1 * 5 + 2], x[:, :, 2, 0::2, 1::2]) and torch.equal(y[:, :, 2 * 5 + 4], x[:, :, 4, 1::2, 0::2])
    # planar-8 round trip (bf16 rounding only), halo rows stay zero
    p = to_planar(x)
    g = Geom(3, 8, 6)
    assert p.shape == (2, 1, g.rows, 8)
    torch.testing.assert_close(from_planar(p, 3, 8, 6, 5), x.bfloat16().float())


def test_native_executor_planning_on_cpu():
    """Planning logic of the native ResNet executor needs no GPU: layer table, space-to-depth blocks, evaluation
    pass sizes, slab width ladder."""
    import yaml
    from dba_mod_b200.ops import native
    if not native.available():
        pytest.skip("libdba_b200.so not built")
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import ImageData
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.models import initial_state
    raw = yaml.safe_load(open(os.path.join(ROOT, "utils", "cifar_params.yaml")))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=640, synthetic_test_size=100)
    p = Params(raw)
    data = ImageData(p, torch.device("cpu"))
    spec, _, _ = initial_state("cifar", 1)
    nx = NativeResNet(p, spec, data, torch.device("cpu"))
    assert len(nx.layers) == 20 and len(nx.blocks) == 8
    assert [l.name for l in nx.layers if l.d2s] == [f"layer{i}.0.conv1.weight" for i in (2, 3, 4)]
    assert [nx._s2d_block(b) for b in range(8)] == [False, False, True, False, True, False, True, False]
    # equal passes of at most native_eval_chunk images, multiples of 32
    assert [nx._chunk_for(n) for n in (10000, 9000, 1250, 100)] == [1000, 1000, 640, 128]
    assert all(nx._chunk_for(n) * -(-n // nx._chunk_for(n)) - n < 32 * -(-n // nx._chunk_for(n)) for n in (1250, 777, 4999))
    # width ladder: at most one third idle rows, never wider than the round
    for K in (3, 10, 80):
        for k in range(1, K + 1):
            w = NativeResNet._width(k, K)
            assert k <= w <= K and (w - k) * 3 <= w
