"""Native ResNet executor (tcgen05 conv chain) vs the batched torch formulation (fp32, TF32 off) on a B200, in both
compute modes: ``tf32`` (fp32 activations, kind::tf32 MMA - the default, the reference's precision class) is held to
tight bands, ``bf16`` (the opt-in fast mode) to the bands torch's own bf16 autocast achieves on this network."""
import pytest
import torch
import yaml

pytestmark = pytest.mark.gpu
MODES = ["tf32", "bf16"]


def _setup(train=1500, test=700, **over):
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import ImageData
    from dba_mod_b200.models import initial_state
    raw = yaml.safe_load(open("utils/cifar_params.yaml"))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=train, synthetic_test_size=test,
               number_of_total_participants=20, **over)
    p = Params(raw)
    dev = torch.device("cuda", 0)
    data = ImageData(p, dev)
    spec, flat, cnt = initial_state("cifar", 1)
    return p, dev, data, spec, flat.to(dev)


@pytest.mark.parametrize("mode", MODES)
def test_native_eval_matches_torch(native_lib, mode):
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    p, dev, data, spec, flat = _setup(compute_dtype=mode)
    torch.manual_seed(0)
    M = 3
    models = flat[None].repeat(M, 1)
    models[:, :spec.Dp] += 0.02 * torch.randn(M, spec.Dp, device=dev)
    models[:, spec.Dp:] = torch.rand(M, spec.Db, device=dev) * 0.5 + 0.25     # running mean / var
    tb = TorchBackend(p, spec, data, dev)
    nx = NativeResNet(p, spec, data, dev)
    nx.eval_chunk = 256
    for split, trig in [("clean", 0), ("poison", 0), ("poison", 2)]:
        l0, c0, n0 = tb.evaluate(models, split, trig)
        l1, c1, n1 = nx.evaluate(models, split, trig)
        torch.cuda.synchronize()
        assert n0 == n1
        rel = ((l1 - l0).abs() / l0.abs()).max().item()
        assert rel < (2e-3 if mode == "tf32" else 0.03), (split, trig, l0.tolist(), l1.tolist())
        assert (c1 - c0).abs().max().item() <= (2 if mode == "tf32" else max(3, 0.03 * n0)), (c0.tolist(), c1.tolist())


def _one_step_schedule(p, data, names, dev, T=1):
    from dba_mod_b200.fl.schedule import build_schedule, plan_clients
    plans = plan_clients(p, data, names, epoch=1)
    g = torch.Generator(device=dev).manual_seed(5)
    return build_schedule(p, data, plans, 1, dev, g)


@pytest.mark.parametrize("mode", MODES)
def test_native_train_gradients_match_autograd(native_lib, mode):
    """One SGD step of K=3 clients (one of them poisoning): per-tensor gradients of the native pipeline vs TRUE fp32
    autograd (cuDNN TF32 off) of the batched torch formulation.

    A random-init ResNet with batch-statistics BatchNorm amplifies operand rounding, so the yardstick for the tf32
    mode is the reference's own precision class measured on the same batches: torch with cuDNN TF32 convolutions
    deviates from true fp32 by 1 - cos = 5e-3 (stem) ... 1e-4 (layer4) (tools/gpu_calib_modes.py,
    profiles/r2/calib_modes.txt; the native tf32 mode measures 1e-2 ... 1e-4); the native mode must stay within 5x of
    the yardstick per tensor (run-to-run atomics noise moves the stem's BatchNorm between 2x and 4.5x).  The bf16 fast mode
    is held to what torch's own bf16 autocast achieves (profiles/calib_bf16_vs_fp32.txt)."""
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    p, dev, data, spec, flat = _setup(train=1280, test=200, sampling_dirichlet=False, compute_dtype=mode,
                                      **{"0_poison_epochs": [1], "internal_epochs": 1, "internal_poison_epochs": 1,
                                         "lr": 0.0, "poison_lr": 0.0})
    names = [17, 3, 5]
    sch = _one_step_schedule(p, data, names, dev)
    assert sch.T == 1 and bool((sch.idx >= 0).all())
    K = len(names)
    torch.manual_seed(0)
    W0 = flat[None].repeat(K, 1)
    W0[:, :spec.Dp] += 0.01 * torch.randn(K, spec.Dp, device=dev)
    cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    # reference: grad_sum with lr = 0 is exactly the sum of raw gradients over the T steps
    tb = TorchBackend(p, spec, data, dev)
    Wt = W0.clone()
    rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), True)
    nx = NativeResNet(p, spec, data, dev)
    Wn = W0.clone()
    rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), True)
    torch.cuda.synchronize()
    yard = None
    if mode == "tf32":           # the reference's precision class on the same batches: cuDNN TF32 convolutions
        torch.backends.cudnn.allow_tf32 = True
        yard = TorchBackend(p, spec, data, dev).train(sch, W0.clone(), cnt.clone(), W0.clone(), True).grad_sum
        torch.backends.cudnn.allow_tf32 = False
    cosf = torch.nn.functional.cosine_similarity
    bad = []
    for name in spec.param_names:
        e = spec.entries[name]
        a = rt.grad_sum[:, e.offset:e.offset + e.numel]
        b = rn.grad_sum[:, e.offset:e.offset + e.numel]
        cos = cosf(a, b, dim=1).min().item()
        ratio = (b.norm(dim=1) / a.norm(dim=1).clamp(min=1e-12))
        dev_r = (ratio - 1).abs().max().item()
        # BatchNorm gamma / beta gradients are 32..256-element sums with heavy cancellation -> looser band
        small = len(e.shape) == 1 and name != "linear.bias"
        if mode == "tf32":
            y = yard[:, e.offset:e.offset + e.numel]
            err_y = 1.0 - cosf(a, y, dim=1).min().item()
            dev_y = ((y.norm(dim=1) / a.norm(dim=1).clamp(min=1e-12)) - 1).abs().max().item()
            # the yardstick itself moves run to run (cuDNN picks TF32 algorithms non-deterministically: the stem
            # BatchNorm's err_y was seen between 3e-3 and 9e-3 on identical inputs), so the per-channel-sum tensors also
            # get an absolute floor at the calibrated worst case of the native mode (cos 0.9803, profiles/r2/calib_modes.txt)
            band = max(5.0 * err_y + 5e-4, 0.025 if small else 0.0)
            # norm ratio: the stem BatchNorm's gamma gradient (32 per-channel sums with heavy cancellation) sits at
            # 1.05-1.065 against fp32 on these inputs in every run, while the yardstick's own deviation moves
            # between 0.005 and 0.02 -> absolute allowance 0.08 for the per-channel-sum tensors
            ok = (1.0 - cos) <= band and dev_r <= 2.5 * dev_y + (0.08 if small else 0.015)
        else:                    # bf16 activations and gradients in HBM (profiles/calib_bf16_vs_fp32.txt)
            cmin, rmax = (0.75, 0.4) if small else (0.90, 0.1)
            ok = cos >= cmin and dev_r <= rmax
        if not ok:
            bad.append((name, round(cos, 4), [round(float(v), 3) for v in ratio]))
    assert not bad, bad
    # metrics of the step and BatchNorm running statistics
    torch.testing.assert_close(rn.count, rt.count)
    tol = dict(rtol=3e-3, atol=3e-3) if mode == "tf32" else dict(rtol=3e-2, atol=3e-2)
    torch.testing.assert_close(rn.loss_sum, rt.loss_sum, **tol)
    tol = dict(rtol=5e-3, atol=5e-4) if mode == "tf32" else dict(rtol=5e-2, atol=5e-3)
    torch.testing.assert_close(Wn[:, spec.Dp:], Wt[:, spec.Dp:], **tol)


@pytest.mark.parametrize("mode", MODES)
def test_native_train_multi_step_tracks_torch(native_lib, mode):
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    p, dev, data, spec, flat = _setup(train=1500, test=200, compute_dtype=mode,
                                      **{"0_poison_epochs": [1], "internal_epochs": 2, "internal_poison_epochs": 3,
                                         "lr": 0.05})
    names = [17, 3]
    sch = _one_step_schedule(p, data, names, dev)
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    tb = TorchBackend(p, spec, data, dev)
    nx = NativeResNet(p, spec, data, dev)
    Wt, Wn, ct, cn = W0.clone(), W0.clone(), cnt.clone(), cnt.clone()
    rt = tb.train(sch, Wt, ct, W0.clone(), False)
    rn = nx.train(sch, Wn, cn, W0.clone(), False)
    torch.cuda.synchronize()
    assert torch.equal(ct, cn)
    dt, dn = (Wt - W0)[:, :spec.Dp], (Wn - W0)[:, :spec.Dp]
    cos = torch.nn.functional.cosine_similarity(dt, dn, dim=1)
    # a few dozen SGD steps of a random-init network amplify rounding differences (lr 0.05, momentum 0.9)
    cmin, rmax = (0.90, 0.06) if mode == "tf32" else (0.8, 0.15)
    assert bool(torch.isfinite(dn).all()) and cos.min().item() > cmin, cos.tolist()
    assert ((dn.norm(dim=1) / dt.norm(dim=1)) - 1).abs().max().item() < rmax
    torch.testing.assert_close(rn.count, rt.count)
    torch.testing.assert_close(rn.poison_count, rt.poison_count)


def test_tail_compaction_matches_full_width_run(native_lib):
    """An attacker that trains 3x longer than its benign neighbours finishes on a narrower slab.

    (a) bookkeeping: with a deterministic stand-in for the step body, the segmented run equals the full-width run
    bit for bit (state rows, momentum, counters and metrics are carried across slabs);  (b) real kernels: the
    segmented run is as close to a full-width run as two full-width runs are to each other (fp32 atomics in the
    BatchNorm statistics make a random-init bf16 ResNet visibly non-reproducible run to run)."""
    from dba_mod_b200.fl.native_resnet import NativeResNet
    over = {"0_poison_epochs": [1], "internal_epochs": 1, "internal_poison_epochs": 3, "lr": 0.02,
            "compute_dtype": "bf16"}
    p, dev, data, spec, flat = _setup(train=1500, test=200, **over)
    names = [17, 3, 5]
    sch = _one_step_schedule(p, data, names, dev)
    assert len(NativeResNet._segments(list(sch.steps_host))) > 1
    K = len(names)
    W0 = flat[None].repeat(K, 1)

    def fake_step(K_, t, want_gs, with_flip, use_dist=False):
        act = t["st_active"].float()[:, None]
        t["V"][:, :8] += act * (1.0 + t["st_lr"][:, None])
        t["W"][:, :8] += act * (t["st_idx"].clamp(min=0).sum(1, keepdim=True).float() * 1e-6 + t["V"][:, :8])
        t["gs"][:, :4] += act * t["st_nvalid"].float()[:, None]
        t["mf"][:, 0] += act[:, 0]
        t["mi"][:, 0, 0] += t["st_nvalid"] * t["st_active"].int()
        t["cnt"] += t["st_active"].long()[:, None]

    def run(compact, fake):
        p["compact_tail"], p["cuda_graphs"] = compact, not fake
        nx = NativeResNet(p, spec, data, dev)
        if fake:
            nx._step_body = fake_step
        W, c = W0.clone(), torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
        r = nx.train(sch, W, c, W0.clone(), True)
        torch.cuda.synchronize()
        return W, c, r

    (Wa, ca, ra), (Wb, cb, rb) = run(False, True), run(True, True)
    assert torch.equal(Wa, Wb) and torch.equal(ca, cb) and torch.equal(ra.grad_sum, rb.grad_sum)
    assert torch.equal(ra.loss_sum, rb.loss_sum) and torch.equal(ra.count, rb.count)
    assert ca[:, 0].tolist() == list(sch.steps_host)

    (W1, c1, r1), (W2, c2, r2), (W3, c3, r3) = run(False, False), run(False, False), run(True, False)
    assert torch.equal(c1, c3)
    torch.testing.assert_close(r1.count, r3.count)
    torch.testing.assert_close(r1.poison_count, r3.poison_count)
    d1, d2, d3 = ((w - W0)[:, :spec.Dp] for w in (W1, W2, W3))
    cos = torch.nn.functional.cosine_similarity
    floor, got = cos(d1, d2, dim=1), cos(d1, d3, dim=1)
    assert bool((got > floor - 0.1).all()) and got.min().item() > 0.7, (floor.tolist(), got.tolist())
    torch.testing.assert_close(r1.loss_sum, r3.loss_sum, rtol=5e-2, atol=5e-2)
    assert ((d3.norm(dim=1) / d1.norm(dim=1)) - 1).abs().max().item() < 0.1


def test_native_eval_pairs_and_jobs_match_plain_eval(native_lib):
    """Batched (model, input-set) pairs and one-model-many-input-sets give the same sums as one evaluate() per job."""
    from dba_mod_b200.fl.native_resnet import NativeResNet
    p, dev, data, spec, flat = _setup()
    torch.manual_seed(1)
    M = 3
    models = flat[None].repeat(M, 1)
    models[:, :spec.Dp] += 0.02 * torch.randn(M, spec.Dp, device=dev)
    models[:, spec.Dp:] = torch.rand(M, spec.Db, device=dev) * 0.5 + 0.25
    nx = NativeResNet(p, spec, data, dev)
    nx.eval_chunk = 256
    jobs = [("clean", 0), ("poison", 0), ("poison", 3)]
    plain = []
    for m, (sp, tr) in zip(range(M), jobs):
        l, c, n = nx.evaluate(models[m:m + 1], sp, tr)
        plain.append((float(l[0]), int(c[0]), n))
    pairs = nx.evaluate_pairs(models, jobs)
    for a, b in zip(plain, pairs):
        assert a[2] == b[2] and abs(a[1] - b[1]) <= 1 and abs(a[0] - b[0]) / abs(a[0]) < 2e-3, (plain, pairs)
    one = nx.evaluate_jobs(models[1], jobs)
    for (sp, tr), b in zip(jobs, one):
        l, c, n = nx.evaluate(models[1:2], sp, tr)
        assert n == b[2] and abs(int(c[0]) - b[1]) <= 1 and abs(float(l[0]) - b[0]) / abs(b[0]) < 2e-3
    # rank sharding partitions the split
    parts = [nx.evaluate_jobs(models[0], [("poison", 1)], shard=(r, 3))[0] for r in range(3)]
    whole = nx.evaluate_jobs(models[0], [("poison", 1)])[0]
    assert sum(x[2] for x in parts) == whole[2] and abs(sum(x[1] for x in parts) - whole[1]) <= 1


def _setup_tiny(train=640, test=200, **over):
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import ImageData
    from dba_mod_b200.models import initial_state
    raw = yaml.safe_load(open("utils/tiny_params.yaml"))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=train, synthetic_test_size=test,
               number_of_total_participants=10, sampling_dirichlet=False, batch_size=32, **over)
    p = Params(raw)
    dev = torch.device("cuda", 0)
    data = ImageData(p, dev)
    spec, flat, cnt = initial_state("tiny-imagenet-200", 1)
    return p, dev, data, spec, flat.to(dev)


@pytest.mark.parametrize("mode", MODES)
def test_tiny_native_eval_and_train(native_lib, mode):
    """Tiny-ImageNet ResNet-18 (7x7 stride-2 stem as a tensor-core GEMM over im2col'd patches, max-pool, 64..512
    channels, 200-way head) on the native executor vs the batched torch formulation."""
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    p, dev, data, spec, flat = _setup_tiny(compute_dtype=mode,
                                           **{"0_poison_epochs": [1], "internal_epochs": 1, "internal_poison_epochs": 1,
                                              "lr": 0.0, "poison_lr": 0.0})
    torch.manual_seed(0)
    M = 2
    models = flat[None].repeat(M, 1)
    models[:, :spec.Dp] += 0.01 * torch.randn(M, spec.Dp, device=dev)
    models[:, spec.Dp:] = torch.rand(M, spec.Db, device=dev) * 0.5 + 0.25
    tb = TorchBackend(p, spec, data, dev)
    nx = NativeResNet(p, spec, data, dev)
    nx.eval_chunk = 100
    for split, trig in [("clean", 0), ("poison", 1)]:
        l0, c0, n0 = tb.evaluate(models, split, trig)
        l1, c1, n1 = nx.evaluate(models, split, trig)
        assert n0 == n1 and ((l1 - l0).abs() / l0.abs()).max().item() < (2e-3 if mode == "tf32" else 0.03), \
            (l0.tolist(), l1.tolist())
    names = [0, 3]
    sch = _one_step_schedule(p, data, names, dev)
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    W0[:, :spec.Dp] += 0.01 * torch.randn(K, spec.Dp, device=dev)
    cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    Wt, Wn = W0.clone(), W0.clone()
    rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), True)
    rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), True)
    torch.cuda.synchronize()
    torch.testing.assert_close(rn.count, rt.count)
    torch.testing.assert_close(rn.loss_sum, rt.loss_sum, **(dict(rtol=2e-3, atol=2e-3) if mode == "tf32"
                                                             else dict(rtol=3e-2, atol=3e-2)))
    bad = []
    cmin, rmax = (0.97, 0.05) if mode == "tf32" else (0.85, 0.15)
    for name in spec.param_names:
        e = spec.entries[name]
        if len(e.shape) == 1 and name != "fc.bias":
            continue
        a = rt.grad_sum[:, e.offset:e.offset + e.numel]
        b = rn.grad_sum[:, e.offset:e.offset + e.numel]
        cos = torch.nn.functional.cosine_similarity(a, b, dim=1).min().item()
        ratio = (b.norm(dim=1) / a.norm(dim=1).clamp(min=1e-12))
        if not (cos >= cmin) or not ((ratio - 1).abs().max().item() <= rmax):
            bad.append((name, round(cos, 4), [round(float(v), 3) for v in ratio]))
    assert not bad, bad


def test_native_distance_regularised_poison_loss(native_lib):
    """alpha_loss = 0.5: gradient of alpha*CE + (1-alpha)*||w - w_anchor|| on the native path vs autograd."""
    from dba_mod_b200.fl.native_resnet import NativeResNet
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    p, dev, data, spec, flat = _setup(train=1280, test=200, sampling_dirichlet=False, alpha_loss=0.5,
                                      **{"0_poison_epochs": [1], "internal_epochs": 1, "internal_poison_epochs": 2,
                                         "lr": 0.01, "poison_lr": 0.01})
    names = [17, 3]
    sch = _one_step_schedule(p, data, names, dev)
    assert sch.T == 2
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    tb, nx = TorchBackend(p, spec, data, dev), NativeResNet(p, spec, data, dev)
    Wt, Wn = W0.clone(), W0.clone()
    rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), True)
    rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), True)
    torch.cuda.synchronize()
    torch.testing.assert_close(rn.loss_sum, rt.loss_sum, rtol=3e-2, atol=3e-2)
    dt, dn = (Wt - W0)[:, :spec.Dp], (Wn - W0)[:, :spec.Dp]
    cos = torch.nn.functional.cosine_similarity(dt, dn, dim=1)
    assert cos.min().item() > 0.9 and ((dn.norm(dim=1) / dt.norm(dim=1)) - 1).abs().max().item() < 0.1, cos.tolist()
