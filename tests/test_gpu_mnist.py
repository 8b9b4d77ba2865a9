"""Native MnistNet executor (tcgen05 convolutions incl. 5x5 taps, fused pools / ReLU / bias gradients; tf32 mode =
fp32 activations + kind::tf32 + split-bf16 weight gradients, bf16 mode = the fast path) vs the batched torch formulation
(fp32) on a B200."""
import pytest
import torch
import yaml

pytestmark = pytest.mark.gpu
MODES = ["tf32", "bf16"]


def _setup(train=1500, test=700, **over):
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import ImageData
    from dba_mod_b200.models import initial_state
    raw = yaml.safe_load(open("utils/mnist_params.yaml"))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=train, synthetic_test_size=test,
               number_of_total_participants=20, **over)
    p = Params(raw)
    dev = torch.device("cuda", 0)
    data = ImageData(p, dev)
    spec, flat, cnt = initial_state("mnist", 1)
    return p, dev, data, spec, flat.to(dev)


def _schedule(p, data, names, dev):
    from dba_mod_b200.fl.schedule import build_schedule, plan_clients
    plans = plan_clients(p, data, names, epoch=1)
    return build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev).manual_seed(5))


@pytest.mark.parametrize("mode", MODES)
def test_mnist_native_eval_matches_torch(native_lib, mode):
    from dba_mod_b200.fl.native_mnist import NativeMnist
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    p, dev, data, spec, flat = _setup(compute_dtype=mode)
    tol = 2e-2 if mode == "bf16" else 3e-3
    M = 3
    g = torch.Generator(device=dev).manual_seed(3)
    W = flat[None].repeat(M, 1) * (1 + 0.1 * torch.randn(M, spec.Df, device=dev, generator=g))
    tb, nx = TorchBackend(p, spec, data, dev), NativeMnist(p, spec, data, dev)
    nx.eval_chunk = 256
    for split, trig in (("clean", 0), ("poison", 0), ("poison", 2)):
        l0, c0, n0 = tb.evaluate(W, split, trig)
        l1, c1, n1 = nx.evaluate(W, split, trig)
        assert n0 == n1
        torch.testing.assert_close(l1 / n1, l0.double() / n0, rtol=tol, atol=tol)
        assert (c1 - c0).abs().max().item() <= max(3, 0.03 * n0), (c0.tolist(), c1.tolist())
    jobs = [("clean", 0), ("poison", 0), ("poison", 1)]
    res = nx.evaluate_jobs(W[1], jobs)
    for (split, trig), (loss, corr, n) in zip(jobs, res):
        l0, c0, n0 = tb.evaluate(W[1:2], split, trig)
        assert n == n0 and abs(loss / n - float(l0[0]) / n0) < tol and abs(corr - int(c0[0])) <= max(3, 0.03 * n0)


@pytest.mark.parametrize("mode", MODES)
def test_mnist_native_gradients_match_autograd(native_lib, mode):
    """One SGD step with lr = 1, momentum 0, no weight decay: W0 - W1 is the gradient.  No BatchNorm in this network, so
    the tf32 mode tracks fp32 autograd closely (cosine > 0.998 per tensor); bf16 mode: > 0.98."""
    from dba_mod_b200.fl.native_mnist import NativeMnist
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    p, dev, data, spec, flat = _setup(train=2600, **{"0_poison_epochs": [1], "internal_epochs": 1,
                                                     "internal_poison_epochs": 1, "lr": 1.0, "poison_lr": 1.0,
                                                     "momentum": 0.0, "decay": 0.0, "batch_size": 64,
                                                     "compute_dtype": mode})
    cmin, rmax = (0.98, 0.05) if mode == "bf16" else (0.998, 0.02)
    names = [17, 3]
    sch = _schedule(p, data, names, dev)
    sch.idx[:, 1:] = -1
    sch.active[:, 1:] = False
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, 0, dtype=torch.long, device=dev)
    tb, nx = TorchBackend(p, spec, data, dev), NativeMnist(p, spec, data, dev)
    p["compact_tail"] = False
    Wt, Wn = W0.clone(), W0.clone()
    rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), False)
    rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), False)
    torch.cuda.synchronize()
    gt, gn = (W0 - Wt), (W0 - Wn)
    cos = torch.nn.functional.cosine_similarity
    report = []
    for name, e in spec.entries.items():
        a, b = gt[:, e.offset:e.offset + e.numel], gn[:, e.offset:e.offset + e.numel]
        c = cos(a, b, dim=1).min().item()
        ratio = (b.norm(dim=1) / a.norm(dim=1).clamp_min(1e-12))
        report.append((name, round(c, 4), [round(x, 3) for x in ratio.tolist()]))
        assert c > cmin, report
        assert (ratio - 1).abs().max().item() < rmax, report
    ltol = 2e-2 if mode == "bf16" else 3e-3
    torch.testing.assert_close(rn.loss_sum[:, 0], rt.loss_sum[:, 0], rtol=ltol, atol=ltol)
    torch.testing.assert_close(rn.count, rt.count)
    torch.testing.assert_close(rn.poison_count, rt.poison_count)


@pytest.mark.parametrize("mode", MODES)
def test_mnist_native_multi_step_tracks_torch(native_lib, mode):
    from dba_mod_b200.fl.native_mnist import NativeMnist
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cudnn.allow_tf32 = False
    p, dev, data, spec, flat = _setup(train=3000, **{"0_poison_epochs": [1], "internal_epochs": 1,
                                                     "internal_poison_epochs": 3, "compute_dtype": mode})
    cmin = 0.95 if mode == "bf16" else 0.99
    names = [17, 3, 5]
    sch = _schedule(p, data, names, dev)
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, 0, dtype=torch.long, device=dev)
    tb, nx = TorchBackend(p, spec, data, dev), NativeMnist(p, spec, data, dev)
    Wt, Wn = W0.clone(), W0.clone()
    rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), True)
    rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), True)
    torch.cuda.synchronize()
    dt, dn = (Wt - W0), (Wn - W0)
    cos = torch.nn.functional.cosine_similarity(dt, dn, dim=1)
    assert bool(torch.isfinite(dn).all()) and cos.min().item() > cmin, cos.tolist()
    assert ((dn.norm(dim=1) / dt.norm(dim=1)) - 1).abs().max().item() < 0.1
    cg = torch.nn.functional.cosine_similarity(rt.grad_sum, rn.grad_sum, dim=1)
    assert cg.min().item() > cmin, cg.tolist()
    torch.testing.assert_close(rn.count, rt.count)
    torch.testing.assert_close(rn.poison_count, rt.poison_count)
    assert not bool(rn.nan_flag.any())
