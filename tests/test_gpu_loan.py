"""LOAN persistent-CTA trainer / evaluator vs the batched torch formulation."""
import pytest
import torch
import yaml

pytestmark = pytest.mark.gpu


def _setup(**over):
    from dba_mod_b200.config import Params
    from dba_mod_b200.data import LoanData
    from dba_mod_b200.models import initial_state
    raw = yaml.safe_load(open("utils/loan_params.yaml"))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=8000, **over)
    p = Params(raw)
    dev = torch.device("cuda", 0)
    data = LoanData(p, dev)
    spec, flat, cnt = initial_state("loan", 1)
    return p, dev, data, spec, flat.to(dev)


def test_loan_eval_matches_torch(native_lib):
    from dba_mod_b200.fl.native_loan import NativeLoan
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cuda.matmul.allow_tf32 = False
    p, dev, data, spec, flat = _setup()
    torch.manual_seed(0)
    models = flat[None].repeat(4, 1) + 0.05 * torch.randn(4, spec.Df, device=dev)
    tb, nx = TorchBackend(p, spec, data, dev), NativeLoan(p, spec, data, dev)
    for split, trig in [("clean", 0), ("poison", 0), ("poison", 2)]:
        l0, c0, n0 = tb.evaluate(models, split, trig)
        l1, c1, n1 = nx.evaluate(models, split, trig)
        assert n0 == n1
        torch.testing.assert_close(l1, l0, rtol=1e-4, atol=1e-2)
        assert (c1 - c0).abs().max().item() <= 2


def _train_both(dropout):
    from dba_mod_b200.fl.native_loan import NativeLoan
    from dba_mod_b200.fl.schedule import build_schedule, plan_clients
    from dba_mod_b200.fl.torch_backend import TorchBackend
    torch.backends.cuda.matmul.allow_tf32 = False
    p, dev, data, spec, flat = _setup(**{"0_poison_epochs": [1], "internal_poison_epochs": 3, "lr": 0.01,
                                         "poison_lr": 0.01})
    names = ["CT", "CA", "NY", "TX", "FL", "IL"]
    plans = plan_clients(p, data, names, 1)
    sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev))
    K = len(names)
    W0 = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, 0, dtype=torch.long, device=dev)
    tb, nx = TorchBackend(p, spec, data, dev), NativeLoan(p, spec, data, dev)
    if not dropout:
        tb.net.dropout_p = 0.0
        nx.dropout = False
    Wt, Wn = W0.clone(), W0.clone()
    rt = tb.train(sch, Wt, cnt, W0.clone(), True, generator=torch.Generator(device=dev).manual_seed(3))
    rn = nx.train(sch, Wn, cnt, W0.clone(), True)
    torch.cuda.synchronize()
    return W0, Wt, Wn, rt, rn


def test_loan_train_matches_torch_with_dropout_off(native_lib):
    """Same batches, same trigger, no dropout -> the persistent-CTA trainer must reproduce the torch path (fp32)."""
    W0, Wt, Wn, rt, rn = _train_both(dropout=False)
    torch.testing.assert_close(rn.count, rt.count)
    torch.testing.assert_close(rn.poison_count, rt.poison_count)
    torch.testing.assert_close(rn.correct, rt.correct)
    torch.testing.assert_close(rn.loss_sum, rt.loss_sum, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(Wn, Wt, rtol=2e-3, atol=2e-5)
    torch.testing.assert_close(rn.grad_sum, rt.grad_sum, rtol=5e-3, atol=1e-4)
    assert not bool(rn.nan_flag.any())


def test_loan_train_with_dropout_is_statistically_alike(native_lib):
    """With Dropout(0.5) the two paths draw different masks; losses and update sizes agree within sampling noise."""
    W0, Wt, Wn, rt, rn = _train_both(dropout=True)
    torch.testing.assert_close(rn.count, rt.count)
    lt = rt.loss_sum.sum(1) / rt.count.sum(1).clamp(min=1)
    ln = rn.loss_sum.sum(1) / rn.count.sum(1).clamp(min=1)
    assert ((ln - lt).abs() / lt).max().item() < 0.15, (lt.tolist(), ln.tolist())
    dt, dn = (Wt - W0).norm(dim=1), (Wn - W0).norm(dim=1)
    assert ((dn / dt) - 1).abs().median().item() < 0.15, (dt.tolist(), dn.tolist())
