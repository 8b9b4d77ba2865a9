"""Flat-slab kernels vs plain PyTorch fp32/fp64 restatements (run on B200 through gpurun)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


@pytest.mark.parametrize("K,Dp,extra", [(1, 5529, 0), (3, 5529, 0), (3, 5528, 0), (10, 2797608, 4800), (4, 431080, 0)])
def test_sgd_step(native_lib, K, Dp, extra):
    from dba_mod_b200.ops import flat
    torch.manual_seed(0)
    d = _dev()
    W = torch.randn(K, Dp + extra, device=d)
    V = torch.randn(K, Dp, device=d)
    G = torch.randn(K, Dp, device=d)
    S = torch.randn(K, Dp, device=d)
    lr = torch.rand(K, device=d) * 0.1
    act = (torch.arange(K, device=d) % 3 != 1).to(torch.uint8)
    W0, V0, S0 = W.clone(), V.clone(), S.clone()
    flat.sgd_step_(W, V, G, lr, act, Dp, 0.9, 5e-4, S)
    a = act.bool()[:, None]
    g = G + 5e-4 * W0[:, :Dp]
    Vr = torch.where(a, 0.9 * V0 + g, V0)
    Wr = W0.clone()
    Wr[:, :Dp] = torch.where(a, W0[:, :Dp] - lr[:, None] * Vr, W0[:, :Dp])
    Sr = torch.where(a, S0 + G, S0)
    torch.testing.assert_close(V, Vr, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(W, Wr, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(S, Sr, rtol=1e-5, atol=1e-6)


def test_scale_and_dist(native_lib):
    from dba_mod_b200.ops import flat
    torch.manual_seed(1)
    d = _dev()
    K, D = 5, 2802410
    W, A = torch.randn(K, D, device=d), torch.randn(K, D, device=d)
    s = torch.tensor([1.0, 100.0, 30.0, 0.5, 2.0], device=d)
    ref = A + (W - A) * s[:, None]
    dist_ref = (W - A).double().norm(dim=1)
    dist = flat.rows_dist(W, A)
    torch.testing.assert_close(dist, dist_ref, rtol=1e-6, atol=1e-6)
    dist_b = flat.rows_dist(W, A[:1], n=D - 4800)
    torch.testing.assert_close(dist_b, (W[:, :D - 4800] - A[:1, :D - 4800]).double().norm(dim=1), rtol=1e-6, atol=1e-6)
    flat.scale_rows_(W, A, s)
    torch.testing.assert_close(W, ref, rtol=1e-5, atol=1e-5)


def test_fedavg(native_lib):
    from dba_mod_b200.ops import flat
    torch.manual_seed(2)
    d = _dev()
    K, D = 10, 2802410
    G = torch.randn(D, device=d)
    W = G[None] + 0.01 * torch.randn(K, D, device=d)
    ref = G + 0.01 * (W - G[None]).sum(0)
    G1 = G.clone()
    flat.fedavg_update_(G1, W, 0.01)
    torch.testing.assert_close(G1, ref, rtol=1e-5, atol=1e-6)
    # broadcast + DP noise: noise has the requested std and every row receives the new global model
    G2, W2 = G.clone(), W.clone()
    flat.fedavg_update_(G2, W2, 0.01, sigma=0.01, seed=7, bcast=True)
    noise = G2 - ref
    assert abs(float(noise.std()) - 0.01) < 5e-4 and abs(float(noise.mean())) < 1e-4
    assert bool((W2 == G2[None]).all())


def test_weiszfeld_matches_torch(native_lib):
    from dba_mod_b200.fl import aggregation as agg
    from dba_mod_b200.ops import flat
    torch.manual_seed(3)
    d = _dev()
    n, D = 10, 431080
    G = torch.randn(D, device=d)
    delta = 0.01 * torch.randn(n, D, device=d)
    delta[0] *= 100.0          # a scaled attacker
    W = G[None] + delta
    ns = torch.randint(300, 700, (n,), device=d).double()
    res = agg.geometric_median(W - G[None], ns, ns.tolist(), [n], maxiter=10)
    alphas = (ns / ns.sum()).float()
    med, dist, wv, obj, state = flat.weiszfeld(W, G, alphas, 10)
    torch.cuda.synchronize()
    assert int(state[1]) == res.iters
    torch.testing.assert_close(med, res.median, rtol=2e-3, atol=2e-5)
    torch.testing.assert_close(dist.float().cpu(), torch.tensor(res.distances), rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(wv.cpu(), torch.tensor(res.wv), rtol=1e-3, atol=1e-6)
    assert float(wv[0]) < 0.02          # the scaled attacker is down-weighted


def test_foolsgold(native_lib):
    from dba_mod_b200.fl import aggregation as agg
    from dba_mod_b200.ops import flat
    torch.manual_seed(4)
    d = _dev()
    n, L, N = 10, 2560, 100
    feat = torch.randn(n, L, device=d)
    feat[3] = feat[7] * 1.5 + 0.01 * torch.randn(L, device=d)      # two sybils
    hist = torch.zeros(N, L, dtype=torch.float64, device=d)
    ids = torch.tensor([5, 17, 33, 77, 11, 2, 9, 50, 99, 0], dtype=torch.int32, device=d)
    st = agg.FoolsGoldState(use_memory=True)
    for rnd in range(2):
        fm = st.features(ids.tolist(), feat)
        wv_ref, al_ref = agg.foolsgold_weights(fm)
        wv, al = flat.foolsgold_weights(hist, ids, feat, True)
        torch.testing.assert_close(wv, wv_ref, rtol=1e-9, atol=1e-9)
        torch.testing.assert_close(al, al_ref, rtol=1e-9, atol=1e-9)
    assert float(wv[3]) == 0.0 and float(wv[7]) == 0.0
    Dp = 2797610 - 2
    G = torch.randn(Dp + 4800, device=d)
    S = torch.randn(n, Dp, device=d)
    ref = G.clone()
    ref[:Dp] -= 0.1 * (0.1 / n * (wv.float()[:, None] * S).sum(0) + 5e-4 * ref[:Dp])
    flat.foolsgold_apply_(G, S, wv, Dp, 0.1 / n, 0.1, 5e-4)
    torch.testing.assert_close(G, ref, rtol=1e-5, atol=1e-6)
