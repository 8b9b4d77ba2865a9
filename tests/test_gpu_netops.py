"""Op-level tests of the network kernels around the convolution, in BOTH activation modes (bf16 planar-8 / fp32
planar-4 with tf32 MMA), each against a plain PyTorch fp32 statement of the same op: weight gradients (csrc/wgrad_tc.cu),
training BatchNorm forward / backward incl. ragged batches at C = 32 ... 512 (csrc/bn_ops.cu), the fused classifier head,
max-pooling and the fused input stage (csrc/net_ops.cu)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _planar(x, act, cpad=None):
    from dba_mod_b200.ops import layout
    return layout.to_planar(x, cpad, dtype=act.dtype)


def _split(x):
    """bf16 hi/lo split copy [K, 2, C/8, R, 8] of x [K, B, C, H, W]: the tf32 mode's weight-gradient operand form."""
    from dba_mod_b200.ops import layout
    hi = x.bfloat16()
    lo = (x - hi.float()).bfloat16()
    return torch.stack([layout.to_planar(hi.float()), layout.to_planar(lo.float())], 1).contiguous()


def _unsplit(s, B, H, W):
    from dba_mod_b200.ops import layout
    return layout.from_planar(s[:, 0].contiguous(), B, H, W) + layout.from_planar(s[:, 1].contiguous(), B, H, W)


def _wgrad_operand(x, act):
    return _split(x) if act.name == "tf32" else _planar(x, act)


# ----------------------------------------------------------------------------------------------- weight gradient
WG_SHAPES = [  # K, B, H, Cin, Cout, ksize, cin_real
    (2, 5, 8, 64, 128, 3, 64),
    (1, 3, 32, 32, 32, 3, 32),           # Cout <= 32, 3x3: three vertical taps stacked along M (and the plain form)
    (2, 9, 8, 64, 32, 3, 64),
    (1, 5, 16, 32, 24, 3, 20),           # ragged real channel counts inside the padded planes
    (2, 64, 32, 32, 32, 1, 27),          # CIFAR stem as a folded 1x1 convolution over 27 (padded to 32) columns
    (1, 4, 4, 256, 256, 3, 256),
    (1, 3, 2, 512, 512, 3, 512),
    (1, 2, 32, 192, 64, 1, 147),         # Tiny-ImageNet stem: 7x7x3 = 147 columns padded to 192
    (2, 7, 16, 64, 64, 1, 64),
]


@pytest.mark.parametrize("vstack", [1, 0], ids=["vstack", "plain"])
@pytest.mark.parametrize("K,B,H,Cin,Cout,ks,cin_real", WG_SHAPES)
def test_wgrad_matches_autograd(native_lib, act, K, B, H, Cin, Cout, ks, cin_real, vstack):
    from dba_mod_b200.ops import native
    lib = native.load()
    if vstack == 0 and not (ks == 3 and Cout <= 32):
        pytest.skip("the stacked form only exists for 3x3 layers with Cout <= 32")
    lib.dba_set_wgrad_vstack(vstack)
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    # bf16 mode: operands are bf16 values; tf32 mode: arbitrary fp32 values carried as bf16 hi + lo (16 mantissa bits)
    rnd = act.round if act.name == "bf16" else (lambda t: t)
    x = rnd(torch.randn(K, B, Cin, H, H, device=DEV))
    x[:, :, cin_real:] = 0
    dy = rnd(torch.randn(K, B, Cout, H, H, device=DEV))
    ref = []
    for k in range(K):
        w = torch.zeros(Cout, Cin, ks, ks, device=DEV, requires_grad=True)
        y = F.conv2d(x[k], w, None, 1, ks // 2)
        ref.append(torch.autograd.grad(y, w, dy[k])[0])
    ref = torch.stack(ref)[:, :, :cin_real]                                  # [K, Cout, cin_real, ks, ks]
    ld = (Cout * ks * ks * cin_real + 3) // 4 * 4 + 8
    G = torch.zeros(K, ld, device=DEV)
    dyp, xp = _wgrad_operand(dy, act), _wgrad_operand(x, act)      # keep the operands alive across the launch
    try:
        native.check(lib.dba_wgrad_tc(native.ptr(dyp), native.ptr(xp), native.ptr(G), G.stride(0), 4, K, B, H, H, Cin, Cout,
                                      cin_real, Cout, ks, native.stream()))
        torch.cuda.synchronize()
    finally:
        lib.dba_set_wgrad_vstack(1)
    got = G[:, 4:4 + Cout * ks * ks * cin_real].view(K, Cout, ks, ks, cin_real).permute(0, 1, 4, 2, 3)
    scale = ref.abs().max().item()
    tol = dict(rtol=1e-4, atol=2e-5 * scale) if act.name == "tf32" else dict(rtol=act.rtol, atol=act.atol * scale)
    torch.testing.assert_close(got, ref, **tol)
    assert float(G[:, :4].abs().max()) == 0 and float(G[:, 4 + Cout * ks * ks * cin_real:].abs().max()) == 0


# ------------------------------------------------------------------------------------------------- BatchNorm
@pytest.mark.parametrize("fused", [1, 0], ids=["cluster", "twokernel"])
@pytest.mark.parametrize("C,H,B,nvalid", [(32, 8, 6, [6, 3]), (64, 16, 5, [5, 5]), (512, 2, 7, [7, 4]),
                                          (256, 4, 9, [1, 9]), (32, 32, 20, [20, 13])])
def test_batchnorm_forward_backward(native_lib, act, C, H, B, nvalid, fused):
    """BatchNorm(train) + residual + ReLU forward from the conv-epilogue statistics, and its backward, for K = 2
    clients with ragged batches (samples b >= nvalid[k] are padding).  The backward runs as one thread-block-cluster
    launch (partial sums over distributed shared memory; the last shape uses clusters of 8) and as reduce + apply."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(1)
    K = len(nvalid)
    y = act.store(torch.randn(K, B, C, H, H, device=DEV) * 1.5 + 0.3)
    res = act.store(torch.randn(K, B, C, H, H, device=DEV))
    gup = act.store(torch.randn(K, B, C, H, H, device=DEV))
    gamma, beta = torch.rand(K, C, device=DEV) + 0.5, torch.randn(K, C, device=DEV)
    # flat state rows: [gamma | beta | running_mean | running_var]
    Wf = torch.cat([gamma, beta, torch.zeros(K, C, device=DEV), torch.ones(K, C, device=DEV)], 1).contiguous()
    stats = torch.zeros(K, 2 * C, device=DEV)
    for k in range(K):
        v = y[k, :nvalid[k]]
        stats[k, 0::2] = v.sum((0, 2, 3))
        stats[k, 1::2] = (v * v).sum((0, 2, 3))
    saved = torch.zeros(K, 2 * C, device=DEV)
    nv = torch.tensor(nvalid, dtype=torch.int32, device=DEV)
    g = layout.Geom(B, H, H)
    yp, rp = _planar(y, act), _planar(res, act)
    out = torch.zeros_like(yp)
    osplit = torch.zeros(K, 2, C // 8, g.rows, 8, dtype=torch.bfloat16, device=DEV) if act.name == "tf32" else None
    native.check(lib.dba_bn_apply_split(native.ptr(yp), native.ptr(rp), native.ptr(out), None, 0, native.ptr(osplit), None,
                                        native.ptr(stats), stats.stride(0), 0, native.ptr(saved), saved.stride(0), 0,
                                        native.ptr(Wf), Wf.stride(0), 0, C, 2 * C, 3 * C, native.ptr(nv), None, K, C, B, H,
                                        H, g.rows, 1, 0.1, 1e-5, native.stream()))
    torch.cuda.synchronize()
    got = layout.from_planar(out, B, H, H)
    if osplit is not None:          # hi + lo carries the unrounded value (16 mantissa bits), `got` is on the tf32 grid
        torch.testing.assert_close(_unsplit(osplit, B, H, H), got, rtol=6e-4, atol=1e-6)    # half a tf32 ulp = 2^-11
    refs, dys, dgs, dbs = [], [], [], []
    for k in range(K):
        n = nvalid[k]
        yk = y[k, :n].clone().requires_grad_(True)
        gk, bk = gamma[k].clone().requires_grad_(True), beta[k].clone().requires_grad_(True)
        o = F.relu(F.batch_norm(yk, None, None, gk, bk, True, 0.1, 1e-5) + res[k, :n])
        d = torch.autograd.grad(o, (yk, gk, bk), gup[k, :n])
        refs.append(o.detach()); dys.append(d[0]); dgs.append(d[1]); dbs.append(d[2])
        torch.testing.assert_close(got[k, :n], act.store(o.detach()), rtol=act.rtol, atol=act.atol)
        assert float(got[k, n:].abs().max()) == 0 if n < B else True
        m = y[k, :n].mean((0, 2, 3))
        var_unb = y[k, :n].var((0, 2, 3), unbiased=True)
        torch.testing.assert_close(Wf[k, 2 * C:3 * C], 0.1 * m, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(Wf[k, 3 * C:], 0.9 + 0.1 * var_unb, rtol=1e-3, atol=1e-4)
    assert layout.halo_is_zero(out, B, H, H)
    # backward: ReLU mask from the stored activation, dy / d gamma / d beta
    sums = torch.zeros(K, 2 * C, device=DEV)
    Gf = torch.zeros(K, 4 * C, device=DEV)
    dy = torch.zeros_like(yp)
    gmask = torch.zeros_like(yp)
    dsplit = torch.zeros_like(osplit) if osplit is not None else None
    gp = _planar(gup, act)
    was = lib.dba_get_bn_fused()
    lib.dba_set_bn_fused(fused)
    try:
        native.check(lib.dba_bn_backward_split(native.ptr(gp), native.ptr(yp), native.ptr(out), native.ptr(dy),
                                               native.ptr(dsplit), native.ptr(gmask), native.ptr(sums), sums.stride(0), 0,
                                               native.ptr(saved), saved.stride(0), 0, native.ptr(Wf), Wf.stride(0),
                                               native.ptr(Gf), Gf.stride(0), 0, C, native.ptr(nv), K, C, B, H, H, g.rows,
                                               0, g.rows, native.stream()))
        torch.cuda.synchronize()
    finally:
        lib.dba_set_bn_fused(was)
    gdy, gm = layout.from_planar(dy, B, H, H), layout.from_planar(gmask, B, H, H)
    if dsplit is not None:
        torch.testing.assert_close(_unsplit(dsplit, B, H, H), gdy, rtol=6e-4, atol=1e-7)
    # bf16 mode: the ReLU mask can flip where the rounded activation is exactly 0 but the fp32 one is not, and the
    # gradient itself is stored in bf16 -> compare in aggregate there, element-wise in tf32 mode
    for k in range(K):
        n = nvalid[k]
        if act.name == "tf32":
            torch.testing.assert_close(gdy[k, :n], dys[k], rtol=2e-3, atol=2e-4)
            torch.testing.assert_close(Gf[k, :C], dgs[k], rtol=2e-3, atol=2e-3)
            torch.testing.assert_close(Gf[k, C:2 * C], dbs[k], rtol=2e-3, atol=2e-3)
            torch.testing.assert_close(gm[k, :n], gup[k, :n] * (refs[k] > 0), rtol=0, atol=0)
        else:
            cos = F.cosine_similarity(gdy[k, :n].flatten(), dys[k].flatten(), dim=0).item()
            assert cos > 0.995, cos
            torch.testing.assert_close(Gf[k, C:2 * C], dbs[k], rtol=5e-2, atol=0.3)
        assert float(gdy[k, n:].abs().max()) == 0 if n < B else True


# ------------------------------------------------------------------------------------------------ classifier head
@pytest.mark.parametrize("C,H,ncls,B", [(256, 4, 10, 11), (512, 2, 200, 5)])
def test_head_forward_backward(native_lib, act, C, H, ncls, B):
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(2)
    K = 2
    x = act.store(torch.rand(K, B, C, H, H, device=DEV))
    Wl, bl = torch.randn(K, ncls, C, device=DEV) * 0.05, torch.randn(K, ncls, device=DEV) * 0.1
    labels = torch.randint(0, ncls, (K, B), device=DEV, dtype=torch.int32)
    labels[1, B - 2:] = -1                                   # padding samples of a ragged batch
    nvalid = torch.tensor([B, B - 2], dtype=torch.int32, device=DEV)
    Wf = torch.cat([Wl.reshape(K, -1), bl], 1).contiguous()
    Gf = torch.zeros_like(Wf)
    g = layout.Geom(B, H, H)
    xp = _planar(x, act)
    dact = torch.zeros_like(xp)
    mf = torch.zeros(K, 2, device=DEV)
    mi = torch.zeros(K, 2, 2, dtype=torch.int32, device=DEV)
    nan = torch.zeros(K, dtype=torch.uint8, device=DEV)
    native.check(lib.dba_head(native.ptr(xp), native.ptr(Wf), Wf.stride(0), 0, ncls * C, native.ptr(Gf), Gf.stride(0),
                              native.ptr(dact), native.ptr(labels), 0, native.ptr(nvalid), None, native.ptr(mf),
                              native.ptr(mi), 2, native.ptr(nan), K, B, C, H, H, ncls, g.rows, native.stream()))
    torch.cuda.synchronize()
    gd = layout.from_planar(dact, B, H, H)
    for k in range(K):
        n = int(nvalid[k])
        xk = x[k, :n].clone().requires_grad_(True)
        wk, bk = Wl[k].clone().requires_grad_(True), bl[k].clone().requires_grad_(True)
        logits = F.linear(F.adaptive_avg_pool2d(xk, 1).flatten(1), wk, bk)
        loss = F.cross_entropy(logits, labels[k, :n].long())
        gx, gw, gb = torch.autograd.grad(loss, (xk, wk, bk))
        assert abs(float(mf[k, 0]) - float(loss)) < 2e-3 * max(1.0, float(loss))
        assert int(mi[k, 0, 1]) == n and int(mi[k, 0, 0]) == int((logits.argmax(1) == labels[k, :n]).sum())
        torch.testing.assert_close(Gf[k, :ncls * C].view(ncls, C), gw, rtol=2e-3, atol=1e-5)
        torch.testing.assert_close(Gf[k, ncls * C:], gb, rtol=2e-3, atol=1e-5)
        tol = dict(rtol=2e-3, atol=1e-7) if act.name == "tf32" else dict(rtol=1e-2, atol=1e-5)
        torch.testing.assert_close(gd[k, :n], gx, **tol)
        assert float(gd[k, n:].abs().max()) == 0 if n < B else True
    assert int(nan.sum()) == 0


# -------------------------------------------------------------------------------------------------- max pooling
def test_maxpool_3x3_stride2(native_lib, act):
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(3)
    K, B, C, H = 2, 3, 64, 16
    x = act.store(torch.rand(K, B, C, H, H, device=DEV))
    x[:, :, :, 2:6, 2:6] = 0.5                              # ties: the first maximum in scan order takes the gradient
    gy = act.store(torch.randn(K, B, C, H // 2, H // 2, device=DEV))
    xp, gyp = _planar(x, act), _planar(gy, act)
    gi, go = layout.Geom(B, H, H), layout.Geom(B, H // 2, H // 2)
    yp = torch.zeros(K, C // act.V, go.rows, act.V, dtype=act.dtype, device=DEV)
    gxp = torch.zeros_like(xp)
    native.check(lib.dba_maxpool_fwd(native.ptr(xp), native.ptr(yp), K, C, B, H, H, gi.rows, go.rows, native.stream()))
    native.check(lib.dba_maxpool_bwd(native.ptr(xp), native.ptr(gyp), native.ptr(gxp), K, C, B, H, H, gi.rows, go.rows,
                                     native.stream()))
    torch.cuda.synchronize()
    xr = x.reshape(K * B, C, H, H).clone().requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 2, 1)
    gxr = torch.autograd.grad(yr, xr, gy.reshape(K * B, C, H // 2, H // 2))[0]
    torch.testing.assert_close(layout.from_planar(yp, B, H // 2, H // 2).reshape(K * B, C, H // 2, H // 2), yr.detach(),
                               rtol=0, atol=0)
    torch.testing.assert_close(layout.from_planar(gxp, B, H, H).reshape(K * B, C, H, H), act.store(gxr),
                               rtol=act.rtol, atol=act.atol)


# -------------------------------------------------------------------------------------------------- input stage
@pytest.mark.parametrize("C,Hi,ks,stride,pad,KP", [(3, 32, 3, 1, 1, 32), (3, 64, 7, 2, 3, 192), (1, 28, 5, 1, 0, 32)])
def test_input_stage_gather_stamp_im2col(native_lib, act, C, Hi, ks, stride, pad, KP):
    """uint8 gather + /255 + trigger stamp (first poison_n samples) + label swap + flip + stem im2col, vs torch unfold."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(4)
    K, B, N = 2, 6, 40
    Ho = (Hi + 2 * pad - ks) // stride + 1
    data = torch.randint(0, 230, (N, Hi, Hi, C), dtype=torch.uint8, device=DEV)
    labels = torch.randint(0, 10, (N,), device=DEV)
    idx = torch.randint(0, N, (K, B), device=DEV)
    idx[1, B - 1] = -1
    trig = torch.zeros(3, Hi, Hi, dtype=torch.uint8, device=DEV)
    trig[1, 0, 0:4] = 1
    trig[2, 3, 6:10] = 1
    trig_row = torch.tensor([1, 2], device=DEV)
    poison_n = torch.tensor([2, 0], device=DEV)
    flip = torch.zeros(K, B, dtype=torch.uint8, device=DEV)
    flip[0, 1] = 1
    flip[1, 0] = 1
    g = layout.Geom(B, Ho, Ho)
    out = torch.zeros(K, KP // act.V, g.rows, act.V, dtype=act.dtype, device=DEV)
    lab = torch.zeros(K, B, dtype=torch.int32, device=DEV)
    native.check(lib.dba_input_stage_image(native.ptr(data), native.ptr(labels), native.ptr(idx), idx.stride(0),
                                           native.ptr(trig), native.ptr(trig_row), native.ptr(poison_n), native.ptr(flip),
                                           native.ptr(out), native.ptr(lab), K, B, Hi, Hi, C, ks, stride, pad, Ho, Ho, KP,
                                           7, g.rows, native.stream()))
    torch.cuda.synchronize()
    got = layout.from_planar(out, B, Ho, Ho)                                  # [K, B, KP, Ho, Ho]
    for k in range(K):
        for b in range(B):
            i = int(idx[k, b])
            if i < 0:
                assert float(got[k, b].abs().max()) == 0 and int(lab[k, b]) == -1
                continue
            img = data[i].float().div(255).permute(2, 0, 1)                   # [C, H, W]
            if int(flip[k, b]):
                img = img.flip(-1)
            poisoned = b < int(poison_n[k])
            if poisoned:
                img = torch.where(trig[int(trig_row[k])].bool()[None], torch.ones_like(img), img)
            cols = F.unfold(img[None], ks, padding=pad, stride=stride)[0]     # [C*ks*ks, Ho*Ho], order (c, kh, kw)
            cols = cols.view(C, ks * ks, Ho, Ho).permute(1, 0, 2, 3).reshape(ks * ks * C, Ho, Ho)   # (kh, kw, c)
            torch.testing.assert_close(got[k, b, :ks * ks * C], act.store(cols), rtol=0, atol=1e-6)
            assert float(got[k, b, ks * ks * C:].abs().max()) == 0
            assert int(lab[k, b]) == (7 if poisoned else int(labels[i]))
