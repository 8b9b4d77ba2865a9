"""tcgen05 implicit-GEMM convolution vs torch fp32 conv2d on the same operand values (rounded to what the tensor core
reads: bf16, or tf32 = fp32 with the low 13 mantissa bits dropped), in both activation modes."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

SHAPES = [  # K, B, H, W, Cin, Cout, ksize, sub
    (2, 3, 8, 8, 32, 32, 3, 1),
    (1, 2, 32, 32, 32, 64, 3, 2),
    (3, 5, 4, 4, 256, 256, 3, 1),
    (2, 4, 16, 16, 64, 128, 1, 2),
    (2, 7, 16, 16, 64, 64, 3, 1),
    (1, 3, 8, 8, 128, 128, 3, 1),
    (1, 2, 2, 2, 512, 512, 3, 1),
    (2, 64, 32, 32, 32, 32, 3, 1),
    (2, 5, 32, 32, 32, 32, 1, 1),          # CIFAR stem: 3x3x3 patch folded to a 1x1 convolution over 32 columns
    (1, 3, 16, 16, 16, 32, 3, 1),          # Cin = 16 (one bf16 k-step / two tf32 k-steps)
    (1, 3, 32, 32, 192, 64, 1, 1),         # Tiny-ImageNet stem: 7x7x3 patch folded to 192 columns
    (2, 3, 2, 2, 512, 512, 1, 2),          # Tiny-ImageNet layer4 shortcut
]


def _ref(x, w, ksize, sub, act):
    K = x.shape[0]
    xb, wb = act.round(x), act.round(w)
    return torch.stack([F.conv2d(xb[k], wb[k], None, sub, ksize // 2) for k in range(K)])


@pytest.mark.parametrize("K,B,H,W,Cin,Cout,ksize,sub", SHAPES)
def test_conv_raw(native_lib, act, K, B, H, W, Cin, Cout, ksize, sub):
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    d = torch.device("cuda", 0)
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, ksize, ksize, device=d) / (Cin * ksize * ksize) ** 0.5
    y = conv.conv_tc(layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w, dtype=act.dtype), K, B, H, W, ksize, sub)
    torch.cuda.synchronize()
    got = layout.from_planar(y, B, H // sub, W // sub)
    ref = _ref(x, w, ksize, sub, act)
    torch.testing.assert_close(got, ref, rtol=act.rtol, atol=act.atol)
    assert layout.halo_is_zero(y, B, H // sub, W // sub)


def test_conv_eval_epilogue_and_shared_input(native_lib, act):
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(1)
    d = torch.device("cuda", 0)
    K, B, H, W, Cin, Cout = 3, 6, 16, 16, 64, 64
    x = torch.randn(1, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5
    res = torch.randn(K, B, Cout, H, W, device=d)
    scale, shift = torch.rand(K, Cout, device=d) + 0.5, torch.randn(K, Cout, device=d)
    y = conv.conv_tc(layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w, dtype=act.dtype), K, B, H, W, 3, 1, relu=True,
                     res=layout.to_planar(res, dtype=act.dtype), scale=scale, shift=shift, x_shared=True)
    got = layout.from_planar(y, B, H, W)
    ref = _ref(x.expand(K, -1, -1, -1, -1), w, 3, 1, act)
    ref = F.relu(ref * scale[:, None, :, None, None] + shift[:, None, :, None, None] + act.store(res))
    torch.testing.assert_close(got, ref, rtol=act.rtol, atol=1.5 * act.atol)


@pytest.mark.parametrize("K,B,H,Cin,Cout,with_act,with_res", [(2, 3, 8, 32, 32, True, False), (1, 5, 4, 256, 256, True, True),
                                                              (2, 4, 16, 64, 64, False, True), (1, 64, 32, 32, 32, True, True)])
def test_dgrad_epilogue_accumulates_batchnorm_backward_sums(native_lib, act, K, B, H, Cin, Cout, with_act, with_res):
    """dba_conv_tc_dgrad_bn: g = conv(dY, Wt) (+ Res) as usual, and sums[k][c] = (sum g', sum g' * xhat) of the BatchNorm
    that consumes g (g' = g where the layer's ReLU output is > 0), compared against the same sums formed from the
    kernel's own stored g."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(7)
    d = torch.device("cuda", 0)
    x = torch.randn(K, B, Cin, H, H, device=d)
    w = torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5
    res = torch.randn(K, B, Cout, H, H, device=d) if with_res else None
    bny = torch.randn(K, B, Cout, H, H, device=d) * 1.3 + 0.2
    bna = F.relu(torch.randn(K, B, Cout, H, H, device=d)) if with_act else None
    saved = torch.stack([torch.randn(K, Cout, device=d) * 0.3, torch.rand(K, Cout, device=d) + 0.5], -1).contiguous()
    sums = torch.zeros(K, Cout, 2, device=d)
    xp, wp = layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w, dtype=act.dtype)
    rp = layout.to_planar(res, dtype=act.dtype) if with_res else None
    yp = layout.to_planar(bny, dtype=act.dtype)
    ap = layout.to_planar(bna, dtype=act.dtype) if with_act else None
    out = torch.zeros(K, Cout // act.V, layout.Geom(B, H, H).rows, act.V, dtype=act.dtype, device=d)
    native.set_act_dtype(act.name)
    native.check(lib.dba_conv_tc_dgrad_bn(native.ptr(xp), native.ptr(wp), wp[0].numel(), native.ptr(out), native.ptr(rp),
                                          native.ptr(yp), native.ptr(ap), native.ptr(saved), saved.stride(0),
                                          native.ptr(sums), sums.stride(0), K, B, H, H, Cin, Cout, 3, native.stream()))
    torch.cuda.synchronize()
    got = layout.from_planar(out, B, H, H)
    ref = _ref(x, w, 3, 1, act) + (act.store(res) if with_res else 0)
    torch.testing.assert_close(got, ref, rtol=act.rtol, atol=1.5 * act.atol)
    ys = layout.from_planar(yp, B, H, H)
    gm = got * (layout.from_planar(ap, B, H, H) > 0) if with_act else got
    mean, rstd = saved[..., 0][:, None, :, None, None], saved[..., 1][:, None, :, None, None]
    torch.testing.assert_close(sums[..., 0], gm.sum(dim=(1, 3, 4)), rtol=2e-3, atol=2e-2)
    torch.testing.assert_close(sums[..., 1], (gm * (ys - mean) * rstd).sum(dim=(1, 3, 4)), rtol=2e-3, atol=5e-2)
    assert layout.halo_is_zero(out, B, H, H)


def test_conv_stats(native_lib, act):
    from dba_mod_b200.ops import conv, layout
    torch.manual_seed(2)
    d = torch.device("cuda", 0)
    K, B, H, W, Cin, Cout = 2, 9, 8, 8, 64, 128
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5
    stats = torch.zeros(K, Cout, 2, device=d)
    y = conv.conv_tc(layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w, dtype=act.dtype), K, B, H, W, 3, 1, stats=stats)
    got = layout.from_planar(y, B, H, W)
    s1 = got.sum(dim=(1, 3, 4))
    s2 = (got * got).sum(dim=(1, 3, 4))
    torch.testing.assert_close(stats[..., 0], s1, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(stats[..., 1], s2, rtol=1e-3, atol=1e-2)


S2D_SHAPES = [  # K, B, H, W (input grid), Cin, Cout, ksize
    (2, 3, 32, 32, 32, 64, 3),
    (1, 5, 16, 16, 64, 128, 3),
    (2, 4, 8, 8, 128, 256, 3),
    (2, 3, 16, 16, 64, 128, 1),
    (1, 70, 32, 32, 32, 64, 3),
]


@pytest.mark.parametrize("K,B,H,W,Cin,Cout,ksize", S2D_SHAPES)
def test_conv_stride2_from_space_to_depth_input(native_lib, act, K, B, H, W, Cin, Cout, ksize):
    """Stride-2 convolutions computed on the output grid only, from the parity-plane form of their input."""
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(2)
    d = torch.device("cuda", 0)
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, ksize, ksize, device=d) / (Cin * ksize * ksize) ** 0.5
    xs = layout.to_planar(layout.space_to_depth(x), dtype=act.dtype)
    stats = torch.zeros(K, Cout, 2, device=d)
    y = conv.conv_tc_s2d(xs, layout.pack_conv_weight(w, dtype=act.dtype), K, B, H // 2, W // 2, ksize, in_s2d=True, stats=stats)
    torch.cuda.synchronize()
    got = layout.from_planar(y, B, H // 2, W // 2)
    ref = _ref(x, w, ksize, 2, act)
    torch.testing.assert_close(got, ref, rtol=act.rtol, atol=act.atol)
    assert layout.halo_is_zero(y, B, H // 2, W // 2)
    torch.testing.assert_close(stats[..., 0], got.sum((1, 3, 4)), rtol=1e-3, atol=2e-2)


def test_conv_space_to_depth_output_with_residual(native_lib, act):
    """The producer side: folded-BN + residual + ReLU epilogue storing the parity-plane form, then a stride-2
    convolution consuming it equals the plain two-layer computation."""
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(3)
    d = torch.device("cuda", 0)
    K, B, H, W, C, C2 = 2, 5, 16, 16, 64, 128
    x = torch.randn(K, B, C, H, W, device=d)
    w1 = torch.randn(K, C, C, 3, 3, device=d) / (C * 9) ** 0.5
    w2 = torch.randn(K, C2, C, 3, 3, device=d) / (C * 9) ** 0.5
    res = torch.randn(K, B, C, H, W, device=d)
    scale, shift = torch.rand(K, C, device=d) + 0.5, torch.randn(K, C, device=d)
    y1 = conv.conv_tc_s2d(layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w1, dtype=act.dtype), K, B, H, W, 3, out_s2d=True, relu=True,
                          res=layout.to_planar(res, dtype=act.dtype), scale=scale, shift=shift)
    ref1 = F.relu(_ref(x, w1, 3, 1, act) * scale[:, None, :, None, None] + shift[:, None, :, None, None]
                  + act.store(res))
    got1 = layout.depth_to_space(layout.from_planar(y1, B, H // 2, W // 2))
    torch.testing.assert_close(got1, ref1, rtol=act.rtol, atol=1.5 * act.atol)
    assert layout.halo_is_zero(y1, B, H // 2, W // 2)
    y2 = conv.conv_tc_s2d(y1, layout.pack_conv_weight(w2, dtype=act.dtype), K, B, H // 2, W // 2, 3, in_s2d=True)
    ref2 = _ref(got1, w2, 3, 2, act)
    torch.testing.assert_close(layout.from_planar(y2, B, H // 2, W // 2), ref2, rtol=act.rtol, atol=1.5 * act.atol)


def _pack_with_kernel(w, mode, act):
    """Runs dba_pack_weights on OHWI-flattened weights w [K, Cout, Cin, 3, 3]; returns (pack, fwd_off, t_off)."""
    from dba_mod_b200.ops import native
    lib = native.load()
    K, Co, Ci = w.shape[:3]
    flat = w.permute(0, 1, 3, 4, 2).reshape(K, -1).contiguous().float()             # [Cout][taps][Cin] per client
    n_f, n_t = 9 * Ci * Co, (16 if mode == 3 else 9) * Ci * Co
    table = torch.tensor([[0, Co, Ci, 9, 0, n_f, Ci, Co, mode]], dtype=torch.int32, device=w.device)
    pack = torch.zeros(K, n_f + n_t, dtype=act.dtype, device=w.device)
    native.check(lib.dba_pack_weights(native.ptr(flat), flat.stride(0), native.ptr(table), 1, K, native.ptr(pack),
                                      pack.stride(0), native.stream()))
    return pack, 0, n_f


@pytest.mark.parametrize("K,B,H,Cin,Cout", [(2, 3, 16, 32, 64), (1, 5, 8, 64, 128), (2, 2, 4, 128, 256)])
def test_stride2_backward_on_the_output_grid(native_lib, act, K, B, H, Cin, Cout):
    """3x3 stride-2 convolution, backward without the full grid: the input gradient as a 2x2-tap convolution of the
    compact dy with depth-to-space store, the weight gradient against the space-to-depth input - both vs autograd."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(4)
    d = torch.device("cuda", 0)
    Ho = H // 2
    x = act.round(torch.randn(K, B, Cin, H, H, device=d))
    w = (act.round(torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5))
    dy = act.round(torch.randn(K, B, Cout, Ho, Ho, device=d))
    ref_dx, ref_dw = [], []
    for k in range(K):
        xk, wk = x[k].clone().requires_grad_(True), w[k].clone().requires_grad_(True)
        y = F.conv2d(xk, wk, None, 2, 1)
        gx, gw = torch.autograd.grad(y, (xk, wk), dy[k])
        ref_dx.append(gx); ref_dw.append(gw)
    ref_dx, ref_dw = torch.stack(ref_dx), torch.stack(ref_dw)
    pack, _, t_off = _pack_with_kernel(w, 3, act)
    dyp = layout.to_planar(dy, dtype=act.dtype)
    # input gradient
    go = layout.Geom(B, H, H)
    dx = torch.zeros(K, Cin // act.V, go.rows, act.V, dtype=act.dtype, device=d)
    wd = pack[:, t_off:]
    native.check(lib.dba_conv_tc_ex(native.ptr(dyp), native.ptr(wd), pack.stride(0), native.ptr(dx), None, None, None, 0,
                                    None, 0, K, B, Ho, Ho, Cout, 4 * Cin, 2, 1, 0, 0, 0, 3, native.stream()))
    torch.cuda.synchronize()
    got_dx = layout.from_planar(dx, B, H, H)
    torch.testing.assert_close(got_dx, ref_dx, rtol=act.rtol, atol=1.5 * act.atol)
    assert layout.halo_is_zero(dx, B, H, H)
    # weight gradient
    xs = layout.to_planar(layout.space_to_depth(x), dtype=act.dtype)
    G = torch.zeros(K, 9 * Cin * Cout, device=d)
    if act.name == "tf32":      # weight-gradient operands are the bf16 hi/lo split copies in this mode
        def split(t):
            hi = t.bfloat16()
            return torch.stack([layout.to_planar(hi.float()), layout.to_planar((t - hi.float()).bfloat16().float())],
                               1).contiguous()
        dyp, xs = split(dy), split(layout.space_to_depth(x))
    native.check(lib.dba_wgrad_tc_s2d(native.ptr(dyp), native.ptr(xs), native.ptr(G), G.stride(0), 0, K, B, Ho, Ho, Cin,
                                      Cout, Cin, Cout, native.stream()))
    torch.cuda.synchronize()
    got_dw = G.view(K, Cout, 3, 3, Cin).permute(0, 1, 4, 2, 3)
    scale = ref_dw.abs().max().item()
    torch.testing.assert_close(got_dw, ref_dw, rtol=2e-2, atol=act.atol * scale)


def test_shortcut_input_gradient_added_in_place_at_even_pixels(native_lib, act):
    """1x1 stride-2 convolution: dx[2h, 2w] += W^T dy[h, w] on the full-resolution gradient, other pixels untouched."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(5)
    d = torch.device("cuda", 0)
    K, B, H, Cin, Cout = 2, 3, 16, 32, 64
    Ho = H // 2
    w = (act.round(torch.randn(K, Cout, Cin, 1, 1, device=d) / Cin ** 0.5))
    dy = act.round(torch.randn(K, B, Cout, Ho, Ho, device=d))
    base = act.round(torch.randn(K, B, Cin, H, H, device=d))
    ref = base.clone()
    ref[:, :, :, 0::2, 0::2] += torch.einsum("koc,kbohw->kbchw", w[..., 0, 0], dy)
    wt = layout.pack_conv_weight(w.transpose(1, 2), dtype=act.dtype)                    # Cout -> Cin, 1x1: no tap flip needed
    dx = layout.to_planar(base, dtype=act.dtype)
    dyp = layout.to_planar(dy, dtype=act.dtype)
    native.check(lib.dba_conv_tc_ex(native.ptr(dyp), native.ptr(wt), wt[0].numel(), native.ptr(dx),
                                    native.ptr(dx), None, None, 0, None, 0, K, B, Ho, Ho, Cout, Cin, 1, 1, 0, 0, 0, 2,
                                    native.stream()))
    torch.cuda.synchronize()
    torch.testing.assert_close(layout.from_planar(dx, B, H, H), ref, rtol=act.rtol, atol=1.5 * act.atol)


def _guarded(shape, dtype, dev, guard=8192, fill=0.0):
    """A zero tensor of ``shape`` carved out of a larger allocation whose guard bands (``guard`` elements before and
    after) hold a sentinel: any out-of-bounds store of a kernel shows up as a changed sentinel."""
    n = 1
    for s_ in shape:
        n *= s_
    base = torch.full((n + 2 * guard,), 777.0, dtype=dtype, device=dev)
    view = base[guard:guard + n].view(shape)
    view.fill_(fill)
    return base, view, guard


def _guards_intact(base, guard):
    return bool((base[:guard] == 777.0).all()) and bool((base[-guard:] == 777.0).all())


@pytest.mark.parametrize("K,B,H,Cin,Cout,ks,sub", [(2, 3, 8, 32, 64, 3, 1), (1, 5, 4, 256, 256, 3, 1),
                                                    (2, 7, 16, 64, 128, 1, 2), (1, 64, 32, 32, 32, 3, 1)])
def test_conv_and_wgrad_write_nothing_outside_their_buffers(native_lib, act, K, B, H, Cin, Cout, ks, sub):
    """Out-of-bounds canary (compute-sanitizer is closed on this pool): output / gradient buffers sit between sentinel
    bands, inputs are followed by sentinels too (an over-read would put 777s into the result and fail the numerics)."""
    from dba_mod_b200.ops import layout, native
    lib = native.load()
    torch.manual_seed(7)
    d = torch.device("cuda", 0)
    x = act.round(torch.randn(K, B, Cin, H, H, device=d))
    w = act.round(torch.randn(K, Cout, Cin, ks, ks, device=d) / (Cin * ks * ks) ** 0.5)
    xp0, wt0 = layout.to_planar(x, dtype=act.dtype), layout.pack_conv_weight(w, dtype=act.dtype)
    bx, xp, gx = _guarded(tuple(xp0.shape), act.dtype, d)
    xp.copy_(xp0)
    bw, wt, gw = _guarded(tuple(wt0.shape), act.dtype, d)
    wt.copy_(wt0)
    Ho = H // sub
    go = layout.Geom(B, Ho, Ho)
    by, y, gy = _guarded((K, Cout // act.V, go.rows, act.V), act.dtype, d)
    stats_b, stats, gs = _guarded((K, Cout, 2), torch.float32, d)
    native.check(lib.dba_conv_tc(native.ptr(xp), native.ptr(wt), native.ptr(y), None, None, None, native.ptr(stats), K, B,
                                 H, H, Cin, Cout, ks, sub, 0, 0, 0, native.stream()))
    torch.cuda.synchronize()
    assert _guards_intact(by, gy) and _guards_intact(stats_b, gs) and _guards_intact(bx, gx) and _guards_intact(bw, gw)
    got = layout.from_planar(y, B, Ho, Ho)
    torch.testing.assert_close(got, _ref(x, w, ks, sub, act), rtol=act.rtol, atol=act.atol)
    assert layout.halo_is_zero(y, B, Ho, Ho)
    if sub == 1:        # weight gradient of the same layer, gradient slab between sentinels
        dy = act.round(torch.randn(K, B, Cout, H, H, device=d))

        def operand(t):
            if act.name == "tf32":
                hi = t.bfloat16()
                return torch.stack([layout.to_planar(hi.float()), layout.to_planar((t - hi.float()).bfloat16().float())],
                                   1).contiguous()
            return layout.to_planar(t, dtype=act.dtype)
        dy0, x0 = operand(dy), operand(x)
        bd, dyp, gd = _guarded(tuple(dy0.shape), dy0.dtype, d)
        dyp.copy_(dy0)
        bx2, xp2, gx2 = _guarded(tuple(x0.shape), x0.dtype, d)
        xp2.copy_(x0)
        nG = Cout * ks * ks * Cin
        bG, G, gG = _guarded((K, nG), torch.float32, d)
        native.check(lib.dba_wgrad_tc(native.ptr(dyp), native.ptr(xp2), native.ptr(G), G.stride(0), 0, K, B, H, H, Cin,
                                      Cout, Cin, Cout, ks, native.stream()))
        torch.cuda.synchronize()
        assert _guards_intact(bG, gG) and _guards_intact(bd, gd) and _guards_intact(bx2, gx2)
        assert bool(torch.isfinite(G).all()) and float(G.abs().max()) < 1e4      # no sentinel leaked into the sums
