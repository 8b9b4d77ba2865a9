"""tcgen05 implicit-GEMM convolution vs torch fp32 conv2d on the same bf16-rounded operands."""
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
]


def _ref(x, w, ksize, sub):
    K = x.shape[0]
    xb, wb = x.bfloat16().float(), w.bfloat16().float()
    return torch.stack([F.conv2d(xb[k], wb[k], None, sub, ksize // 2) for k in range(K)])


@pytest.mark.parametrize("K,B,H,W,Cin,Cout,ksize,sub", SHAPES)
def test_conv_raw(native_lib, K, B, H, W, Cin, Cout, ksize, sub):
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    d = torch.device("cuda", 0)
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, ksize, ksize, device=d) / (Cin * ksize * ksize) ** 0.5
    y = conv.conv_tc(layout.to_planar(x), layout.pack_conv_weight(w), K, B, H, W, ksize, sub)
    torch.cuda.synchronize()
    got = layout.from_planar(y, B, H // sub, W // sub)
    ref = _ref(x, w, ksize, sub)
    torch.testing.assert_close(got, ref, rtol=2e-2, atol=2e-2)
    assert layout.halo_is_zero(y, B, H // sub, W // sub)


def test_conv_eval_epilogue_and_shared_input(native_lib):
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(1)
    d = torch.device("cuda", 0)
    K, B, H, W, Cin, Cout = 3, 6, 16, 16, 64, 64
    x = torch.randn(1, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5
    res = torch.randn(K, B, Cout, H, W, device=d)
    scale, shift = torch.rand(K, Cout, device=d) + 0.5, torch.randn(K, Cout, device=d)
    y = conv.conv_tc(layout.to_planar(x), layout.pack_conv_weight(w), K, B, H, W, 3, 1, relu=True,
                     res=layout.to_planar(res), scale=scale, shift=shift, x_shared=True)
    got = layout.from_planar(y, B, H, W)
    ref = _ref(x.expand(K, -1, -1, -1, -1), w, 3, 1)
    ref = F.relu(ref * scale[:, None, :, None, None] + shift[:, None, :, None, None] + res.bfloat16().float())
    torch.testing.assert_close(got, ref, rtol=2e-2, atol=3e-2)


def test_conv_stats(native_lib):
    from dba_mod_b200.ops import conv, layout
    torch.manual_seed(2)
    d = torch.device("cuda", 0)
    K, B, H, W, Cin, Cout = 2, 9, 8, 8, 64, 128
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, 3, 3, device=d) / (Cin * 9) ** 0.5
    stats = torch.zeros(K, Cout, 2, device=d)
    y = conv.conv_tc(layout.to_planar(x), layout.pack_conv_weight(w), K, B, H, W, 3, 1, stats=stats)
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
def test_conv_stride2_from_space_to_depth_input(native_lib, K, B, H, W, Cin, Cout, ksize):
    """Stride-2 convolutions computed on the output grid only, from the parity-plane form of their input."""
    from dba_mod_b200.ops import conv, layout
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(2)
    d = torch.device("cuda", 0)
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, ksize, ksize, device=d) / (Cin * ksize * ksize) ** 0.5
    xs = layout.to_planar(layout.space_to_depth(x))
    stats = torch.zeros(K, Cout, 2, device=d)
    y = conv.conv_tc_s2d(xs, layout.pack_conv_weight(w), K, B, H // 2, W // 2, ksize, in_s2d=True, stats=stats)
    torch.cuda.synchronize()
    got = layout.from_planar(y, B, H // 2, W // 2)
    ref = _ref(x, w, ksize, 2)
    torch.testing.assert_close(got, ref, rtol=2e-2, atol=2e-2)
    assert layout.halo_is_zero(y, B, H // 2, W // 2)
    torch.testing.assert_close(stats[..., 0], got.sum((1, 3, 4)), rtol=1e-3, atol=2e-2)


def test_conv_space_to_depth_output_with_residual(native_lib):
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
    y1 = conv.conv_tc_s2d(layout.to_planar(x), layout.pack_conv_weight(w1), K, B, H, W, 3, out_s2d=True, relu=True,
                          res=layout.to_planar(res), scale=scale, shift=shift)
    ref1 = F.relu(_ref(x, w1, 3, 1) * scale[:, None, :, None, None] + shift[:, None, :, None, None]
                  + res.bfloat16().float())
    got1 = layout.depth_to_space(layout.from_planar(y1, B, H // 2, W // 2))
    torch.testing.assert_close(got1, ref1, rtol=2e-2, atol=3e-2)
    assert layout.halo_is_zero(y1, B, H // 2, W // 2)
    y2 = conv.conv_tc_s2d(y1, layout.pack_conv_weight(w2), K, B, H // 2, W // 2, 3, in_s2d=True)
    ref2 = _ref(got1, w2, 3, 2)
    torch.testing.assert_close(layout.from_planar(y2, B, H // 2, W // 2), ref2, rtol=2e-2, atol=3e-2)
