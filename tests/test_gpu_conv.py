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
