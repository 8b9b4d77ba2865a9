"""Tensor-level wrapper of the tcgen05 implicit-GEMM convolution (csrc/conv_tc.cu)."""
from __future__ import annotations

from typing import Optional

import torch

from . import native
from .layout import Geom


def _mode_from(x: torch.Tensor, wt: torch.Tensor) -> int:
    """Sets the library's activation mode from the operand dtype; returns channels per 16-byte chunk."""
    assert x.dtype == wt.dtype and x.dtype in (torch.bfloat16, torch.float32), (x.dtype, wt.dtype)
    assert x.is_contiguous() and wt.is_contiguous()
    native.set_act_dtype("tf32" if x.dtype == torch.float32 else "bf16")
    return 4 if x.dtype == torch.float32 else 8


def conv_tc(x: torch.Tensor, wt: torch.Tensor, K: int, B: int, H: int, W: int, ksize: int, sub: int = 1,
            relu: bool = False, res: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None,
            shift: Optional[torch.Tensor] = None, stats: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None, x_shared: bool = False, debug_swap: int = 0) -> torch.Tensor:
    """x planar [Kx, Cin/V, R, V]; wt packed [K, taps, Cin/V, Cout, V] -> planar [K, Cout/V, Ro, V]; bf16 tensors
    (V = 8) run on ``kind::f16``, fp32 tensors (V = 4) on ``kind::tf32`` - the mode is set from the operand dtype.

    ``out`` must be a buffer whose halo rows are zero (they are never written).  ``stats`` [K, Cout, 2] fp32 is
    accumulated into (caller zeroes it)."""
    lib = native.load()
    V = _mode_from(x, wt)
    Cin, Cout = wt.shape[2] * V, wt.shape[3]
    assert x.shape[1] * V == Cin and x.shape[2] == Geom(B, H, W).rows, (x.shape, Cin, Geom(B, H, W).rows)
    go = Geom(B, H // sub, W // sub)
    if out is None:
        out = torch.zeros(K, Cout // V, go.rows, V, dtype=x.dtype, device=x.device)
    assert out.shape == (K, Cout // V, go.rows, V) and out.is_contiguous()
    native.check(lib.dba_conv_tc(native.ptr(x), native.ptr(wt), native.ptr(out), native.ptr(res), native.ptr(scale),
                                 native.ptr(shift), native.ptr(stats), K, B, H, W, Cin, Cout, ksize, sub, int(relu),
                                 int(x_shared), int(debug_swap), native.stream()))
    return out


def conv_tc_s2d(x: torch.Tensor, wt: torch.Tensor, K: int, B: int, H: int, W: int, ksize: int,
                in_s2d: bool = False, out_s2d: bool = False, relu: bool = False,
                res: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None,
                shift: Optional[torch.Tensor] = None, stats: Optional[torch.Tensor] = None,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Space-to-depth variants (``dba_conv_tc_ex``).  (H, W) is the grid the convolution is COMPUTED on: with
    ``in_s2d`` that is the stride-2 layer's output grid and x is ``layout.to_planar(layout.space_to_depth(input))``
    with 4*Cin channels; with ``out_s2d`` the result has 4*Cout channels on the (H/2, W/2) grid."""
    lib = native.load()
    V = _mode_from(x, wt)
    Cin, Cout = wt.shape[2] * V, wt.shape[3]
    assert x.shape[1] * V == (4 if in_s2d else 1) * Cin and x.shape[2] == Geom(B, H, W).rows
    go = Geom(B, H // 2, W // 2) if out_s2d else Geom(B, H, W)
    oshape = (K, (4 if out_s2d else 1) * Cout // V, go.rows, V)
    if out is None:
        out = torch.zeros(oshape, dtype=x.dtype, device=x.device)
    assert tuple(out.shape) == oshape and out.is_contiguous()
    native.check(lib.dba_conv_tc_ex(
        native.ptr(x), native.ptr(wt), wt[0].numel(), native.ptr(out), native.ptr(res), native.ptr(scale),
        native.ptr(shift), Cout, native.ptr(stats), 2 * Cout, K, B, H, W, Cin, Cout, ksize, 1, int(relu), 0,
        int(in_s2d), int(out_s2d), native.stream()))
    return out
