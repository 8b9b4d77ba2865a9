"""Planar-8 / shared-halo activation layout and packed weight layouts used by the tensor-core kernels
(csrc/conv_tc.cu has the rationale).  These torch implementations are the reference packers used by the tests and by
cold paths (checkpoint import, oracle comparisons); the hot path packs with kernels.

  activation  X[k][c8][row][8] bf16,  row(b,h,w) = lead + b*S + h*Wp + w,  Wp = W+1, S = (H+1)*Wp, lead = Wp+1
  conv weight [k][tap][cin8][cout][8] bf16  (tap = kh*3+kw), i.e. K-major rows of 8 input channels per output channel
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class Geom:
    B: int
    H: int
    W: int

    @property
    def Wp(self) -> int:
        return self.W + 1

    @property
    def S(self) -> int:
        return (self.H + 1) * self.Wp

    @property
    def lead(self) -> int:
        return self.Wp + 1

    @property
    def rows(self) -> int:
        body = (self.B * self.S + 511) // 512 * 512
        return self.lead + body + self.Wp + 2


def _v(dtype) -> int:
    """Channels per 16-byte chunk: 8 for bf16 (planar-8), 4 for fp32 (planar-4, the tf32 mode)."""
    return 4 if dtype == torch.float32 else 8


def to_planar(x: torch.Tensor, cpad: int = None, dtype=torch.bfloat16) -> torch.Tensor:
    """x [K, B, C, H, W] float -> ``dtype`` [K, Cp/V, R, V] (channels zero-padded to ``cpad``)."""
    K, B, C, H, W = x.shape
    V = _v(dtype)
    Cp = cpad or ((C + 7) // 8 * 8)
    g = Geom(B, H, W)
    buf = torch.zeros(K, Cp // V, g.rows, V, dtype=dtype, device=x.device)
    body = torch.zeros(K, B, H + 1, g.Wp, Cp, dtype=dtype, device=x.device)
    body[:, :, :H, :W, :C] = x.permute(0, 1, 3, 4, 2).to(dtype)
    body = body.reshape(K, B * g.S, Cp // V, V).permute(0, 2, 1, 3)
    buf[:, :, g.lead:g.lead + B * g.S] = body
    return buf


def from_planar(y: torch.Tensor, B: int, H: int, W: int, C: int = None) -> torch.Tensor:
    """[K, Cp/V, R, V] (bf16 or fp32) -> float [K, B, C, H, W]."""
    K, C8 = y.shape[:2]
    g = Geom(B, H, W)
    body = y[:, :, g.lead:g.lead + B * g.S].permute(0, 2, 1, 3).reshape(K, B, H + 1, g.Wp, C8 * y.shape[3])
    out = body[:, :, :H, :W].permute(0, 1, 4, 2, 3).float()
    return out[:, :, :C] if C is not None else out


def space_to_depth(x: torch.Tensor) -> torch.Tensor:
    """[K, B, C, H, W] -> [K, B, 4C, H/2, W/2], channel p*C + c with parity plane p = 2*(h & 1) + (w & 1): the operand
    form of stride-2 convolutions (csrc/conv_tc.cu ``in_s2d``)."""
    K, B, C, H, W = x.shape
    v = x.reshape(K, B, C, H // 2, 2, W // 2, 2).permute(0, 1, 4, 6, 2, 3, 5)       # [K, B, ph, pw, C, H/2, W/2]
    return v.reshape(K, B, 4 * C, H // 2, W // 2)


def depth_to_space(x: torch.Tensor) -> torch.Tensor:
    """Inverse of :func:`space_to_depth`."""
    K, B, C4, H2, W2 = x.shape
    C = C4 // 4
    v = x.reshape(K, B, 2, 2, C, H2, W2).permute(0, 1, 4, 5, 2, 6, 3)               # [K, B, C, H/2, ph, W/2, pw]
    return v.reshape(K, B, C, 2 * H2, 2 * W2)


def halo_is_zero(y: torch.Tensor, B: int, H: int, W: int) -> bool:
    K, C8 = y.shape[:2]
    g = Geom(B, H, W)
    t = y.clone()
    body = t[:, :, g.lead:g.lead + B * g.S].reshape(K, C8, B, H + 1, g.Wp, y.shape[3])
    body[:, :, :, :H, :W] = 0
    return bool((t == 0).all())


def pack_conv_weight(w: torch.Tensor, cin_pad: int = None, cout_pad: int = None, dtype=torch.bfloat16) -> torch.Tensor:
    """w [K, Cout, Cin, kh, kw] -> ``dtype`` [K, kh*kw, Cin_p/V, Cout_p, V]."""
    K, Co, Ci, kh, kw = w.shape
    V = _v(dtype)
    Cip = cin_pad or ((Ci + 15) // 16 * 16)
    Cop = cout_pad or ((Co + 31) // 32 * 32)
    wp = torch.zeros(K, Cop, Cip, kh, kw, dtype=torch.float32, device=w.device)
    wp[:, :Co, :Ci] = w
    t = wp.permute(0, 3, 4, 2, 1).reshape(K, kh * kw, Cip // V, V, Cop).permute(0, 1, 2, 4, 3)
    return t.contiguous().to(dtype)


def pack_conv_weight_T(w: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """Weights for the input-gradient (dgrad) pass run as a forward convolution: roles of Cin/Cout swapped and the
    3x3 taps flipped:  w'[k][ci][co][2-kh][2-kw] = w[k][co][ci][kh][kw]."""
    return pack_conv_weight(w.transpose(1, 2).flip(-1, -2), dtype=dtype)
