"""ctypes binding of the in-tree native library ``libdba_b200.so`` (built by ``__graft_entry__.build()`` /
``tools/build_native.py`` with nvcc for sm_100a).  Declarations mirror include/dba_b200.h.

On a machine with a GPU a missing library is an error (no silent fallback to library kernels); on a CPU-only
machine ``available()`` is simply False and the torch backend is used.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_float, c_int, c_longlong, c_ulonglong, c_void_p
from typing import Optional

import torch

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libdba_b200.so")


class NativeError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _sig(lib, name, argtypes, restype=c_int):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


def load() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not available():
        raise NativeError(f"{_LIB_PATH} not found - run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = ctypes.CDLL(_LIB_PATH)
    P, I, L, F, D, U = c_void_p, c_int, c_longlong, c_float, c_double, c_ulonglong
    _sig(lib, "dba_last_error", [], ctypes.c_char_p)
    _sig(lib, "dba_launch_count", [], c_ulonglong)
    _sig(lib, "dba_reset_launch_count", [], None)
    _sig(lib, "dba_add_launch_count", [c_ulonglong], None)
    _sig(lib, "dba_stage_step", [P, P, P, P, P, P, P, I, I, I, I, P, P, P, P, P, P, P])
    _sig(lib, "dba_set_pdl", [I], None)
    _sig(lib, "dba_get_pdl", [], c_int)
    _sig(lib, "dba_set_bn_fused", [I], None)
    _sig(lib, "dba_set_bn_pair", [I], None)
    _sig(lib, "dba_set_wgrad_vstack", [I], None)
    _sig(lib, "dba_get_bn_fused", [], c_int)
    _sig(lib, "dba_set_act_dtype", [I])
    _sig(lib, "dba_get_act_dtype", [], c_int)
    _sig(lib, "dba_sgd_step", [P, L, P, P, P, L, P, P, I, L, F, F, P])
    _sig(lib, "dba_scale_rows", [P, L, P, L, P, I, L, P])
    _sig(lib, "dba_dist_grad", [P, L, P, L, P, L, P, P, P, P, P, I, F, I, L, P])
    _sig(lib, "dba_rows_sqdist_scratch_elems", [I], c_longlong)
    _sig(lib, "dba_rows_sqdist", [P, L, P, L, P, I, L, P, I, P, L, P])
    _sig(lib, "dba_rows_sqdist_noalloc_max_rows", [], c_int)
    _sig(lib, "dba_rows_sqdist_noalloc", [P, L, P, L, P, I, L, P, I, P])
    _sig(lib, "dba_fedavg_update", [P, P, L, I, L, F, F, U, I, P])
    _sig(lib, "dba_weiszfeld", [P, L, P, I, L, P, I, F, D, P, P, P, P, P, P, P, L, P])
    _sig(lib, "dba_foolsgold_weights", [P, L, P, P, L, I, L, I, P, P, P, P, P, P])
    _sig(lib, "dba_foolsgold_weights_ptrs", [P, L, P, P, I, L, I, P, P, P, P, P, P])
    _sig(lib, "dba_foolsgold_apply", [P, P, L, P, I, L, F, F, F, P])
    _sig(lib, "dba_conv_rows", [I, I, I], c_longlong)
    _sig(lib, "dba_conv_set_debug", [P], None)
    _sig(lib, "dba_conv_set_debug_shift", [c_int], None)
    _sig(lib, "dba_conv_tc", [P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_conv_tc_strided", [P, P, L, P, P, P, P, L, P, L, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_conv_tc_ex", [P, P, L, P, P, P, P, L, P, L, I, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_conv_tc_valid", [P, P, L, P, P, P, P, L, P, L, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_wgrad_tc_ex", [P, P, L, P, L, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_wgrad_tc_s2d", [P, P, P, L, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_wgrad_linear_chw", [P, P, P, L, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_maxpool2_fwd", [P, P, I, I, I, I, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_maxpool2_fwd_split", [P, P, P, I, I, I, I, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_maxpool2_bwd_relu_bias", [P, P, P, P, L, I, I, I, I, I, I, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_maxpool2_bwd_relu_bias_split", [P, P, P, P, P, L, I, I, I, I, I, I, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_relu_bwd_bias", [P, P, P, P, L, I, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_relu_bwd_bias_split", [P, P, P, P, P, L, I, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_pack_weights", [P, L, P, I, I, P, L, P])
    _sig(lib, "dba_bn_fold", [P, L, P, I, I, P, P, L, F, P])
    _sig(lib, "dba_input_stage_image", [P, P, P, L, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_head", [P, P, L, I, I, P, L, P, P, I, P, P, P, P, I, P, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_head_ex", [P, P, L, I, I, P, L, P, P, I, P, P, P, P, I, P, I, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_bn_apply", [P, P, P, P, L, I, P, L, I, P, L, I, I, I, I, P, P, I, I, I, I, I, L, I, F, F, P])
    _sig(lib, "dba_bn_apply_ex", [P, P, P, P, L, P, L, I, P, L, I, P, L, I, I, I, I, P, P, I, I, I, I, I, L, I, F, F, P])
    _sig(lib, "dba_bn_apply_split", [P, P, P, P, L, P, P, P, L, I, P, L, I, P, L, I, I, I, I, P, P, I, I, I, I, I, L, I, F, F,
                                     P])
    _sig(lib, "dba_bn_backward_split", [P, P, P, P, P, P, P, L, I, P, L, I, P, L, P, L, I, I, P, I, I, I, I, I, L, I, L, P])
    _sig(lib, "dba_bn_backward_ex", [P, P, P, P, P, P, P, L, I, P, L, I, P, L, P, L, I, I, P, I, I, I, I, I, L, I, L, I, P])
    _sig(lib, "dba_conv_tc_dgrad_bn", [P, P, L, P, P, P, P, P, L, P, L, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_maxpool_fwd_split", [P, P, P, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_input_stage_image_split", [P, P, P, L, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, I, L, P])
    _sig(lib, "dba_bn_backward", [P, P, P, P, P, P, L, I, P, L, I, P, L, P, L, I, I, P, I, I, I, I, I, L, I, L, P])
    _sig(lib, "dba_maxpool_fwd", [P, P, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_maxpool_bwd", [P, P, P, I, I, I, I, I, L, L, P])
    _sig(lib, "dba_wgrad_tc", [P, P, P, L, I, I, I, I, I, I, I, I, I, I, P])
    _sig(lib, "dba_peer_init", [I, I, P, P, P, P, P, P, I, P, P, P, P, P, P])
    _sig(lib, "dba_peer_error", [], c_int)
    _sig(lib, "dba_peer_fedavg", [P, L, I, L, F, F, U, I, P])
    _sig(lib, "dba_peer_axpby", [P, L, I, P, L, F, F, I, I, P])
    _sig(lib, "dba_peer_weiszfeld", [P, L, I, L, P, P, I, F, D, I, I, P, I, P, P, P, P, P])
    _sig(lib, "dba_peer_barrier", [P])
    _sig(lib, "dba_gather_rows_ptrs", [P, I, L, P, L, P])
    _sig(lib, "dba_gram_ptrs", [P, I, L, P, P])
    _sig(lib, "dba_loan_train", [P, L, P, L, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, F, F,
                                 ctypes.c_uint, I, P])
    _sig(lib, "dba_loan_eval", [P, L, I, P, P, P, L, P, P, I, P, P, P])
    if os.environ.get("DBA_PDL") is not None:      # A/B switch for programmatic dependent launch (default: on)
        lib.dba_set_pdl(int(os.environ["DBA_PDL"]))
    if os.environ.get("DBA_WGRAD_VSTACK") is not None:   # A/B switch: tap-stacked weight gradient of the 32-channel layers
        lib.dba_set_wgrad_vstack(int(os.environ["DBA_WGRAD_VSTACK"]))
    if os.environ.get("DBA_BN_PAIR") is not None:       # A/B switch: BatchNorm apply on 8-channel planes in tf32 mode
        lib.dba_set_bn_pair(int(os.environ["DBA_BN_PAIR"]))
    if os.environ.get("DBA_BN_FUSED") is not None:  # A/B switch: cluster-fused BatchNorm backward (default: off)
        lib.dba_set_bn_fused(int(os.environ["DBA_BN_FUSED"]))
    _LIB = lib
    return lib


ACT_DTYPES = {"bf16": 0, "tf32": 1}


def set_act_dtype(name: str) -> None:
    """Activation element mode of the network kernels (library-wide, like a BLAS math mode): ``bf16`` = bf16
    activations, 8 channels per 16-byte chunk, ``tcgen05.mma.kind::f16``; ``tf32`` = fp32 activations, 4 channels per
    chunk, ``kind::tf32`` - the reference's precision class (fp32 tensors, TF32 convolutions)."""
    check(load().dba_set_act_dtype(ACT_DTYPES[name]))


def act_dtype() -> str:
    return "tf32" if load().dba_get_act_dtype() == 1 else "bf16"


def act_layout(name: str):
    """(torch dtype, channels per 16-byte chunk) of the planar activation / packed weight buffers of a mode."""
    return (torch.float32, 4) if name == "tf32" else (torch.bfloat16, 8)


def check(rc: int) -> None:
    if rc != 0:
        msg = load().dba_last_error().decode(errors="replace")
        raise NativeError(f"native call failed (code {rc}): {msg}")


def ptr(t: Optional[torch.Tensor]) -> c_void_p:
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    return int(load().dba_launch_count()) if _LIB is not None or available() else 0
