"""The scratch-free row-distance entry point, called through the C ABI the way an outside caller would.

Who needs it: code that has no device buffer to offer as scratch and may not create one - for instance a process
whose device memory is entirely owned by its own pool, or a monitoring thread for which cudaMalloc is ruled out because
it stalls every stream of the device.  ``dba_rows_sqdist_noalloc`` keeps its partial sums in an array that is part of
the library image, so it has a row limit (64) and all calls share that array.

Checked here: the result equals ``dba_rows_sqdist`` (caller-supplied scratch) within the tolerance the other
flat-kernel tests use, on two consecutive calls with different shapes and data (the second call finds the first
call's partial sums still in the array and must not be disturbed by them), and an oversized call is refused with a
message instead of being launched.  The file name sorts last on purpose: a run that stops at the first failure has
then already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else 0)


def test_noalloc_matches_scratch_version_on_two_calls(native_lib):
    lib = native_lib
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    max_rows = int(lib.dba_rows_sqdist_noalloc_max_rows())
    assert max_rows == 64
    torch.manual_seed(11)
    scratch = torch.empty(int(lib.dba_rows_sqdist_scratch_elems(max_rows)), dtype=torch.float64, device=dev)
    # (rows, n, per-row Y, Z, sqrt); the last call fills the whole built-in array (64 rows = the limit)
    calls = [(10, 2797610, True, True, True), (3, 5529, False, True, False), (max_rows, 70001, True, False, True)]
    for K, n, per_row_y, use_z, take_sqrt in calls:
        X = torch.randn(K, n, device=dev)
        Y = torch.randn(K if per_row_y else 1, n, device=dev)
        Z = torch.randn(n, device=dev) if use_z else None
        ldy = Y.stride(0) if per_row_y else 0
        ref = torch.full((K,), -1.0, dtype=torch.float64, device=dev)
        got = torch.full((K,), -2.0, dtype=torch.float64, device=dev)
        rc = lib.dba_rows_sqdist(_ptr(X), X.stride(0), _ptr(Y), ldy, _ptr(Z), K, n, _ptr(ref), int(take_sqrt),
                                 _ptr(scratch), scratch.numel(), stream)
        assert rc == 0, lib.dba_last_error()
        rc = lib.dba_rows_sqdist_noalloc(_ptr(X), X.stride(0), _ptr(Y), ldy, _ptr(Z), K, n, _ptr(got),
                                         int(take_sqrt), stream)
        assert rc == 0, lib.dba_last_error()
        torch.cuda.synchronize()
        torch.testing.assert_close(got, ref, rtol=1e-6, atol=1e-6)
        # and both agree with a plain PyTorch restatement
        d = X.double() - Y.double() - (Z.double() if Z is not None else 0.0)
        plain = d.pow(2).sum(1)
        torch.testing.assert_close(got, plain.sqrt() if take_sqrt else plain, rtol=1e-6, atol=1e-6)


def test_noalloc_refuses_more_rows_than_its_scratch_holds(native_lib):
    lib = native_lib
    dev = torch.device("cuda", 0)
    K = int(lib.dba_rows_sqdist_noalloc_max_rows()) + 1
    X = torch.zeros(K, 8, device=dev)
    out = torch.zeros(K, dtype=torch.float64, device=dev)
    rc = lib.dba_rows_sqdist_noalloc(_ptr(X), X.stride(0), _ptr(None), 0, _ptr(None), K, 8, _ptr(out), 1,
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.dba_last_error().decode()
    assert "64" in msg and "exceed" in msg
