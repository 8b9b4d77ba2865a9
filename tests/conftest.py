import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 via gpurun)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def native_lib():
    """The in-tree native library; GPU tests must fail loudly when it is missing (no silent fallback)."""
    from dba_mod_b200.ops import native
    return native.load()


class _Act:
    """One activation element mode of the network kernels: operand dtype, channels per 16-byte chunk, the rounding the
    tensor core applies to operands (so that the fp32 oracle sees the same operand values) and the tolerance band."""

    def __init__(self, name):
        import torch
        self.name = name
        self.dtype = torch.float32 if name == "tf32" else torch.bfloat16
        self.V = 4 if name == "tf32" else 8
        # bf16: 8-bit mantissa operands AND bf16-stored outputs; tf32: 10-bit mantissa operands, fp32 outputs
        self.rtol, self.atol = (2e-3, 2e-3) if name == "tf32" else (2e-2, 2e-2)

    def round(self, t):
        import torch
        if self.name == "bf16":
            return t.bfloat16().float()
        return (t.float().contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)    # tf32: low 13 bits dropped

    def store(self, t):
        """Rounding applied when a value is stored as a tensor-core operand tensor: bf16 conversion, or round to
        the nearest tf32 value (ties away from zero, ``cvt.rna.tf32.f32``) kept in fp32."""
        import torch
        if self.name == "bf16":
            return t.bfloat16().float()
        return ((t.float().contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


@pytest.fixture(params=["bf16", "tf32"])
def act(request):
    """Parametrises a GPU kernel test over both activation modes and selects the mode in the native library."""
    from dba_mod_b200.ops import native
    native.set_act_dtype(request.param)
    yield _Act(request.param)
    native.set_act_dtype("bf16")


@pytest.fixture(autouse=True)
def _reset_act_mode(request):
    """The activation mode is a library-wide switch: every test starts from bf16 unless it asks otherwise."""
    if "gpu" in request.keywords:
        from dba_mod_b200.ops import native
        if native.available():
            native.set_act_dtype("bf16")
    yield
