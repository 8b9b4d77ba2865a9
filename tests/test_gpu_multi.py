"""Multi-GPU tests (need >= 2 GPUs on the box; skipped otherwise): fused peer-memory collectives and world-size
invariance of a short CIFAR simulation on the native backend."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _need2():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")


def _torchrun(n, script, *args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(29600 + os.getpid() % 300), script, *args]
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_p2p_collectives_two_ranks(native_lib):
    _need2()
    r = _torchrun(2, "tools/gpu_test_p2p.py")
    assert r.returncode == 0 and "P2P_OK" in r.stdout, (r.stdout[-1500:], r.stderr[-1500:])


def test_bench_two_ranks_runs(native_lib):
    _need2()
    r = _torchrun(2, "bench.py", "--gpus", "2", "--steps", "2", "--warmup", "3", timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["n_gpus"] == 2 and out["value"] > 0 and "peer-memory" in out["config"]["collectives"]


@pytest.mark.parametrize("method", ["mean", "geom_median", "foolsgold"])
def test_world_size_invariance_native(native_lib, method):
    """1 rank vs 2 ranks (clients sharded, fused peer-memory collectives): same global model up to reduction-order
    noise (fp32 atomics in the BatchNorm statistics / weight gradients), same aggregation weights, zero NCCL calls
    inside the rounds."""
    _need2()
    one = subprocess.run([sys.executable, "tools/gpu_sim_worker.py", method, "2"], cwd=ROOT, capture_output=True,
                         text=True, timeout=600, env=dict(os.environ, WORLD_SIZE="1"))
    assert one.returncode == 0, one.stderr[-1500:]
    two = _torchrun(2, "tools/gpu_sim_worker.py", method, "2")
    assert two.returncode == 0, two.stderr[-1500:]
    a = json.loads([l for l in one.stdout.splitlines() if l.startswith("RESULT ")][0][7:])
    b = json.loads([l for l in two.stdout.splitlines() if l.startswith("RESULT ")][0][7:])
    assert "peer-memory" in b["path"]
    # training, per-client evaluation exchange, aggregation and global evaluation of every round ran without a single
    # NCCL collective (fused peer-memory kernels only; torch.distributed is start-up plumbing)
    assert b["nccl_calls_in_rounds"] == 0, b
    assert abs(a["norm"] - b["norm"]) / a["norm"] < 2e-2, (a, b)
    if a["w"] is not None:
        # weights are reported in rank-concatenated client order, which depends on the placement -> compare sorted
        # FoolsGold's logit re-scaling is steep and training from a random init is not reproducible run to run
        # (fp32 atomics order + different batch widths per placement; two 1-rank runs differ the same way), so only
        # gross disagreement fails here - exact 1-vs-2-rank agreement of the aggregators is asserted with the
        # deterministic fp32 backend in tests/test_cpu_sim.py::test_world_size_invariance_gloo (weights to 1e-3)
        assert max(abs(x - y) for x, y in zip(sorted(a["w"]), sorted(b["w"]))) < 0.35, (a["w"], b["w"])
