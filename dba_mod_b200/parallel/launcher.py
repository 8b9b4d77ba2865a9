"""Process-per-GPU bootstrap.

The reference is a single process on ``cuda:0`` (config.py:2) with no distributed code at all; its CLI
is ``python main.py --params X``.  To keep that contract while using all GPUs of the box, the parent process
re-executes the same command line once per GPU with RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* set (what ``torchrun``
would do) and waits.  Under an external ``torchrun`` (env already set) nothing is spawned.
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
from typing import List

import torch


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def world_from_env() -> int:
    return int(os.environ.get("WORLD_SIZE", "1"))


def maybe_spawn(params, argv: List[str]) -> bool:
    """Returns True in the parent after the children finished (caller should exit)."""
    if "RANK" in os.environ:
        return False
    n = int(params.get("n_gpus") or 0)
    avail = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n == 0:
        n = avail
    n = min(n, avail) if avail else 1
    if n <= 1:
        return False
    port = _free_port()
    procs = []
    for r in range(n):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(n), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable] + argv, env=env))
    # poll all children: the first failure takes the others down (their kernels would otherwise wait for a dead
    # peer until the barrier time-out)
    import time
    rc = 0
    alive = list(procs)
    while alive and rc == 0:
        for p in list(alive):
            r = p.poll()
            if r is not None:
                alive.remove(p)
                rc = rc or r
        if alive and rc == 0:
            time.sleep(0.2)
    if rc:
        for p in alive:
            p.terminate()
        for p in alive:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
        raise SystemExit(rc)
    return True


def make_comm(backend: str = None):
    """LocalComm for one process; a ``torch.distributed`` communicator when launched with WORLD_SIZE > 1."""
    from ..fl.aggregation import LocalComm, TorchDistComm
    if world_from_env() <= 1:
        return LocalComm()
    import torch.distributed as dist
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend=backend)
    return TorchDistComm()
