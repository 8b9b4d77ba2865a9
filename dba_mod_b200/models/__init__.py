"""Model families: ParamSpec + reference-layout modules + batched functional networks."""
from __future__ import annotations

import torch

from .batched import BatchedNet  # noqa: F401
from .nn_compat import build_module, display_name  # noqa: F401
from .spec import ParamSpec  # noqa: F401


def initial_state(kind: str, seed: int = 1):
    """(spec, flat[Df], counters) of a freshly initialised global model.

    The reference seeds torch with 1 (main.py:36-37) and constructs Local then Target (image_helper.py:36-52); the
    Target model — the second construction — is the global model.  Doing the same here gives identical initial
    weights for the same torch version."""
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        build_module(kind)
        target = build_module(kind)
    spec = ParamSpec(target)
    flat, cnt = spec.flatten(target.state_dict())
    return spec, flat, cnt
