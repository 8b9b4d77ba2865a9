"""Backdoor trigger tables.

The reference stamps triggers sample by sample on the host: ``add_pixel_pattern`` deep-copies one image and sets the
listed ``[row, col]`` pixels to 1 on every channel (image_helper.py:328-350); the LOAN trainer / tests overwrite named
feature columns with fixed values (loan_train.py:47-57,99-107; test.py:62-81).  Here every trigger becomes one ROW of
a small device-resident table, built once:

  images : ``masks[T + 1, H, W]`` bool      row 0 = union of all ``trigger_num`` patterns (the global trigger,
                                             ``adversarial_index == -1``), row j + 1 = local pattern j
  LOAN   : ``mask[T + 1, F]`` bool and ``val[T + 1, F]`` float, same row convention

so poisoning a batch is a predicated store inside the input stage of the first layer (``csrc/net_ops.cu``
``input_stage_kernel``, ``csrc/loan_mlp.cu``) or one ``masked_fill`` on the torch backend.  The ``stamp_*`` functions
are the plain-torch statement of the same thing and the oracle the kernels are tested against.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch


def pixel_mask(params, adv_index: int, H: int, W: int) -> torch.Tensor:
    """[H, W] bool mask of one pattern (``adv_index == -1``: all patterns together)."""
    m = torch.zeros(H, W, dtype=torch.bool)
    for pos in params.pixel_pattern(adv_index):
        r, c = int(pos[0]), int(pos[1])
        if not (0 <= r < H and 0 <= c < W):
            raise ValueError(f"trigger pixel {pos} outside a {H}x{W} image")
        m[r, c] = True
    return m


def all_pixel_masks(params, H: int, W: int) -> torch.Tensor:
    """[trigger_num + 1, H, W]: row 0 global, row j + 1 local trigger j."""
    T = int(params.get("trigger_num", 0))
    if any(f"{j}_poison_pattern" not in params for j in range(T)):
        T = 0                                  # a clean run may name triggers without defining them
    if T == 0:
        return torch.zeros(1, H, W, dtype=torch.bool)
    rows = [pixel_mask(params, -1, H, W)] + [pixel_mask(params, j, H, W) for j in range(T)]
    return torch.stack(rows)


def stamp_images_(x: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """In place: pixels under ``mask[H, W]`` become 1.0 on all channels of ``x[..., C, H, W]``."""
    x.masked_fill_(mask.to(x.device).expand_as(x), 1.0)
    return x


def all_feature_triggers(params, feature_dict: Dict[str, int], n_features: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """(mask, val), both [trigger_num + 1, n_features]; row 0 applies every trigger's columns."""
    T = int(params.get("trigger_num", 0))
    if any(f"{j}_poison_trigger_names" not in params for j in range(T)):
        T = 0
    mask = torch.zeros(T + 1, n_features, dtype=torch.bool)
    val = torch.zeros(T + 1, n_features, dtype=torch.float32)
    for row, adv in enumerate([-1] + list(range(T))):
        if T == 0:
            break
        names, values = params.feature_trigger(adv)
        for nm, v in zip(names, values):
            if nm not in feature_dict:
                raise KeyError(f"trigger feature {nm!r} is not a column of the LOAN table")
            mask[row, feature_dict[nm]] = True
            val[row, feature_dict[nm]] = float(v)
    return mask, val


def stamp_features_(x: torch.Tensor, mask_row: torch.Tensor, val_row: torch.Tensor) -> torch.Tensor:
    """In place: columns under ``mask_row[F]`` of ``x[..., F]`` take the values of ``val_row[F]``."""
    m = mask_row.to(x.device).expand_as(x)
    x.copy_(torch.where(m, val_row.to(x.device, x.dtype).expand_as(x), x))
    return x
