"""Checkpoint save / resume in the reference's on-disk format.

Format (helper.py:420-435): ``torch.save({'state_dict': OrderedDict[str, Tensor], 'epoch': int, 'lr': float})``
to ``{folder}/model_last.pt.tar`` (+ ``.epoch_{e}`` when ``e in save_on_epochs``; + ``.best`` when
``val_loss < best_loss``, where val_loss is whatever loss main.py computed last — Q7).  Resume
(image_helper.py:56-67, loan_helper.py:94-106): ``saved_models/{resumed_model_name}`` -> load state_dict,
``start_epoch = epoch + 1``, ``lr = ckpt.get('lr', lr)``.

The state_dict uses the reference's key names / shapes / dtypes (incl. int64 ``num_batches_tracked``) so files
move between the two code bases unchanged; the flat-buffer <-> state_dict mapping is ParamSpec's job.
"""
from __future__ import annotations

import logging
import math
import os
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

logger = logging.getLogger("logger")


class CheckpointWriter:
    def __init__(self, params) -> None:
        self.params = params
        self.best_loss = math.inf

    def save_model(self, state_dict: "OrderedDict[str, torch.Tensor]", epoch: int, val_loss: float) -> list:
        written = []
        if not self.params["save_model"]:
            return written
        logger.info("saving model")
        base = os.path.join(self.params["folder_path"], "model_last.pt.tar")
        payload = {"state_dict": OrderedDict((k, v.detach().cpu()) for k, v in state_dict.items()),
                   "epoch": int(epoch), "lr": self.params["lr"]}
        torch.save(payload, base)
        written.append(base)
        if epoch in self.params["save_on_epochs"]:
            logger.info(f"Saving model on epoch {epoch}")
            torch.save(payload, f"{base}.epoch_{epoch}")
            written.append(f"{base}.epoch_{epoch}")
        if val_loss < self.best_loss:
            torch.save(payload, f"{base}.best")
            written.append(f"{base}.best")
            self.best_loss = val_loss
        return written


def load_checkpoint(path: str) -> Tuple[Dict[str, torch.Tensor], int, Optional[float]]:
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    return ckpt["state_dict"], int(ckpt["epoch"]), ckpt.get("lr")


def resume_path(params, root: str = "saved_models") -> str:
    return os.path.join(root, str(params["resumed_model_name"]))
