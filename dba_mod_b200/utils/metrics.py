"""Headless replacement for the reference's Visdom dashboard (models/simple.py:18-200, main.py:34,122).

There is no web / graphics stack on the target boxes, so every "window" of the reference becomes a record in a
JSON-lines stream ``{folder}/metrics.jsonl``: one object per plotted point, keyed by the window name the
reference used (``train_acc``, ``test_acc``, ``poison_test_acc``, ``trigger_test_acc``, ``global_dist``,
``aggregation_weight``, ``fg_alpha``, ``train_batch_loss``).  The toggles ``vis_train``,
``vis_train_batch_loss``, ``vis_trigger_split_test`` and ``batch_track_distance`` gate the same series.
``dict_html`` reproduces utils/utils.py:8-19 (the params table) and is saved as ``params.html``.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional

_SKIP_IN_HTML = {"poisoning_test", "test_batch_size", "discount_size", "folder_path", "log_interval",
                 "coefficient_transfer", "grad_threshold"}


def dict_html(dict_obj: Dict[str, Any], current_time: str) -> str:
    rows = "".join(f"<tr><td>{k}</td><td>{v}</td></tr>" for k, v in dict_obj.items() if k not in _SKIP_IN_HTML)
    return f"<h4>Params for model: {current_time}:</h4><table>{rows}</table>"


class MetricsStream:
    def __init__(self, folder_path: Optional[str], enabled: bool = True) -> None:
        self.path = os.path.join(folder_path, "metrics.jsonl") if folder_path else None
        self.enabled = enabled and self.path is not None
        self._fh = open(self.path, "a") if self.enabled else None

    def point(self, window: str, x, y, name: str = "", **extra) -> None:
        if not self.enabled:
            return
        rec = {"win": window, "x": _num(x), "y": _num(y), "name": str(name)}
        rec.update({k: _num(v) for k, v in extra.items()})
        self._fh.write(json.dumps(rec) + "\n")

    # --- the ten plotting methods of SimpleNet, as series -------------------------------------------------
    def train_vis(self, epoch, acc, loss=None, name="", is_poisoned=False):
        self.point("train_acc", epoch, acc, name, poisoned=is_poisoned)
        if loss is not None:
            self.point("train_loss", epoch, loss, name, poisoned=is_poisoned)

    def train_batch_vis(self, epoch, data_len, batch, loss, name="", is_poisoned=False):
        self.point("train_batch_loss", (epoch - 1) * data_len + batch, loss, name, poisoned=is_poisoned)

    def track_distance_batch_vis(self, epoch, data_len, batch, distance, name="", is_poisoned=False):
        self.point("global_dist", (epoch - 1) * data_len + batch + 1, distance, f"Model_{name}", poisoned=is_poisoned)

    def test_vis(self, epoch, acc, loss=None, name=""):
        self.point("test_acc", epoch, acc, name)
        if loss is not None:
            self.point("test_loss", epoch, loss, name)

    def poison_test_vis(self, epoch, acc, loss=None, name=""):
        self.point("poison_test_acc", epoch, acc, name)
        if loss is not None:
            self.point("poison_test_loss", epoch, loss, name)

    def trigger_agent_test_vis(self, epoch, acc, name=""):
        self.point("trigger_test_acc", epoch, acc, name)

    def weight_vis(self, epoch, weight, name="", is_poisoned=False):
        self.point("aggregation_weight", epoch, weight, name, poisoned=is_poisoned)

    def alpha_vis(self, epoch, alpha, name="", is_poisoned=False):
        self.point("fg_alpha", epoch, alpha, name, poisoned=is_poisoned)

    def flush(self) -> None:
        if self._fh:
            self._fh.flush()

    def close(self) -> None:
        if self._fh:
            self._fh.close()
            self._fh = None
            self.enabled = False


def _num(v):
    if hasattr(v, "item"):
        try:
            return v.item()
        except Exception:  # noqa: BLE001
            return str(v)
    return v
