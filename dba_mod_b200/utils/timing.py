"""Device-side phase timers (the reference only has wall-clock ``time.time()`` prints)."""
from __future__ import annotations

import time
from contextlib import contextmanager
from typing import Dict, List

import torch


class PhaseTimer:
    """CUDA-event (or perf_counter on CPU) timers per named phase; resolved lazily, no sync while running."""

    def __init__(self, device: torch.device, enabled: bool = True) -> None:
        self.cuda = device.type == "cuda"
        self.enabled = enabled
        self._pending: Dict[str, List] = {}
        self.totals_ms: Dict[str, float] = {}

    @contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(f"dba/{name}")       # visible in Nsight timelines
            a.record()
            try:
                yield
            finally:
                b.record()
                torch.cuda.nvtx.range_pop()
                self._pending.setdefault(name, []).append((a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self.totals_ms[name] = self.totals_ms.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def resolve(self) -> Dict[str, float]:
        if self.cuda and self._pending:
            torch.cuda.synchronize()
            for name, evs in self._pending.items():
                self.totals_ms[name] = self.totals_ms.get(name, 0.0) + sum(a.elapsed_time(b) for a, b in evs)
            self._pending.clear()
        return dict(self.totals_ms)

    def reset(self) -> None:
        self._pending.clear()
        self.totals_ms.clear()
