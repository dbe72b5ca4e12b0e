"""Phase timers: CUDA events on the launching stream on a GPU, ``perf_counter`` on CPU.

The reference prints ``time.time()`` deltas that are not synchronised with the device (SURVEY
5.1, C27).  ``PhaseTimer`` keeps named accumulators; on CUDA each phase is bracketed by events and
resolved lazily (no sync inside the step), so it can stay enabled in production runs.
"""
from __future__ import annotations

import time
from collections import defaultdict
from contextlib import contextmanager
from typing import Dict, List, Tuple

import torch


class PhaseTimer:
    def __init__(self, device: torch.device, enabled: bool = True):
        self.cuda = device.type == "cuda"
        self.enabled = enabled
        self.totals_ms: Dict[str, float] = defaultdict(float)
        self.counts: Dict[str, int] = defaultdict(int)
        self._pending: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    @contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                self._pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self.totals_ms[name] += (time.perf_counter() - t0) * 1e3
                self.counts[name] += 1

    def resolve(self) -> None:
        if self._pending:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.totals_ms[name] += a.elapsed_time(b)
                self.counts[name] += 1
            self._pending.clear()

    def mean_ms(self, name: str) -> float:
        self.resolve()
        return self.totals_ms[name] / max(self.counts[name], 1)

    def summary(self) -> Dict[str, float]:
        self.resolve()
        return {k: self.totals_ms[k] / max(self.counts[k], 1) for k in self.totals_ms}

    def reset(self) -> None:
        self.resolve()
        self.totals_ms.clear()
        self.counts.clear()


@contextmanager
def nvtx_range(name: str):
    """NVTX range around a phase of the step (visible in Nsight Systems / ncu --nvtx); a no-op without CUDA.

    The reference marks its phases with ``print(time.time() - start)`` pairs (R:265, R:317, R:322, R:388,
    R:397, R:440); these are the same phases as ranges a profiler can attribute kernels to."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()
