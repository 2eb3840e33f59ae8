"""Timing and metrics.

The reference's only instrument is two ``datetime.now()`` calls around ten iterations, divided
by nine, unsynchronised (ref: master/part2a/part2a.py:54-59).  Here every step
is bracketed by CUDA events on the launching stream (host clock on CPU), phases can be timed
separately, and results go out as one JSON line.
"""
from __future__ import annotations

import json
import statistics
import sys
import time
from typing import Dict, List

import torch


class StepClock:
    """Per-step timer: CUDA events on a GPU (no host sync inside the loop), perf_counter on CPU."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        self._begin: Dict[int, object] = {}
        self._end: Dict[int, object] = {}

    def step_begin(self, i: int) -> None:
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._begin[i] = ev
        else:
            self._begin[i] = time.perf_counter()

    def step_end(self, i: int) -> None:
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._end[i] = ev
        else:
            self._end[i] = time.perf_counter()

    def durations(self) -> List[float]:
        """Seconds per step, begin(i) -> begin(i+1) where available (includes inter-step gaps)."""
        keys = sorted(self._begin)
        out = []
        for a, b in zip(keys, keys[1:]):
            if self.cuda:
                out.append(self._begin[a].elapsed_time(self._begin[b]) * 1e-3)
            else:
                out.append(self._begin[b] - self._begin[a])
        if keys:
            last = keys[-1]
            if self.cuda:
                out.append(self._begin[last].elapsed_time(self._end[last]) * 1e-3)
            else:
                out.append(self._end[last] - self._begin[last])
        return out

    def summary(self, skip: int = 3) -> Dict[str, float]:
        if self.cuda:
            torch.cuda.synchronize()
        d = self.durations()
        steady = d[skip:] if len(d) > skip else d
        if not steady:
            return {}
        return {"s_per_iter": sum(steady) / len(steady), "s_per_iter_median": statistics.median(steady),
                "timed_iters": len(steady)}


class PhaseTimer:
    """Named phase timing with CUDA events (forward / backward / sync / optimizer), used by the
    profiling helpers to report exposed communication."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        self._open: Dict[str, object] = {}
        self.totals: Dict[str, float] = {}
        self._pending: List[tuple] = []

    def start(self, name: str) -> None:
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._open[name] = ev
        else:
            self._open[name] = time.perf_counter()

    def stop(self, name: str) -> None:
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._pending.append((name, self._open.pop(name), ev))
        else:
            self.totals[name] = self.totals.get(name, 0.0) + time.perf_counter() - self._open.pop(name)

    def collect(self) -> Dict[str, float]:
        if self.cuda:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.totals[name] = self.totals.get(name, 0.0) + a.elapsed_time(b) * 1e-3
            self._pending = []
        return dict(self.totals)


def emit_json(obj: dict, stream=None) -> None:
    (stream or sys.stdout).write(json.dumps(obj) + "\n")
    (stream or sys.stdout).flush()


_NVTX = None


def nvtx_range(name: str):
    """NVTX range around a phase of the step (forward / loss / backward / sync) when ``B200DP_NVTX=1`` and a
    GPU is present; otherwise a no-op context.  The ranges show up in Nsight Systems / ``torch.profiler``
    traces of an eager run (``--no-cuda-graph``); a captured step replays without host code, so there the
    phases are told apart by kernel names (scripts/timeline_step.py).  SURVEY.md 5.1."""
    global _NVTX
    if _NVTX is None:
        import os
        _NVTX = os.environ.get("B200DP_NVTX", "0") != "0" and torch.cuda.is_available()
    if not _NVTX:
        import contextlib
        return contextlib.nullcontext()
    return torch.cuda.nvtx.range(name)
