"""Tracing (the reference has only DEBUG entry/exit log lines).

With ``LoaderConfig(trace=True)`` every batch assembly and every producer iteration is bracketed
by CUDA events on the stream that runs it (device time, not host time) and by an NVTX range, so
the loader shows up by name in Nsight timelines.  ``summary()`` synchronises and reduces the
recorded pairs; with tracing off nothing is recorded and no event is created.
"""
from __future__ import annotations

from collections import defaultdict
from contextlib import contextmanager
from typing import Dict, List, Tuple

import torch


class Tracer:
    def __init__(self, enabled: bool, cuda: bool, cap: int = 8192):
        self.enabled = bool(enabled) and cuda
        self.cap = cap
        self.pairs: Dict[str, List[Tuple[torch.cuda.Event, torch.cuda.Event]]] = defaultdict(list)

    @contextmanager
    def span(self, name: str, stream):
        if not self.enabled:
            yield
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.nvtx.range_push("ddl_b200." + name)
        a.record(stream)
        try:
            yield
        finally:
            b.record(stream)
            torch.cuda.nvtx.range_pop()
            lst = self.pairs[name]
            if len(lst) < self.cap:
                lst.append((a, b))

    def summary(self) -> dict:
        out = {}
        if not self.enabled:
            return out
        torch.cuda.synchronize()
        for name, lst in self.pairs.items():
            ms = sorted(a.elapsed_time(b) for a, b in lst)
            if ms:
                out[name] = {"count": len(ms), "mean_ms": sum(ms) / len(ms), "median_ms": ms[len(ms) // 2],
                             "max_ms": ms[-1], "total_ms": sum(ms)}
        return out
