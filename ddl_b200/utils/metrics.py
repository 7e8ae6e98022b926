"""Loader counters (the reference has none -- only log lines and prints).

Counting is a handful of integer additions per batch; with ``enabled=False`` it is one branch.
``timing=True`` additionally measures the host-visible loader stall: how long ``loader[i]`` had
to wait for the batch to be assembled (it synchronises on the batch event, so keep it off when
benchmarking throughput).
"""
from __future__ import annotations

import json
import time
from typing import Optional


class LoaderMetrics:
    def __init__(self, enabled: bool = True, timing: bool = False):
        self.enabled, self.timing = enabled, timing
        self.batches = self.samples = self.bytes_in = self.bytes_out = self.epochs = 0
        self.stall_s = 0.0
        self.stall_max_s = 0.0
        self.t0 = time.perf_counter()

    def batch(self, samples: int, bytes_in: int, bytes_out: int) -> None:
        if self.enabled:
            self.batches += 1
            self.samples += samples
            self.bytes_in += bytes_in
            self.bytes_out += bytes_out

    def stall(self, seconds: float) -> None:
        self.stall_s += seconds
        self.stall_max_s = max(self.stall_max_s, seconds)

    def epoch(self) -> None:
        if self.enabled:
            self.epochs += 1

    def as_dict(self) -> dict:
        wall = time.perf_counter() - self.t0
        return {"batches": self.batches, "samples": self.samples, "bytes_in": self.bytes_in,
                "bytes_out": self.bytes_out, "epochs": self.epochs, "stall_s": self.stall_s,
                "stall_max_s": self.stall_max_s, "wall_s": wall,
                "samples_per_s": self.samples / wall if wall > 0 else 0.0}

    def dump_jsonl(self, path: str, extra: Optional[dict] = None) -> None:
        rec = self.as_dict()
        if extra:
            rec.update(extra)
        with open(path, "a") as fh:
            fh.write(json.dumps(rec) + "\n")
