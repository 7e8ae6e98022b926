"""Device-side batch fingerprint (csrc/checksum.cu) and its numpy statement.

The reference has no way to check what a consumer received (its only test asserts an exit code,
reference tests/test_ddl.py:8-28).  ``BatchChecksum`` fingerprints a batch where it lies in HBM;
``checksum_reference`` is the same arithmetic on the host.  Integer sums modulo 2^64 do not depend
on the order of summation, so device and host agree exactly.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from .. import _native as N

_GOLD = np.uint64(0x9E3779B97F4A7C15)


def _mix64(x: np.ndarray) -> np.ndarray:
    x = x.copy()
    x ^= x >> np.uint64(33)
    x *= np.uint64(0xFF51AFD7ED558CCD)
    x ^= x >> np.uint64(33)
    x *= np.uint64(0xC4CEB9FE1A85EC53)
    x ^= x >> np.uint64(33)
    return x


def checksum_reference(batch: np.ndarray, payload_bytes: Optional[int] = None) -> int:
    """batch: [rows, row_bytes] uint8 (or any 2-D array viewed as bytes)."""
    b = np.ascontiguousarray(batch).view(np.uint8).reshape(batch.shape[0], -1)
    payload = b.shape[1] if payload_bytes is None else payload_bytes
    words = np.ascontiguousarray(b[:, :payload]).view(np.uint64).reshape(-1)
    with np.errstate(over="ignore"):
        pos = (np.arange(1, words.size + 1, dtype=np.uint64)) * _GOLD
        return int(_mix64(words ^ pos).sum(dtype=np.uint64))


class BatchChecksum:
    """Reusable launcher; owns its scratch and result words on the device."""

    def __init__(self, device: torch.device, max_rows: int, payload_bytes: int):
        self.device = torch.device(device)
        words = int(N.lib().bddl_batch_checksum_scratch_words(max_rows, payload_bytes))
        self.scratch = torch.zeros(words, dtype=torch.int64, device=self.device)
        self.out = torch.zeros(1, dtype=torch.int64, device=self.device)

    def run(self, batch: torch.Tensor, payload_bytes: Optional[int] = None, stream=None) -> torch.Tensor:
        """Enqueue the fingerprint of ``batch`` ([rows, row_bytes] bytes, unit inner stride)."""
        raw = batch if batch.dtype == torch.uint8 else batch.view(torch.uint8)
        rows, row_bytes = raw.shape[0], raw.stride(0)
        payload = raw.shape[1] if payload_bytes is None else payload_bytes
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        N.check(N.lib().bddl_batch_checksum(raw.data_ptr(), rows, row_bytes, payload, self.scratch.data_ptr(),
                                            self.scratch.numel(), self.out.data_ptr(), C.c_void_p(s.cuda_stream)))
        return self.out

    def value(self) -> int:
        return int(self.out.item()) & (2**64 - 1)
