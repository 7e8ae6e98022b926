"""numpy-compatible random streams backed by the native PCG64 (csrc/rng_pcg64.h).

Every random decision of the reference comes from ``numpy.random.default_rng(seed)``: the
partner permutation (reference ddl/shuffle.py:28-30, 53) and the user's per-epoch row shuffle
(reference tests/run_ddl.py:122, 167).  ``Pcg64`` reproduces that stream bit for bit without
numpy in the loop, can be checkpointed (`state_dict`), and can be copied to / from the device
where ``perm_gen`` continues it.
"""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import numpy as np

from .. import _native as N


class Pcg64:
    """Drop-in for the subset of ``numpy.random.Generator`` the reference uses."""

    __slots__ = ("_st",)

    def __init__(self, seed: int = 0):
        if seed < 0 or seed >= 1 << 64:
            raise ValueError("seed must fit in 64 bits and be non-negative")
        self._st = N.RngState()
        N.lib().bddl_rng_seed(C.byref(self._st), seed)

    # -- draws ------------------------------------------------------------------------------
    def next32(self) -> int:
        return int(N.lib().bddl_rng_next32(C.byref(self._st)))

    def next64(self) -> int:
        return int(N.lib().bddl_rng_next64(C.byref(self._st)))

    def interval(self, max_inclusive: int) -> int:
        return int(N.lib().bddl_rng_interval(C.byref(self._st), max_inclusive))

    def advance(self, n_outputs64: int) -> None:
        N.lib().bddl_rng_advance(C.byref(self._st), n_outputs64)

    # -- shuffles (same draws and swaps as Generator.shuffle) ------------------------------------
    def shuffle(self, x: np.ndarray) -> None:
        """In-place shuffle along axis 0, identical to ``Generator.shuffle(x)``."""
        if not isinstance(x, np.ndarray):
            raise TypeError("Pcg64.shuffle expects a numpy array")
        n = x.shape[0] if x.ndim else 0
        if n <= 1:
            return
        if not x.flags.c_contiguous or not x.flags.writeable:
            raise ValueError("Pcg64.shuffle needs a writable C-contiguous array")
        lib = N.lib()
        if x.ndim == 1 and x.dtype.itemsize == 8:
            N.check(lib.bddl_host_shuffle_u64(C.byref(self._st), x.ctypes.data, n))
        elif x.ndim == 1 and x.dtype.itemsize == 4:
            N.check(lib.bddl_host_shuffle_i32(C.byref(self._st), x.ctypes.data, n))
        else:
            N.check(lib.bddl_host_shuffle_rows(C.byref(self._st), x.ctypes.data, n, x.strides[0]))

    def permutation(self, n: int) -> np.ndarray:
        """The permutation one ``shuffle`` applies to ``arange(n)`` (as int64)."""
        t = np.arange(n, dtype=np.int64)
        self.shuffle(t)
        return t

    def partner(self, n_instance: int, my_rank: int) -> Tuple[int, int, np.ndarray, int]:
        """(send_to, recv_from, permutation, rejected_tries) -- reference ddl/shuffle.py:32-79."""
        s, r, tries = C.c_int32(), C.c_int32(), C.c_int32()
        perm = np.zeros(max(n_instance, 1), dtype=np.int32)
        N.check(
            N.lib().bddl_partner_select(
                C.byref(self._st), n_instance, my_rank, C.byref(s), C.byref(r), perm.ctypes.data, C.byref(tries)
            )
        )
        return int(s.value), int(r.value), perm, int(tries.value)

    # -- state ------------------------------------------------------------------------------------
    @property
    def draws32(self) -> int:
        return int(self._st.draws32)

    def state_dict(self) -> dict:
        s = self._st
        return {
            "state": (int(s.state_hi) << 64) | int(s.state_lo),
            "inc": (int(s.inc_hi) << 64) | int(s.inc_lo),
            "has_uint32": int(s.has_uint32),
            "uinteger": int(s.uinteger),
            "draws32": int(s.draws32),
        }

    def load_state_dict(self, d: dict) -> None:
        s = self._st
        s.state_lo, s.state_hi = d["state"] & (2**64 - 1), d["state"] >> 64
        s.inc_lo, s.inc_hi = d["inc"] & (2**64 - 1), d["inc"] >> 64
        s.has_uint32, s.uinteger = d["has_uint32"], d["uinteger"]
        s.draws32 = d.get("draws32", 0)

    def to_words(self) -> np.ndarray:
        """The 48-byte native state as six int64 words (the device layout ``perm_gen`` uses)."""
        return np.frombuffer(bytes(self._st), dtype=np.int64).copy()

    def from_words(self, w: np.ndarray) -> None:
        w = np.ascontiguousarray(w, dtype=np.int64)
        if w.size != 6:
            raise ValueError("rng state is six 64-bit words")
        C.memmove(C.byref(self._st), w.ctypes.data, 48)

    def clone(self) -> "Pcg64":
        c = Pcg64(0)
        C.memmove(C.byref(c._st), C.byref(self._st), C.sizeof(N.RngState))
        return c

    def numpy_state(self) -> dict:
        """State in the shape ``numpy.random.PCG64().state`` expects (for cross-checks)."""
        d = self.state_dict()
        return {
            "bit_generator": "PCG64",
            "state": {"state": d["state"], "inc": d["inc"]},
            "has_uint32": d["has_uint32"],
            "uinteger": d["uinteger"],
        }
