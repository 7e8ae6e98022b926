"""Slot-table operations: the index-space form of the reference's window shuffles.

A *window* of the reference is a float32 array whose rows are physically permuted every
producer iteration (reference ddl/datapusher.py:147-170).  Here a window is a *slot table*: an
int64 tensor of sample ids ``source << 48 | row`` on which the same permutations act while the
sample bytes stay where they were loaded.  ``perm_gen`` is the local shuffle
(reference tests/run_ddl.py:163-167), ``index_exchange`` the global one
(reference ddl/shuffle.py:92-108).  Each op has a CPU statement used as oracle and CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _native as N
from .rng import Pcg64


def _stream_ptr(device: torch.device, stream: Optional[torch.cuda.Stream]) -> C.c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream(device)
    return C.c_void_p(s.cuda_stream)


def rng_states_to_tensor(rngs: Sequence[Pcg64], device: torch.device) -> torch.Tensor:
    """[n, 6] int64 tensor holding the native 48-byte states (device layout of perm_gen)."""
    host = np.stack([r.to_words() for r in rngs])
    return torch.from_numpy(host).to(device)


def rng_states_from_tensor(t: torch.Tensor) -> List[Pcg64]:
    host = t.detach().cpu().numpy()
    out = []
    for row in host:
        r = Pcg64(0)
        r.from_words(row)
        out.append(r)
    return out


def slot_iota(table: torch.Tensor, source: int, first_row: int = 0, stream=None) -> None:
    """table[i] = source << 48 | (first_row + i)."""
    if table.dtype != torch.int64 or not table.is_contiguous():
        raise ValueError("slot tables are contiguous int64 tensors")
    if table.is_cuda:
        N.check(N.lib().bddl_slot_iota(table.data_ptr(), table.numel(), source, first_row,
                                       _stream_ptr(table.device, stream)))
    else:
        table.copy_(torch.arange(first_row, first_row + table.numel(), dtype=torch.int64)
                    + (int(source) << N.SLOT_ROW_BITS))


_perm_scratch: dict = {}


def perm_gen_scratch(device: torch.device, n: int, n_tables: int = 1) -> torch.Tensor:
    """Device scratch for ``perm_gen`` (targets, reservation words, work lists: 28 bytes per entry),
    cached per device and grown on demand so the shuffle path never allocates in steady state."""
    need = n_tables * ((int(N.lib().bddl_perm_gen_scratch_bytes(n)) + 63) // 64 * 64)
    key = (device.type, device.index)
    buf = _perm_scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = _perm_scratch[key] = torch.empty(need, dtype=torch.uint8, device=device)
    return buf


def perm_gen(tables: torch.Tensor, states: torch.Tensor, n: Optional[int] = None, stream=None,
             scratch: Optional[torch.Tensor] = None) -> None:
    """Shuffle each row of ``tables`` ([T, stride] int64) in place with ``states`` ([T, 6] int64).

    Device tensors run the sm_100a kernel; CPU tensors run the host Fisher-Yates.  Both advance
    the states exactly as ``numpy.random.Generator.shuffle`` would.
    """
    if tables.dim() == 1:
        tables = tables.unsqueeze(0)
    if states.dim() == 1:
        states = states.unsqueeze(0)
    T, stride = tables.shape
    n = stride if n is None else int(n)
    if tables.dtype != torch.int64 or states.dtype != torch.int64 or states.shape != (T, 6):
        raise ValueError("perm_gen expects int64 tables [T, n] and int64 states [T, 6]")
    if tables.stride(1) != 1 or not states.is_contiguous():
        raise ValueError("perm_gen needs unit-stride tables and contiguous states")
    if tables.is_cuda:
        if scratch is None:
            scratch = perm_gen_scratch(tables.device, n, T)
        N.check(N.lib().bddl_perm_gen(tables.data_ptr(), tables.stride(0), n, states.data_ptr(), T,
                                      scratch.data_ptr(), scratch.numel(), _stream_ptr(tables.device, stream)))
        return
    for t in range(T):
        r = Pcg64(0)
        r.from_words(states[t].numpy())
        view = tables[t, :n].numpy()
        r.shuffle(view)
        states[t].copy_(torch.from_numpy(r.to_words()))


def index_exchange(dst: torch.Tensor, own: torch.Tensor, from_recv: Optional[torch.Tensor],
                   from_send: Optional[torch.Tensor], num_exchange: int, stream=None,
                   from_recv_ptr: Optional[int] = None, from_send_ptr: Optional[int] = None) -> None:
    """dst = own with [0, ne/2) taken from ``from_recv`` and [ne/2, ne) from ``from_send``.

    The partner tables may be given as raw peer-mapped addresses (``*_ptr``) when they live on
    another GPU.
    """
    n = own.numel()
    ne = int(num_exchange)
    if dst.is_cuda:
        fr = from_recv_ptr if from_recv_ptr is not None else (from_recv.data_ptr() if from_recv is not None else None)
        fs = from_send_ptr if from_send_ptr is not None else (from_send.data_ptr() if from_send is not None else None)
        N.check(N.lib().bddl_index_exchange(dst.data_ptr(), own.data_ptr(), fr, fs, n, ne,
                                            _stream_ptr(dst.device, stream)))
        return
    h = ne // 2
    dst.copy_(own)
    if ne > 0:
        dst[:h] = from_recv[:h]
        dst[h:ne] = from_send[h:ne]
