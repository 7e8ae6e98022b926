"""Python face of the fused batch-assembly kernel ``exchange_gather_collate`` (csrc/gather.cu).

The reference hands out zero-copy CPU views of a shuffled window
(reference ddl/mpi_dataloader.py:179-198) and leaves conversion and the host-to-device copy to
the user (reference tests/run_ddl.py:233-235).  Here one kernel launch gathers the batch's
samples from whichever GPU (or pinned host buffer) owns them, converts and collates them into
the batch tensor.  ``gather_reference`` is the plain PyTorch fp32 statement of the same
operation, used as the numerics oracle and as the CPU execution path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _native as N

SLOT_ROW_BITS = N.SLOT_ROW_BITS
SLOT_ROW_MASK = (1 << SLOT_ROW_BITS) - 1

_IN_ES = {N.CVT_COPY: 1, N.CVT_U8_BF16: 1, N.CVT_F32_BF16: 4, N.CVT_U8_F32: 1, N.CVT_I32_I64: 4}
_OUT_ES = {N.CVT_COPY: 1, N.CVT_U8_BF16: 2, N.CVT_F32_BF16: 2, N.CVT_U8_F32: 4, N.CVT_I32_I64: 8}
_OUT_DTYPE = {
    N.CVT_COPY: torch.uint8,
    N.CVT_U8_BF16: torch.bfloat16,
    N.CVT_F32_BF16: torch.bfloat16,
    N.CVT_U8_F32: torch.float32,
    N.CVT_I32_I64: torch.int64,
}
_IN_NP = {N.CVT_COPY: np.uint8, N.CVT_U8_BF16: np.uint8, N.CVT_F32_BF16: np.float32, N.CVT_U8_F32: np.uint8,
          N.CVT_I32_I64: np.int32}


def make_slot(source: int, row: int) -> int:
    return (int(source) << SLOT_ROW_BITS) | int(row)


def split_slot(slot: int) -> Tuple[int, int]:
    slot = int(slot) & (2**64 - 1)
    return slot >> SLOT_ROW_BITS, slot & SLOT_ROW_MASK


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass(frozen=True)
class Convert:
    """What happens to a sample on its way into the batch tensor."""

    mode: int = N.CVT_COPY
    scale: Tuple[float, ...] = (1.0,)
    bias: Tuple[float, ...] = (0.0,)
    chan_stride: int = 0  # elements per channel plane (CHW layouts); 0 = one scale/bias
    pad_value: float = 0.0

    @staticmethod
    def copy() -> "Convert":
        return Convert(N.CVT_COPY)

    @staticmethod
    def normalize_u8(mean: Sequence[float], std: Sequence[float], chan_stride: int, out: str = "bf16") -> "Convert":
        """uint8 image -> ((x/255) - mean) / std, as ``x*scale + bias`` per channel."""
        scale = tuple(float(np.float32(1.0) / (np.float32(255.0) * np.float32(s))) for s in std)
        bias = tuple(float(-np.float32(m) / np.float32(s)) for m, s in zip(mean, std))
        mode = N.CVT_U8_BF16 if out == "bf16" else N.CVT_U8_F32
        return Convert(mode, scale, bias, chan_stride if len(scale) > 1 else 0)

    @staticmethod
    def cast_f32_bf16(scale: Sequence[float] = (1.0,), bias: Sequence[float] = (0.0,), chan_stride: int = 0) -> "Convert":
        return Convert(N.CVT_F32_BF16, tuple(scale), tuple(bias), chan_stride if len(scale) > 1 else 0)

    @staticmethod
    def widen_tokens() -> "Convert":
        return Convert(N.CVT_I32_I64)

    @property
    def in_elem_size(self) -> int:
        return _IN_ES[self.mode]

    @property
    def out_elem_size(self) -> int:
        return _OUT_ES[self.mode]

    @property
    def out_dtype(self) -> torch.dtype:
        return _OUT_DTYPE[self.mode]

    def out_bytes(self, sample_bytes: int) -> int:
        return sample_bytes // self.in_elem_size * self.out_elem_size

    def pad_bits(self) -> int:
        if self.mode in (N.CVT_COPY,):
            return int(self.pad_value) & 0xFF
        if self.mode == N.CVT_I32_I64:
            return int(self.pad_value) & 0xFFFFFFFF
        if self.mode == N.CVT_U8_F32:
            return int(np.float32(self.pad_value).view(np.uint32))
        return int(torch.tensor(self.pad_value, dtype=torch.bfloat16).view(torch.int16).item()) & 0xFFFF


class SourceTable:
    """Device-resident table of shard base addresses (one per source), the kernel's view of the
    symmetric heap.  A source is a local HBM buffer, a peer GPU's buffer mapped over NVLink, or
    a pinned host buffer mapped into the device address space."""

    def __init__(self, base_ptrs: Sequence[int], device: torch.device,
                 offsets: Optional[Sequence[torch.Tensor]] = None,
                 lengths: Optional[Sequence[torch.Tensor]] = None):
        self.device = torch.device(device)
        self.n_sources = len(base_ptrs)
        self.base = torch.tensor([int(p) for p in base_ptrs], dtype=torch.int64).to(self.device)
        self._keep = (offsets, lengths)
        self.offsets = self.lengths = None
        if offsets is not None:
            if lengths is None or len(offsets) != self.n_sources or len(lengths) != self.n_sources:
                raise ValueError("variable-size sources need one offsets and one lengths tensor per source")
            for o, l in zip(offsets, lengths):
                if o.dtype != torch.int64 or l.dtype != torch.int32 or o.device != self.device:
                    raise ValueError("offsets must be int64 and lengths int32 on the table's device")
            self.offsets = torch.tensor([o.data_ptr() for o in offsets], dtype=torch.int64).to(self.device)
            self.lengths = torch.tensor([l.data_ptr() for l in lengths], dtype=torch.int64).to(self.device)

    def with_lengths(self, offsets: Sequence[torch.Tensor], lengths: Sequence[torch.Tensor]) -> "SourceTable":
        """The same sources, described as variable-size (CSR offsets + payload lengths)."""
        t = SourceTable.__new__(SourceTable)
        t.device, t.n_sources, t.base = self.device, self.n_sources, self.base
        t._keep = (offsets, lengths)
        t.offsets = torch.tensor([o.data_ptr() for o in offsets], dtype=torch.int64).to(self.device)
        t.lengths = torch.tensor([l.data_ptr() for l in lengths], dtype=torch.int64).to(self.device)
        return t


class GatherPlan:
    """A reusable launch description: everything but the slot list and the destination."""

    def __init__(self, sources: SourceTable, row_bytes: int, sample_bytes: int, convert: Convert = Convert(),
                 dst_row_bytes: Optional[int] = None, variant: int = N.VARIANT_AUTO, max_ctas: int = 0,
                 tune: int = 0, col_offset_bytes: int = 0):
        self.sources = sources
        self.convert = convert
        self.row_bytes = int(row_bytes)
        self.sample_bytes = int(sample_bytes)
        self.out_sample_bytes = convert.out_bytes(self.sample_bytes)
        self.dst_row_bytes = int(dst_row_bytes) if dst_row_bytes else round_up(self.out_sample_bytes, 16)
        d = N.GatherDesc()
        d.d_src_base = sources.base.data_ptr()
        d.d_src_offsets = sources.offsets.data_ptr() if sources.offsets is not None else None
        d.d_src_lengths = sources.lengths.data_ptr() if sources.lengths is not None else None
        d.row_bytes, d.sample_bytes, d.dst_row_bytes = self.row_bytes, self.sample_bytes, self.dst_row_bytes
        d.chan_stride = convert.chan_stride
        d.n_sources = sources.n_sources
        d.mode = convert.mode
        d.n_channels = len(convert.scale)
        if len(convert.scale) != len(convert.bias) or len(convert.scale) > N.MAX_CHANNELS:
            raise ValueError("scale/bias must have equal length <= %d" % N.MAX_CHANNELS)
        for i in range(N.MAX_CHANNELS):
            d.scale[i] = convert.scale[i] if i < len(convert.scale) else 1.0
            d.bias[i] = convert.bias[i] if i < len(convert.bias) else 0.0
        d.pad_bits = convert.pad_bits()
        d.variant, d.max_ctas, d.tune = variant, max_ctas, tune
        d.col_offset_bytes = int(col_offset_bytes)
        self._desc = d
        self._fn = N.lib().bddl_exchange_gather_collate
        self._step = N.lib().bddl_gather_step

    def push_ptr(self, push_desc, n_samples: int, stream_ptr: int) -> None:
        """Push form (``bddl_exchange_push_collate``): this GPU delivers the samples it owns into
        every consumer's ring slot; ``push_desc`` is a filled ``_native.PushDesc``."""
        d = self._desc
        d.d_slots, d.d_dst, d.n_samples = None, None, n_samples
        rc = N.lib().bddl_exchange_push_collate(C.byref(d), C.byref(push_desc), C.c_void_p(stream_ptr))
        if rc:
            N.check(rc)

    def alloc_batch(self, n_samples: int) -> torch.Tensor:
        """Raw batch buffer [n_samples, dst_row_bytes] uint8 (rows 16-byte aligned)."""
        return torch.empty((n_samples, self.dst_row_bytes), dtype=torch.uint8, device=self.sources.device)

    def view(self, raw: torch.Tensor, sample_shape: Optional[Sequence[int]] = None) -> torch.Tensor:
        """Typed view of a raw batch buffer: [n, *sample_shape] in the output dtype."""
        n = raw.shape[0]
        es = self.convert.out_elem_size
        typed = raw.view(self.convert.out_dtype) if es > 1 else raw
        n_elem = self.out_sample_bytes // es
        typed = typed[:, :n_elem]
        if sample_shape is not None:
            if self.dst_row_bytes == self.out_sample_bytes:
                typed = typed.reshape(n, *sample_shape)
            else:
                typed = typed.unflatten(1, tuple(sample_shape))
        return typed

    def run(self, slots: torch.Tensor, dst: torch.Tensor, n_samples: Optional[int] = None,
            stream: Optional[torch.cuda.Stream] = None) -> None:
        n = int(slots.numel()) if n_samples is None else int(n_samples)
        d = self._desc
        d.d_slots = slots.data_ptr()
        d.d_dst = dst.data_ptr()
        d.n_samples = n
        s = stream if stream is not None else torch.cuda.current_stream(self.sources.device)
        N.check(self._fn(C.byref(d), C.c_void_p(s.cuda_stream)))

    def step_ptr(self, slots_ptr: int, dst_ptr: int, n_samples: int, stream_ptr: int, wait_event: int,
                 record_event: int) -> None:
        """[wait for `wait_event`] -> launch -> [record `record_event`] in one native call."""
        d = self._desc
        d.d_slots, d.d_dst, d.n_samples = slots_ptr, dst_ptr, n_samples
        rc = self._step(C.byref(d), stream_ptr, wait_event, record_event)
        if rc:
            N.check(rc)

    def run_ptr(self, slots_ptr: int, dst_ptr: int, n_samples: int, stream_ptr: int) -> None:
        """Lowest-overhead launch for the per-step hot loop (raw addresses)."""
        d = self._desc
        d.d_slots, d.d_dst, d.n_samples = slots_ptr, dst_ptr, n_samples
        rc = self._fn(C.byref(d), C.c_void_p(stream_ptr))
        if rc:
            N.check(rc)


def convert_reference(raw: torch.Tensor, convert: Convert, lengths: Optional[torch.Tensor] = None,
                      copy_elem_size: int = 1) -> torch.Tensor:
    """Plain PyTorch (fp32 math, separate multiply and add) statement of the conversion epilogue
    on already gathered rows: ``raw`` is [n, sample_bytes] uint8 on any device, ``lengths`` the
    valid elements per row of variable-size windows.  Returns [n, elems] in the output dtype."""
    mode = convert.mode
    n = raw.shape[0]
    if mode == N.CVT_COPY:
        y, pad = raw.clone(), int(convert.pad_value) & 0xFF
    elif mode == N.CVT_I32_I64:
        y, pad = raw.view(torch.int32).to(torch.int64), int(convert.pad_value)
    else:
        t = (raw if convert.in_elem_size == 1 else raw.view(torch.float32)).to(torch.float32)
        n_elem, nch = t.shape[1], len(convert.scale)
        if nch > 1 and convert.chan_stride > 0:
            ch = (torch.arange(n_elem, device=raw.device) // convert.chan_stride) % nch
            sc = torch.tensor(convert.scale, dtype=torch.float32, device=raw.device)[ch]
            bi = torch.tensor(convert.bias, dtype=torch.float32, device=raw.device)[ch]
        else:
            sc = torch.tensor(convert.scale[0], dtype=torch.float32, device=raw.device)
            bi = torch.tensor(convert.bias[0], dtype=torch.float32, device=raw.device)
        y = t * sc
        y = (y + bi).to(convert.out_dtype)
        pad = convert.pad_value
    if lengths is not None:
        lengths = lengths.to(raw.device).long() * (copy_elem_size if mode == N.CVT_COPY else 1)  # copies pad bytes
        mask = torch.arange(y.shape[1], device=raw.device)[None, :] >= lengths[:, None]
        y[mask] = torch.tensor(pad, dtype=y.dtype, device=raw.device)
    return y.view(n, -1)


def gather_reference(shards: Sequence[np.ndarray], slots: np.ndarray, sample_bytes: int, convert: Convert,
                     lengths: Optional[Sequence[np.ndarray]] = None,
                     offsets: Optional[Sequence[np.ndarray]] = None) -> torch.Tensor:
    """Plain PyTorch (fp32 math) statement of exchange_gather_collate on the CPU.

    ``shards[s]`` is a uint8 array: [rows, row_bytes] for fixed-size samples, or flat bytes with
    ``offsets[s]`` / ``lengths[s]`` for variable-size ones.  Returns [n, elems] in the output dtype.
    """
    mode = convert.mode
    in_np = _IN_NP[mode]
    n_elem = sample_bytes // convert.in_elem_size
    rows = []
    for slot in np.asarray(slots).astype(np.uint64).tolist():
        src, row = split_slot(slot)
        if offsets is None:
            raw = np.ascontiguousarray(shards[src][row, :sample_bytes])
            ln = sample_bytes
        else:
            o, ln = int(offsets[src][row]), int(lengths[src][row])
            raw = np.ascontiguousarray(shards[src][o:o + ln])
        rows.append((raw.view(in_np), ln // convert.in_elem_size))
    out = torch.empty((len(rows), n_elem), dtype=convert.out_dtype)
    nch = len(convert.scale)
    if nch > 1 and convert.chan_stride > 0:
        ch = (torch.arange(n_elem) // convert.chan_stride) % nch
        sc = torch.tensor(convert.scale, dtype=torch.float32)[ch]
        bi = torch.tensor(convert.bias, dtype=torch.float32)[ch]
    else:
        sc = torch.full((n_elem,), convert.scale[0], dtype=torch.float32)
        bi = torch.full((n_elem,), convert.bias[0], dtype=torch.float32)
    for i, (vals, k) in enumerate(rows):
        t = torch.from_numpy(vals.copy())
        if mode == N.CVT_COPY:
            out[i, :k] = t
            out[i, k:] = int(convert.pad_value) & 0xFF
        elif mode == N.CVT_I32_I64:
            out[i, :k] = t.to(torch.int64)
            out[i, k:] = int(convert.pad_value)
        else:
            y = t.to(torch.float32) * sc[:k] + bi[:k]
            out[i, :k] = y.to(convert.out_dtype)
            out[i, k:] = torch.tensor(convert.pad_value, dtype=convert.out_dtype)
    return out
