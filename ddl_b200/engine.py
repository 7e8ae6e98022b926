"""Device transport engine: HBM shards, slot tables, streams and the batch ring.

This is the B200-first data plane that replaces the reference's shared-memory windows and
their physical shuffles (reference ddl/connection.py:88-139, ddl/datapusher.py:147-170).  The
idea: **rows never move, indices do**.

* Every producer's data is loaded once into an immutable *shard* in the symmetric heap
  (HBM, mapped by every peer over NVLink).
* Each reference "window" is a double-buffered *slot table* of 64-bit sample ids.  The global
  shuffle (``index_exchange``) and the local shuffle (``perm_gen``) act on the table on a side
  stream, one iteration ahead of the consumer.
* A batch is assembled by one ``exchange_gather_collate`` launch on the fetch stream, straight
  into one of ``ring_depth`` batch buffers, while the training stream works on the previous one.
  Sample bytes cross NVLink exactly once, inside that kernel.

With ``device.type == 'cpu'`` the same engine runs on POSIX shared memory and numpy/torch CPU
ops; that is the executable model the multi-rank CPU tests use.
"""
from __future__ import annotations

import ctypes as C
import logging
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _native as N
from .config import LoaderConfig
from .exceptions import DoesNotMatchError
from .ops import tables as T
from .ops.gather import SLOT_ROW_BITS, SLOT_ROW_MASK, Convert, GatherPlan, SourceTable, round_up
from .ops.rng import Pcg64
from .parallel.heap import SymmetricHeap
from .utils.tracing import Tracer

log = logging.getLogger("ddl_b200")

_TORCH_OF_NP = {np.dtype(np.float32): torch.float32, np.dtype(np.uint8): torch.uint8, np.dtype(np.int32): torch.int32,
                np.dtype(np.int64): torch.int64, np.dtype(np.float16): torch.float16,
                np.dtype(np.float64): torch.float64, np.dtype(np.int16): torch.int16, np.dtype(np.int8): torch.int8}


class NativeEvent:
    """A raw cudaEvent_t (timing disabled) driven through the native library: recording and
    waiting cost one ctypes call each, and ``bddl_gather_step`` can take the handles directly."""

    __slots__ = ("handle",)

    def __init__(self):
        h = C.c_void_p()
        N.check(N.lib().bddl_event_create(C.byref(h)))
        self.handle = h.value

    def record(self, stream=None) -> None:
        s = stream if stream is not None else torch.cuda.current_stream()
        rc = N.lib().bddl_event_record(self.handle, s.cuda_stream)
        if rc:
            N.check(rc)

    def synchronize(self) -> None:
        N.check(N.lib().bddl_event_synchronize(self.handle))

    def __del__(self):
        try:
            N.lib().bddl_event_destroy(self.handle)
        except Exception:
            pass


def stream_wait(stream, ev) -> None:
    """``stream`` waits for ``ev``: native, torch or null events alike."""
    if isinstance(ev, NativeEvent):
        rc = N.lib().bddl_stream_wait_event(stream.cuda_stream, ev.handle)
        if rc:
            N.check(rc)
    elif ev is not None:
        stream.wait_event(ev)


class NullEvent:
    def record(self, stream=None):
        pass

    def synchronize(self):
        pass

    def query(self):
        return True

    def elapsed_time(self, other):
        return 0.0


class NullStream:
    cuda_stream = 0

    def wait_event(self, ev):
        pass

    def wait_stream(self, s):
        pass

    def synchronize(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class WindowSpec:
    """Geometry of one producer's window, derived from its ``on_init`` answer."""

    def __init__(self, meta, seq_rank: int = 0, seq_size: int = 1):
        self.nData = int(meta.nData)
        self.window_values = int(meta.nValues)          # elements per row of the window / shard
        self.np_dtype = np.dtype(meta.dtype if meta.dtype is not None else np.float32)
        self.window_bytes = self.window_values * self.np_dtype.itemsize
        self.row_bytes = round_up(self.window_bytes, 16)
        # what is delivered: the whole row, or this rank's column slice of it (sequence parallelism)
        self.nValues, self.col_off = self.window_values, 0
        if seq_size > 1:
            if len(meta.splits) != 1 or self.window_values % seq_size:
                raise ValueError("sequence sharding needs single-field rows whose length divides by seq_size")
            self.nValues = self.window_values // seq_size
            self.col_off = seq_rank * self.nValues * self.np_dtype.itemsize
            if self.col_off % 16 or (self.nValues * self.np_dtype.itemsize) % 16:
                raise ValueError("sequence shards must be multiples of 16 bytes")
        self.sample_bytes = self.nValues * self.np_dtype.itemsize
        self.convert: Convert = meta.convert if meta.convert is not None else Convert.copy()
        self.splits = tuple(int(s) for s in meta.splits) if seq_size == 1 else (self.nValues,)
        self.var = bool(getattr(meta, "variable_length", False))
        self.replicate = bool(getattr(meta, "replicate", False))
        if self.convert.mode == N.CVT_COPY:
            self.out_dtype = _TORCH_OF_NP[self.np_dtype]
            self.out_elem = self.np_dtype.itemsize
        else:
            self.out_dtype = self.convert.out_dtype
            self.out_elem = self.convert.out_elem_size
            if self.convert.in_elem_size != self.np_dtype.itemsize:
                raise ValueError(f"convert expects {self.convert.in_elem_size}-byte elements, window has {self.np_dtype}")
        self.out_sample_bytes = self.nValues * self.out_elem
        self.dst_row_bytes = round_up(self.out_sample_bytes, 16)
        if sum(self.splits) != self.nValues:
            raise DoesNotMatchError((self.splits, self.nValues), "splits must add up to nValues")

    def key(self):
        return (self.nData, self.sample_bytes, self.row_bytes, str(self.np_dtype), self.convert.mode)


class CpuGather:
    """CPU execution of exchange_gather_collate (vectorised numpy/torch), same semantics."""

    def __init__(self, shards: Sequence[np.ndarray], spec: WindowSpec):
        self.shards, self.spec = shards, spec

    lengths = None  # [n_sources, n_max] valid elements per row (variable-size windows)

    def run(self, slots: np.ndarray, out_raw: torch.Tensor) -> None:
        sp = self.spec
        slots = slots.astype(np.int64)
        src = (slots >> SLOT_ROW_BITS) & 0xFFFF
        row = slots & SLOT_ROW_MASK
        raw = np.empty((slots.size, sp.sample_bytes), dtype=np.uint8)
        for s in np.unique(src):
            m = src == s
            raw[m] = self.shards[int(s)][row[m], sp.col_off:sp.col_off + sp.sample_bytes]
        cv = sp.convert
        if self.lengths is not None:
            self._run_padded(raw, self.lengths[src, row], out_raw)
            return
        dst = out_raw[:slots.size, :sp.out_sample_bytes]
        if cv.mode == N.CVT_COPY:
            dst.copy_(torch.from_numpy(raw))
            return
        in_np = {N.CVT_U8_BF16: np.uint8, N.CVT_U8_F32: np.uint8, N.CVT_F32_BF16: np.float32,
                 N.CVT_I32_I64: np.int32}[cv.mode]
        t = torch.from_numpy(raw.view(in_np))
        if cv.mode == N.CVT_I32_I64:
            y = t.to(torch.int64)
        else:
            n_elem = t.shape[1]
            nch = len(cv.scale)
            if nch > 1 and cv.chan_stride > 0:
                ch = (torch.arange(n_elem) // cv.chan_stride) % nch
                sc = torch.tensor(cv.scale, dtype=torch.float32)[ch]
                bi = torch.tensor(cv.bias, dtype=torch.float32)[ch]
            else:
                sc = torch.tensor(cv.scale[0], dtype=torch.float32)
                bi = torch.tensor(cv.bias[0], dtype=torch.float32)
            y = (t.to(torch.float32) * sc + bi).to(cv.out_dtype)
        dst.copy_(y.contiguous().view(torch.uint8).view(slots.size, -1))


def _cpu_run_padded(self, raw: np.ndarray, lens: np.ndarray, out_raw: torch.Tensor) -> None:
    """Variable-size rows on the CPU path: convert the valid prefix, pad the rest."""
    sp, cv = self.spec, self.spec.convert
    n = raw.shape[0]
    typed = torch.from_numpy(raw.view(sp.np_dtype))
    if cv.mode == N.CVT_COPY:
        y = typed.clone()
        pad = torch.tensor(cv.pad_value).to(y.dtype)
    elif cv.mode == N.CVT_I32_I64:
        y = typed.to(torch.int64)
        pad = torch.tensor(int(cv.pad_value), dtype=torch.int64)
    else:
        y = (typed.to(torch.float32) * cv.scale[0] + cv.bias[0]).to(cv.out_dtype)
        pad = torch.tensor(cv.pad_value, dtype=cv.out_dtype)
    mask = torch.arange(y.shape[1])[None, :] >= torch.from_numpy(lens.astype(np.int64))[:, None]
    y[mask] = pad
    out_raw[:n, :sp.out_sample_bytes].copy_(y.contiguous().view(torch.uint8).view(n, -1))


CpuGather._run_padded = _cpu_run_padded


class DeviceEngine:
    """Owns heap, shards, tables, RNG states, streams and gather plans of one consumer rank."""

    def __init__(self, env, config: LoaderConfig, specs: List[WindowSpec], batch_size: int):
        self.env, self.cfg, self.specs = env, config, specs
        self.device = torch.device(env.device)
        self.cuda = self.device.type == "cuda"
        self.P = len(specs)
        self.B = int(batch_size)
        self.n_instances = env.n_instances
        self.me = env.color
        self.comm = env.comm_nth_pusher if env.transport == "device" else env.comm_global
        # with dedicated storage GPUs the heap spans consumers and storage ranks; shards then live
        # on `env.storage_rank_of(instance)` instead of on the instance's own GPU
        self.remote_storage = getattr(env, "n_storage", 0) > 0
        self.heap_comm = env.comm_global if self.remote_storage else self.comm
        self.launches = 0
        self.tracer = Tracer(config.trace, self.cuda)

        # every instance must describe the same windows for the k-th producer (rows are swapped
        # between them); the reference would fail inside MPI with a truncation error instead
        keys = self.comm.allgather_obj([s.key() for s in specs])
        for k in range(self.P):
            if len({tuple(inst[k]) for inst in keys}) != 1:
                raise DoesNotMatchError([inst[k] for inst in keys],
                                        f"producer {k + 1}: window geometry differs between instances")

        self.n_max = max(s.nData for s in specs)
        self.heap = SymmetricHeap(self.heap_comm, self.device)
        # host residency: shards stay in pinned, node-shared host memory and every step's rows
        # are copied host->device (N9 stage_host) or pulled by the kernel over PCIe
        self.host_resident = config.residency == "host" and self.cuda and not self.remote_storage
        if self.remote_storage:
            self.shards = []
        elif self.host_resident:
            self.shards = [self.heap.alloc_host_shared(f"shard{k}_i{self.me}", s.nData * s.row_bytes)
                           for k, s in enumerate(specs)]
        else:
            self.shards = [self.heap.alloc(f"shard{k}_i{self.me}", s.nData * s.row_bytes, zero=False)
                           for k, s in enumerate(specs)]
        self.tables_buf = self.heap.alloc("tables", self.P * 2 * self.n_max * 8)
        self.flags_buf = self.heap.alloc("flags", N.FLAG_PAGE_BYTES)
        ring_row = max(s.dst_row_bytes for s in specs)
        self.ring_depth = config.ring_depth
        self.ring_buf = self._alloc_ring(self.ring_depth * self.B * ring_row)
        self.heap.publish()

        self.tables = self.tables_buf.tensor(torch.int64, (self.P, 2, self.n_max))
        self.states = torch.zeros((self.P, 6), dtype=torch.int64, device=self.device)
        self.cur = [0] * self.P  # buffer holding the newest produced iteration of window k
        self.ring_row = ring_row
        self.ring = [self.ring_buf.tensor(torch.uint8, (self.B, ring_row), i * self.B * ring_row)
                     for i in range(self.ring_depth)]

        if self.cuda:
            self.main = torch.cuda.current_stream(self.device)
            # table production: few, small kernels that everything else eventually waits for -> high
            # priority, so that they get the next free SM even while batch assembly fills the GPU
            self.side = torch.cuda.Stream(self.device, priority=-1)
            self.fetch_streams = [torch.cuda.Stream(self.device, priority=config.fetch_priority)
                                  for _ in range(config.fetch_streams)]  # batch assembly
            self.fetch = self.fetch_streams[0]
        else:
            self.main = self.side = self.fetch = NullStream()
            self.fetch_streams = [self.fetch]

        # source table: id g = instance * P + k  ->  that instance's shard k
        n_sources = self.n_instances * self.P
        where = env.storage_rank_of if self.remote_storage else (lambda i: i)
        ptrs = [self.heap.peer_ptr(where(g // self.P), f"shard{g % self.P}_i{g // self.P}") for g in range(n_sources)]
        self.src_ptrs = list(ptrs)  # this process's address of every shard (host address for host residency)
        self.plans: List = []
        self.stage = None
        if self.host_resident and config.host_stage in ("memcpy", "gather"):
            self.h_bases = (C.c_void_p * n_sources)(*ptrs)  # host addresses of every shard in this process
            stage_row = max(s.row_bytes for s in specs)
            self.stage = [torch.empty(self.B * stage_row, dtype=torch.uint8, device=self.device)
                          for _ in range(self.ring_depth)]
            self.stage_iota = torch.stack([torch.arange(self.B, dtype=torch.int64) + (i << SLOT_ROW_BITS)
                                           for i in range(self.ring_depth)]).to(self.device)
            self.host_tables = torch.empty((self.P, self.n_max), dtype=torch.int64).pin_memory()
            self.sources = SourceTable([t.data_ptr() for t in self.stage], self.device)
            for s in specs:
                self.plans.append(GatherPlan(self.sources, s.row_bytes, s.sample_bytes, s.convert,
                                             dst_row_bytes=ring_row, variant=config.gather_variant,
                                             max_ctas=config.gather_max_ctas, tune=config.gather_tune,
                                             col_offset_bytes=s.col_off))
            self.h2d_copies = 0
            self.pinned = None
            if config.host_stage == "gather":
                self.pinned = [torch.empty(self.B * stage_row, dtype=torch.uint8).pin_memory()
                               for _ in range(self.ring_depth)]
                self.pinned_done = [None] * self.ring_depth
        elif self.cuda:
            if self.host_resident:
                ptrs = [self.heap.peer_dev_ptr(g // self.P, f"shard{g % self.P}_i{g // self.P}") for g in range(n_sources)]
            self.sources = SourceTable(ptrs, self.device)
            for s in specs:
                self.plans.append(GatherPlan(self.sources, s.row_bytes, s.sample_bytes, s.convert,
                                             dst_row_bytes=ring_row, variant=config.gather_variant,
                                             max_ctas=config.gather_max_ctas, tune=config.gather_tune,
                                             col_offset_bytes=s.col_off))
        else:
            for k, s in enumerate(specs):
                views = [self.heap.peer_buffer(where(g // self.P), f"shard{g % self.P}_i{g // self.P}").numpy(
                    np.uint8, (specs[g % self.P].nData, specs[g % self.P].row_bytes)) for g in range(n_sources)]
                self.plans.append(CpuGather(views, s))
        self._status = None
        self._perm_scratch = None
        self.push = config.exchange == "push" and self.cuda and not self.host_resident
        if self.push:
            self.init_push()
        self._barrier_epoch = 0
        self._barrier_tensor = None
        self._distinct_gpus = None
        self._flag_pages = None
        self._closed = False

    def _alloc_ring(self, nbytes: int):
        """The batch ring.  For the push form it is made with the VMM API where possible, mapped by
        every peer and bound to an NVLS multicast object (replicated windows); otherwise cudaMalloc +
        CUDA IPC like every other heap buffer."""
        want = (self.cfg.exchange == "push" and self.cuda and self.cfg.residency != "host"
                and self.cfg.multicast != "off" and self.n_instances > 1)
        if want:
            try:
                return self.heap.alloc_vmm("ring", nbytes, multicast=True)
            except N.NativeError as e:
                # a box without VMM support raises before any collective step of alloc_vmm, on every
                # rank alike (the capability query is the same on one box): all ranks fall back together
                if self.cfg.multicast == "on":
                    raise
                log.info("VMM ring unavailable (%s); using a cudaMalloc ring", e)
        if self.cfg.multicast == "on" and self.cfg.exchange == "push" and self.n_instances > 1:
            raise RuntimeError("multicast='on' needs CUDA devices with VMM and NVLS multicast support")
        return self.heap.alloc("ring", nbytes, zero=False)

    # -- helpers --------------------------------------------------------------------------------
    def event(self):
        return NativeEvent() if self.cuda else NullEvent()

    @staticmethod
    def wait(stream, ev) -> None:
        stream_wait(stream, ev)

    def source_id(self, k: int, instance: Optional[int] = None) -> int:
        return (self.me if instance is None else instance) * self.P + k

    def table(self, k: int, buf: Optional[int] = None) -> torch.Tensor:
        return self.tables[k, self.cur[k] if buf is None else buf, :self.specs[k].nData]

    def table_ptr(self, k: int, buf: int, first: int = 0) -> int:
        return self.tables_buf.ptr + ((k * 2 + buf) * self.n_max + first) * 8

    # -- K1 shard_load ----------------------------------------------------------------------------
    def alloc_window(self, k: int) -> np.ndarray:
        """Pinned host window handed to the user's post_init (reference: the MPI window)."""
        s = self.specs[k]
        if self.host_resident and s.row_bytes == s.window_bytes:
            return self.shards[k].numpy(s.np_dtype, (s.nData, s.window_values))  # the window is the shard
        if self.cuda:
            buf, _ = self.heap.alloc_pinned(f"window{k}", s.nData * s.window_bytes)
            return buf.numpy(s.np_dtype, (s.nData, s.window_values))
        return np.zeros((s.nData, s.window_values), dtype=s.np_dtype)

    def load_shard(self, k: int, window: np.ndarray, stream=None) -> None:
        """Copy a filled host window into the HBM shard (row stride padded to 16 bytes)."""
        s = self.specs[k]
        if window.ctypes.data == self.shards[k].ptr:
            return
        src = torch.from_numpy(window.reshape(s.nData, -1).view(np.uint8))
        dst = self.shards[k].tensor(torch.uint8, (s.nData, s.row_bytes))
        st = stream or self.side
        if self.host_resident:
            dst[:, :s.window_bytes] = src
        elif self.cuda:
            # K1 shard_load: the window is pinned (cudaHostAlloc), so this is a true asynchronous
            # DMA on the side stream -- one cudaMemcpyAsync, or one strided copy when the shard's
            # rows are padded to 16 bytes
            lib, sp = N.lib(), C.c_void_p(st.cuda_stream)
            if s.row_bytes == s.window_bytes:
                N.check(lib.bddl_memcpy_async(C.c_void_p(self.shards[k].ptr), C.c_void_p(window.ctypes.data),
                                              s.nData * s.window_bytes, 1, sp))
            else:
                N.check(lib.bddl_memcpy2d_async(C.c_void_p(self.shards[k].ptr), s.row_bytes,
                                                C.c_void_p(window.ctypes.data), s.window_bytes, s.window_bytes,
                                                s.nData, 1, sp))
        else:
            dst[:, :s.window_bytes] = src

    def load_shard_from_device(self, k: int, data: torch.Tensor) -> None:
        """Fill shard k from a tensor that already lives on the device (synthetic benches)."""
        s = self.specs[k]
        dst = self.shards[k].tensor(torch.uint8, (s.nData, s.row_bytes))
        src = data.reshape(s.nData, -1).view(torch.uint8)
        with torch.cuda.stream(self.side) if self.cuda else NullStream():
            dst[:, :s.window_bytes].copy_(src)

    # -- variable-size windows ------------------------------------------------------------------------
    def set_lengths(self, k: int, lens_elems: np.ndarray) -> None:
        """Collective over the consumers: every rank learns every instance's row lengths of window
        k (4 bytes per sample), so the kernel finds a sample's length locally and only its valid
        bytes cross NVLink; the rest of the batch row is padded while collating."""
        s = self.specs[k]
        if lens_elems.shape != (s.nData,) or lens_elems.min() < 0 or lens_elems.max() > s.nValues:
            raise ValueError(f"lengths must be {s.nData} values in [0, {s.nValues}]")
        everyone = self.comm.allgather_obj(lens_elems.astype(np.int32))
        n_sources = self.n_instances * self.P
        if not hasattr(self, "lens_all"):
            self.lens_all = torch.zeros((n_sources, self.n_max), dtype=torch.int32)
        for i, l in enumerate(everyone):
            self.lens_all[i * self.P + k, :s.nData] = torch.from_numpy(l)
        if not self.cuda:
            self.plans[k].lengths = self.lens_all.numpy()
            return
        if self.stage is not None:
            raise NotImplementedError("variable-size windows use the kernel paths (HBM or host_stage='pull'), "
                                      "not the memcpy/gather staging")
        es = s.np_dtype.itemsize
        self.lens_all_dev = self.lens_all.to(self.device)                           # elements
        self._lens_dev = (self.lens_all * es).to(self.device)                       # payload bytes
        self._offs_dev = (torch.arange(self.n_max, dtype=torch.int64) * s.row_bytes).to(self.device)
        table = self.sources.with_lengths([self._offs_dev] * n_sources, [self._lens_dev[g] for g in range(n_sources)])
        self.plans[k] = GatherPlan(table, 0, s.sample_bytes, s.convert, dst_row_bytes=self.ring_row,
                                   variant=self.cfg.gather_variant, max_ctas=self.cfg.gather_max_ctas,
                                   tune=self.cfg.gather_tune)

    def batch_lengths(self, k: int, buf: int, first_row: int, n: int) -> torch.Tensor:
        """Valid elements of each sample of a batch (int32 [n]) for variable-size windows."""
        slots = self.tables[k, buf, first_row:first_row + n]
        src, row = (slots >> SLOT_ROW_BITS) & 0xFFFF, slots & SLOT_ROW_MASK
        lens = self.lens_all_dev if self.cuda else self.lens_all
        return lens[src, row]

    # -- table production (side stream) -----------------------------------------------------------
    def init_table(self, k: int, rng: Optional[Pcg64]) -> None:
        self.cur[k] = 0
        with torch.cuda.stream(self.side) if self.cuda else NullStream():
            T.slot_iota(self.table(k, 0), self.source_id(k), 0, stream=self.side if self.cuda else None)
            if rng is not None:
                self.states[k].copy_(torch.from_numpy(rng.to_words()))

    def stream_barrier(self) -> None:
        """All consumers' side streams meet here (orders table reads/writes across GPUs).

        Default on GPUs: the flag-barrier kernel (csrc/flags.cu) -- peer stores into every rank's
        flag page + a bounded acquire-poll, no library collective on the shuffle path.  It needs one
        GPU per rank (kernels of several ranks on one GPU must not wait on one another); ranks that
        share a GPU, and CPU runs, meet on the host instead."""
        if self.n_instances == 1:
            return
        mode = self.cfg.barrier
        if mode == "auto":
            mode = "flags" if self.cuda and self._one_gpu_per_rank() else "host"
        if mode == "nccl":
            import torch.distributed as dist

            if self._barrier_tensor is None:
                self._barrier_tensor = torch.zeros(1, device=self.device)
            with torch.cuda.stream(self.side):
                dist.all_reduce(self._barrier_tensor, group=self.env.loader_group or self.env.consumer_group)
        elif mode == "flags":
            self._flag_barrier()
        else:
            self.side.synchronize()
            self.comm.barrier()

    def _one_gpu_per_rank(self) -> bool:
        if self._distinct_gpus is None:
            import os

            props = torch.cuda.get_device_properties(self.device)
            gpu = str(getattr(props, "uuid", "")) or f"{os.getpid()}:{self.device.index}"
            ids = self.comm.allgather_obj((os.uname().nodename, gpu))
            self._distinct_gpus = len({str(i) for i in ids}) == len(ids)
        return self._distinct_gpus

    def _status_word(self):
        """Mapped host word that bounded device-side waits write when they give up (1 timed out,
        2 stopped); read by the host without synchronising."""
        if self._status is None:
            self._status_buf, self._status_dev = self.heap.alloc_pinned("wait_status", 64)
            self._status = self._status_buf.numpy(np.uint32, (16,))
            self._status[:] = 0
        return self._status

    def check_waits(self) -> None:
        if self._status is not None and self._status[0]:
            from .exceptions import LoaderTimeoutError

            code = int(self._status[0])
            what = {1: "timed out", 2: "was stopped"}.get(code & 15, "failed")
            peer = (code >> 8) & 255
            which = {0: "a peer's flag", 1: f"the `free` word of consumer {peer}",
                     2: f"the `table` word of consumer {peer}", 3: "this rank's `ready` word"}.get((code >> 4) & 15)
            which += f" (block {code >> 16})"
            raise LoaderTimeoutError(f"a bounded device-side wait on {which} {what} (timeout_s={self.cfg.timeout_s}); "
                                     f"flag page: {self.flag_page_summary()}")

    def flag_page_summary(self) -> str:
        """This rank's counters, copied out on a private stream (diagnostics after a failed wait)."""
        try:
            st = torch.cuda.Stream(self.device)
            with torch.cuda.stream(st):
                w = self.flags_buf.tensor(torch.int32, (N.FLAG_PAGE_BYTES // 4,)).to("cpu", non_blocking=True)
            st.synchronize()
            R, D = self.n_instances, self.ring_depth
            M = N.FLAG_MAX_RANKS
            return str({"ready": w[N.FLAG_READY0:N.FLAG_READY0 + D].tolist(),
                        "free": [w[N.FLAG_FREE0 + s * M:N.FLAG_FREE0 + s * M + R].tolist() for s in range(D)],
                        "done": [w[N.FLAG_DONE0 + s * M:N.FLAG_DONE0 + s * M + R].tolist() for s in range(D)],
                        "done_mc": w[N.FLAG_DONE_MC0:N.FLAG_DONE_MC0 + D].tolist(), "rank": self.me,
                        "table": [w[N.FLAG_TABLE0 + k * M:N.FLAG_TABLE0 + k * M + R].tolist() for k in range(self.P)],
                        "use": getattr(self, "_use", None)})
        except Exception as e:  # diagnostics must never mask the timeout itself
            return f"unavailable ({e})"

    def _flag_barrier(self) -> None:
        if self._flag_pages is None:
            pages = [self.heap.peer_ptr(r, "flags") for r in range(self.n_instances)]
            self._flag_pages = torch.tensor(pages, dtype=torch.int64).to(self.device)
            self._status_word()
        self.check_waits()
        self._barrier_epoch += 1
        N.check(N.lib().bddl_flag_barrier(self._flag_pages.data_ptr(), self.n_instances, self.me,
                                          self._barrier_epoch, int(self.cfg.timeout_s * 1e9), None, C.c_void_p(self._status_dev),
                                          C.c_void_p(self.side.cuda_stream)))
        self.launches += 1

    def exchange_tables(self, k: int, send_to: int, recv_from: int, num_exchange: int) -> None:
        """index_exchange for window k: next table = own with the leading rows taken from the
        partners' current tables (reference ddl/shuffle.py:92-108 in index space)."""
        cur, nxt = self.cur[k], 1 - self.cur[k]
        n = self.specs[k].nData
        active = num_exchange > 0 and self.n_instances > 1
        if active:
            self.stream_barrier()  # partners' current tables are complete; nobody reads `nxt` any more
        own = self.table(k, cur)
        dst = self.table(k, nxt)
        if self.cuda:
            fr = self.heap.peer_ptr(recv_from, "tables") + ((k * 2 + cur) * self.n_max) * 8 if active else None
            fs = self.heap.peer_ptr(send_to, "tables") + ((k * 2 + cur) * self.n_max) * 8 if active else None
            with torch.cuda.stream(self.side):
                T.index_exchange(dst, own, None, None, num_exchange if active else 0, stream=self.side,
                                 from_recv_ptr=fr, from_send_ptr=fs)
            self.launches += 1
        else:
            if active:
                fr = self.heap.peer_buffer(recv_from, "tables").tensor(torch.int64, (self.P, 2, self.n_max))[k, cur, :n]
                fs = self.heap.peer_buffer(send_to, "tables").tensor(torch.int64, (self.P, 2, self.n_max))[k, cur, :n]
                T.index_exchange(dst, own, fr, fs, num_exchange)
            else:
                T.index_exchange(dst, own, None, None, 0)
        self.cur[k] = nxt

    def copy_forward(self, k: int) -> None:
        self.exchange_tables(k, self.me, self.me, 0)

    def local_shuffle(self, k: int) -> None:
        """perm_gen on the newest table of window k (reference tests/run_ddl.py:163-167)."""
        n = self.specs[k].nData
        tbl = self.tables[k, self.cur[k]].unsqueeze(0)
        st = self.states[k].unsqueeze(0)
        if self.cuda:
            if self._perm_scratch is None:  # one scratch for all windows: they shuffle in turn on the side stream
                self._perm_scratch = torch.empty(int(N.lib().bddl_perm_gen_scratch_bytes(self.n_max)) + 64,
                                                 dtype=torch.uint8, device=self.device)
            with torch.cuda.stream(self.side):
                T.perm_gen(tbl, st, n=n, stream=self.side, scratch=self._perm_scratch)
            self.launches += 1
        else:
            T.perm_gen(tbl, st, n=n)

    def upload_table(self, k: int, slots: np.ndarray) -> None:
        """Install a host-computed table (trace mode / checkpoint restore) into the next buffer."""
        nxt = 1 - self.cur[k]
        t = torch.from_numpy(np.ascontiguousarray(slots, dtype=np.int64))
        with torch.cuda.stream(self.side) if self.cuda else NullStream():
            self.table(k, nxt).copy_(t, non_blocking=True)
        self.cur[k] = nxt

    # -- batch assembly (fetch stream) --------------------------------------------------------------
    def gather(self, k: int, buf: int, first_row: int, ring_slot: int, n: Optional[int] = None, stream=None) -> None:
        n = self.B if n is None else n
        stream = stream or self.fetch
        if self.stage is not None:
            # N9 stage_host: this step's rows, pinned host -> device, then the convert kernel
            s = self.specs[k]
            copies = C.c_int64()
            slots_ptr = self.host_tables.data_ptr() + (k * self.n_max + first_row) * 8
            lib = N.lib()
            if self.pinned is not None:
                ev = self.pinned_done[ring_slot]
                if ev is not None:
                    ev.synchronize()  # the previous transfer out of this pinned buffer has finished
                N.check(lib.bddl_host_gather_slots(C.c_void_p(self.pinned[ring_slot].data_ptr()), s.row_bytes,
                                                   self.h_bases, len(self.h_bases), s.row_bytes, C.c_void_p(slots_ptr),
                                                   n, s.row_bytes, self.cfg.host_threads))
                N.check(lib.bddl_memcpy_async(C.c_void_p(self.stage[ring_slot].data_ptr()),
                                              C.c_void_p(self.pinned[ring_slot].data_ptr()), n * s.row_bytes, 1,
                                              C.c_void_p(stream.cuda_stream)))
                if ev is None:
                    ev = self.pinned_done[ring_slot] = torch.cuda.Event()
                ev.record(stream)
                copies.value = 1
            elif self.cfg.host_threads > 1:
                N.check(lib.bddl_stage_slots_mt(C.c_void_p(self.stage[ring_slot].data_ptr()), s.row_bytes,
                                                self.h_bases, len(self.h_bases), s.row_bytes, C.c_void_p(slots_ptr),
                                                n, s.row_bytes, C.c_void_p(stream.cuda_stream),
                                                self.cfg.host_threads))
                copies.value = n
            else:
                N.check(lib.bddl_stage_slots(C.c_void_p(self.stage[ring_slot].data_ptr()), s.row_bytes, self.h_bases,
                                             len(self.h_bases), s.row_bytes, C.c_void_p(slots_ptr), n, s.row_bytes,
                                             C.c_void_p(stream.cuda_stream), C.byref(copies)))
            self.h2d_copies += copies.value
            self.plans[k].run_ptr(self.stage_iota[ring_slot].data_ptr(), self.ring[ring_slot].data_ptr(), n,
                                  stream.cuda_stream)
            self.launches += 1
        elif self.cuda:
            self.plans[k].run_ptr(self.table_ptr(k, buf, first_row), self.ring[ring_slot].data_ptr(), n,
                                  stream.cuda_stream)
            self.launches += 1
        else:
            slots = self.tables[k, buf, first_row:first_row + n].numpy()
            self.plans[k].run(slots, self.ring[ring_slot])

    # -- push form of the exchange (owner-side stores over NVLink + in-kernel flags) ---------------------
    # One launch per step and GPU (csrc/gather.cu, push_tma_cvt_kernel): every rank scans every
    # consumer's batch slice, delivers the samples it owns into that consumer's ring slot with
    # stores over NVLink and bumps the consumer's `ready` word from inside the kernel
    # (red.release.sys); the same kernel first waits for the consumer's `free` and `table` words and
    # finally for its own `ready` word -- the reference's lend/return tokens
    # (ddl/connection.py:153-182) as NVLink-visible counters, all waits bounded.
    def init_push(self) -> None:
        R, D = self.n_instances, self.ring_depth
        if R > N.FLAG_MAX_RANKS or D > N.FLAG_MAX_SLOTS or self.P > N.FLAG_MAX_WINDOWS:
            raise ValueError(f"push exchange supports up to {N.FLAG_MAX_RANKS} instances, ring_depth <= "
                             f"{N.FLAG_MAX_SLOTS} and {N.FLAG_MAX_WINDOWS} producers per instance")
        if self.remote_storage:
            raise NotImplementedError("the push form needs the shards on the consumer GPUs (no storage ranks)")
        self._use = [0] * D
        self._fills = {False: [0] * D, True: [0] * D}  # per slot: launches so far, unicast / multicast
        self.mc_ring = int(getattr(self.ring_buf, "mc_ptr", 0) or 0)
        if self.cfg.multicast == "on" and R > 1 and not self.mc_ring:
            raise RuntimeError(f"multicast='on' but the rings could not be bound to a multicast object: "
                               f"{getattr(self.ring_buf, 'mc_error', 'unsupported')}")
        flag_pages = [self.heap.peer_ptr(r, "flags") for r in range(R)]
        dev = lambda xs: torch.tensor(xs, dtype=torch.int64).to(self.device)  # noqa: E731
        self._push_tables = dev([self.heap.peer_ptr(r, "tables") for r in range(R)])
        self._push_rings = dev([self.heap.peer_ptr(r, "ring") for r in range(R)])
        self._push_flags = dev(flag_pages)
        # consumer -> owners: "I consumed fill number `use` of my ring slot"
        self._free_targets = dev([[p + 4 * (N.FLAG_FREE0 + s * N.FLAG_MAX_RANKS + self.me) for p in flag_pages]
                                  for s in range(D)])                           # [D, R]
        # consumer -> owners: "my slot table number `seq` of window k is complete"
        self._table_targets = dev([[p + 4 * (N.FLAG_TABLE0 + k * N.FLAG_MAX_RANKS + self.me) for p in flag_pages]
                                   for k in range(self.P)])                     # [P, R]
        self._status_word()  # bounded waits report here
        self.signal = torch.cuda.Stream(self.device, priority=-1)  # `free` tokens, see release_slot
        g = C.c_int32()
        N.check(N.lib().bddl_push_ctas_per_consumer(R, self.cfg.gather_max_ctas, C.byref(g)))
        d = N.PushDesc()
        d.d_tables, d.d_rings, d.d_flags = (self._push_tables.data_ptr(), self._push_rings.data_ptr(),
                                            self._push_flags.data_ptr())
        d.d_status = self._status_dev
        d.timeout_ns = int(self.cfg.timeout_s * 1e9)
        d.n_ranks, d.me, d.ctas_per_consumer = R, self.me, g.value
        d.owner_lo, d.owner_hi = self.me * self.P, (self.me + 1) * self.P
        self._push_desc = d

    def signal_table_ready(self, k: int, seq: int) -> None:
        """Side stream, after window k's table number ``seq`` was produced: tell every owner."""
        N.check(N.lib().bddl_flag_signal(C.c_void_p(self._table_targets[k].data_ptr()), self.n_instances, seq, 0,
                                         C.c_void_p(self.side.cuda_stream)))
        self.launches += 1

    def push_step(self, k: int, buf: int, first_row: int, ring_slot: int, stream, table_seq: int) -> None:
        self.check_waits()
        d = self._push_desc
        d.table_offset_bytes = ((k * 2 + buf) * self.n_max + first_row) * 8
        d.ring_offset_bytes = ring_slot * self.B * self.ring_row
        d.ring_slot, d.window = ring_slot, k
        d.use, d.table_seq = self._use[ring_slot], table_seq
        rep = self.specs[k].replicate
        mc = rep and self.mc_ring != 0 and self._mc_ok(k)
        d.table_rank = 0 if rep else -1          # replicated window: everybody receives instance 0's
        d.d_mc_ring = self.mc_ring if mc else None
        d.fills = self._fills[mc][ring_slot]
        self.plans[k].push_ptr(d, self.B, stream.cuda_stream)
        self._fills[mc][ring_slot] += 1
        self._use[ring_slot] += 1
        self.launches += 1

    def _mc_ok(self, k: int) -> bool:
        """multimem.st moves whole 16-byte groups: fixed-size rows whose output is a multiple of 16 bytes."""
        s = self.specs[k]
        group_in = {N.CVT_COPY: 16, N.CVT_U8_BF16: 8, N.CVT_F32_BF16: 32, N.CVT_U8_F32: 4, N.CVT_I32_I64: 8}
        return not s.var and s.sample_bytes % group_in[s.convert.mode] == 0

    def release_slot(self, ring_slot: int, consumed_event) -> None:
        """Consumer side of the push form: tell every owner that `ring_slot` may be refilled.

        The signal runs on its own HIGH-priority stream behind the slot's `consumed` event, not on
        the training stream.  Push kernels wait for this word from inside the kernel while they
        occupy SMs, and the block dispatcher serves grids of one priority strictly in order: a
        one-block signal kernel of normal priority can sit behind a partially dispatched push kernel
        whose blocks wait for that very signal (measured: 3 fetch streams deadlocked this way).  A
        higher-priority grid is dispatched first."""
        stream_wait(self.signal, consumed_event)
        N.check(N.lib().bddl_flag_signal(C.c_void_p(self._free_targets[ring_slot].data_ptr()), self.n_instances,
                                         self._use[ring_slot], 0, C.c_void_p(self.signal.cuda_stream)))
        self.launches += 1

    # -- integrity mode -----------------------------------------------------------------------------------
    def verify_batch(self, k: int, buf: int, first_row: int, ring_slot: int, stream) -> torch.Tensor:
        """Assemble the same batch again with the LDG form into a scratch buffer and return the
        two fingerprints ([2] int64 on the device); they must be equal."""
        from .ops.checksum import BatchChecksum

        s = self.specs[k]
        if not hasattr(self, "_verify"):
            self._verify = {}
        if k not in self._verify:
            base = self.plans[k]
            plan = GatherPlan(base.sources, s.row_bytes if not s.var else 0, s.sample_bytes, s.convert,
                              dst_row_bytes=self.ring_row, variant=N.VARIANT_LDG, col_offset_bytes=s.col_off)
            scratch = torch.zeros((self.B, self.ring_row), dtype=torch.uint8, device=self.device)
            payload = max(8, s.out_sample_bytes // 8 * 8)
            self._verify[k] = (plan, scratch, BatchChecksum(self.device, self.B, payload),
                               BatchChecksum(self.device, self.B, payload), payload)
        plan, scratch, ca, cb, payload = self._verify[k]
        plan.run_ptr(self.table_ptr(k, buf, first_row), scratch.data_ptr(), self.B, stream.cuda_stream)
        a = ca.run(self.ring[ring_slot], payload, stream=stream)
        b = cb.run(scratch, payload, stream=stream)
        with torch.cuda.stream(stream):
            out = torch.cat([a, b])
        self.launches += 3
        return out

    def reference_rows(self, k: int, slots: np.ndarray) -> torch.Tensor:
        """The raw bytes of the samples named by ``slots`` ([n, sample_bytes] uint8 on the engine's
        device), fetched without any kernel of this library: one ``cudaMemcpyAsync`` per row from
        the owning shard (peer mapping, local HBM or pinned host memory).  The independent half of
        ``DistributedDataLoader.reference_batch``."""
        s = self.specs[k]
        slots = np.asarray(slots).astype(np.int64)
        src, row = (slots >> SLOT_ROW_BITS) & 0xFFFF, slots & SLOT_ROW_MASK
        if not self.cuda:
            out = np.empty((slots.size, s.sample_bytes), dtype=np.uint8)
            for i in range(slots.size):
                out[i] = self.plans[k].shards[int(src[i])][int(row[i]), s.col_off:s.col_off + s.sample_bytes]
            return torch.from_numpy(out)
        out = torch.empty((slots.size, s.sample_bytes), dtype=torch.uint8, device=self.device)
        st = torch.cuda.current_stream(self.device)
        lib, sp = N.lib(), C.c_void_p(st.cuda_stream)
        for i in range(slots.size):
            a = self.src_ptrs[int(src[i])] + int(row[i]) * s.row_bytes + s.col_off
            N.check(lib.bddl_memcpy_async(C.c_void_p(out[i].data_ptr()), C.c_void_p(a), s.sample_bytes, 4, sp))
        return out

    def on_lend(self, k: int, buf: int) -> None:
        """Called when the consumer borrows window k: host staging needs the table on the host."""
        if self.stage is not None:
            n = self.specs[k].nData
            with torch.cuda.stream(self.fetch):
                self.host_tables[k, :n].copy_(self.tables[k, buf, :n], non_blocking=True)
            self.fetch.synchronize()

    def batch_views(self, k: int, ring_slot: int):
        """Typed [B, nValues] view of a ring slot, split into the sample's fields (dim=1)."""
        s = self.specs[k]
        raw = self.ring[ring_slot]
        typed = raw if s.out_dtype == torch.uint8 else raw.view(s.out_dtype)
        typed = typed[:, :s.nValues]
        return torch.split(typed, list(s.splits), dim=1)

    # -- teardown -------------------------------------------------------------------------------------
    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self.cuda:
            torch.cuda.synchronize(self.device)
        self.heap_comm.barrier()  # storage ranks keep their shards mapped until every consumer is done
        self.tables = None
        self.ring = []
        self.plans = []
        self.heap.close()


def _nccl_ready() -> bool:
    try:
        import torch.distributed as dist

        return dist.is_initialized() and dist.is_nccl_available() and "nccl" in str(dist.get_backend()).lower()
    except Exception:
        return False
