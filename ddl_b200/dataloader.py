"""Consumer role: ``DistributedDataLoader`` (reference ddl/mpi_dataloader.py:31-249).

Same constructor, ``len``/``[]``/``mark`` protocol and producer rotation as the reference:
consumer epoch ``e`` reads the window of producer ``(e mod P) + 1`` after that producer's
iteration ``e div P``; batch ``b`` is rows ``[b*B, (b+1)*B)`` of the window as currently
permuted; the ``nData mod B`` tail is dropped.  What differs is where a batch
comes from:

* device transport -- one ``exchange_gather_collate`` launch per batch on a fetch stream, into a
  ring of batch buffers in HBM, prefetched one batch ahead of the training stream.  The tensors
  handed out are views of a ring slot and stay valid until ``ring_depth - 1`` further batches
  have been marked.
* process transport on a GPU -- the borrowed host window is pinned (cudaHostRegister) and the
  batch's rows are staged with one ``cudaMemcpyAsync`` (or pulled by the kernel) on the fetch
  stream, then converted by the same kernel.
* process transport on the CPU -- zero-copy views of the shared window, exactly the reference.

Unlike the reference a single-rank world works (reference ddl/mpi_dataloader.py:173-174 sets
``len == 0`` and then fails in ``mark``).
"""
from __future__ import annotations

import ctypes as C
import logging
import time
from abc import ABC, abstractmethod
from typing import List, Optional, Tuple

import torch

from . import _native as N
from .config import LoaderConfig
from .exceptions import ShutdownError
from .connection import Connection
from .datapusher import DevicePusher
from .datasetwrapper import ProducerFunctionSkeleton
from .engine import DeviceEngine, NullStream, WindowSpec, stream_wait
from .ops.gather import GatherPlan, SourceTable, make_slot
from .types import Marker, MetaData_Consumer_To_Producer, MetaData_Producer_To_Consumer
from .utils import for_all_methods, with_logging
from .utils.metrics import LoaderMetrics

log = logging.getLogger("ddl_b200")

__all__ = ["DistributedDataLoader", "ProducerFunctionSkeleton", "Marker"]


class DistributedDataloaderABC(ABC):
    @abstractmethod
    def __len__(self): ...

    @abstractmethod
    def __getitem__(self, index): ...

    @abstractmethod
    def _advance_to_next_producer(self): ...

    @abstractmethod
    def _finalize(self): ...

    @abstractmethod
    def _start_access_epoch(self, target_rank: int) -> None: ...

    @abstractmethod
    def _end_access_epoch(self, target_rank: int) -> None: ...

    @abstractmethod
    def _can_continue(self): ...

    @abstractmethod
    def mark(self, mark: Marker): ...


@for_all_methods(with_logging, exclude=["__getitem__", "__iter__", "__len__", "mark", "_on_batch_end"])
class DistributedDataLoader(DistributedDataloaderABC):
    def __init__(self, producer_function: ProducerFunctionSkeleton, batch_size: int, connection: Optional[Connection],
                 n_epochs: int, fraction_exchange: float = 0.0, exchange_method: str = "sendrecv_replace",
                 instance_idx: int = 0, n_instances: int = 1, config: Optional[LoaderConfig] = None):
        self.epoch = 0
        self.batch = 0
        self.target_rank = 1
        self.n_epochs = n_epochs
        self.batch_size = int(batch_size)
        self._finalized = False
        self._lend_seq = 0
        self._pref: dict = {}
        self._fast = False
        self._verify_pending: list = []
        self._cur_slot = -1
        self._next_slot = 0
        self._canary = None

        if connection is None:  # single process, no decorator: build the one-rank world here
            from .parallel.topology import init_env

            env = init_env(1)
            connection = Connection(env.comm_per_gpu_shm, env, config)
        self.connection = connection
        self.env = connection.env
        self.cfg = config or connection.config
        if config is not None:
            connection.config = config
        self.transport = connection.transport
        self.device = torch.device(self.env.device) if self.env is not None else torch.device("cpu")
        self.cuda = self.device.type == "cuda"
        self.metrics = LoaderMetrics(enabled=self.cfg.metrics)

        self.metadata_to_producer = MetaData_Consumer_To_Producer(
            producer_function=producer_function, global_shuffle_fraction_exchange=fraction_exchange,
            global_shuffle_exchange_method=exchange_method, batch_size=self.batch_size)
        connection.send_metadata(self.metadata_to_producer, "consumer")

        if self.transport == "device":
            self._init_device()
        else:
            self._init_process()

        self.mode = "do_not_split_along_epoch"  # the only logically correct mode (reference :149-153)
        self._len = self.batches_per_window[0]
        assert self._len > 0
        self._start_access_epoch()
        for d in range(self.cfg.prefetch_depth):
            self._prefetch(d)

    # -- construction: device transport -------------------------------------------------------------------
    def _init_device(self) -> None:
        conn, env = self.connection, self.env
        remote = [None] * env.n_producers
        if getattr(env, "n_storage", 0) > 0:
            # the storage rank runs on_init/post_init for this instance's producers and answers
            # with their metadata (the reference's producer processes do the same)
            srank = env.storage_rank_of(env.color)
            env.comm_global.send_obj(self.metadata_to_producer, srank, tag=31)
            remote = env.comm_global.recv_obj(srank, tag=32)
        pushers = [DevicePusher(conn, env, k + 1, remote_meta=remote[k]) for k in range(env.n_producers)]
        self.metadata_from_producer: List[MetaData_Producer_To_Consumer] = conn.recv_metadata_as_consumer()
        self.splits = [m.splits for m in self.metadata_from_producer]
        self.batches_per_window = [m.batches_per_window for m in self.metadata_from_producer]
        specs = [WindowSpec(m, self.cfg.seq_rank, self.cfg.seq_size) for m in self.metadata_from_producer]
        self.splits = [s.splits for s in specs]
        if any(s.replicate for s in specs) and self.cfg.exchange != "push" and env.n_instances > 1:
            raise ValueError("replicated windows (replicate=True) are delivered by the push form: "
                             "use LoaderConfig(exchange='push')")
        self.engine = DeviceEngine(env, self.cfg, specs, self.batch_size)
        conn.attach(self.engine, pushers)
        for p in pushers:
            p.bind(self.engine)
        eng = self.engine
        eng.side.synchronize()
        eng.heap_comm.barrier()  # every shard is resident before anyone's kernel may pull from it
        for p in pushers:
            p.start()
        self._consumed = [eng.event() if eng.cuda else None for _ in range(eng.ring_depth)]
        self._done = [eng.event() for _ in range(eng.ring_depth)]
        self._ring_ptr = [r.data_ptr() for r in eng.ring]
        # the one-call hot path covers the product's main configuration; everything else
        # (push form, host staging, tracing, integrity mode, metrics) takes the general path
        self._fast = (eng.cuda and not eng.push and eng.stage is None and not self.cfg.trace
                      and not self.cfg.verify and not self.cfg.canary)
        self._canary = (torch.zeros((eng.ring_depth, 2), dtype=torch.int32, device=self.device)
                        if self.cfg.canary else None)
        self._views = {}
        self.main = eng.main
        self.fetch = eng.fetch
        self.fetch_streams = eng.fetch_streams

    # -- construction: process transport --------------------------------------------------------------------
    def _init_process(self) -> None:
        conn = self.connection
        self.metadata_from_producer = conn.recv_metadata_as_consumer()
        self.splits = [m.splits for m in self.metadata_from_producer]
        self.batches_per_window = [m.batches_per_window for m in self.metadata_from_producer]
        self.specs = [WindowSpec(m) for m in self.metadata_from_producer]
        shapes = [tuple(m.shape) for m in self.metadata_from_producer]
        dtypes = [m.dtype for m in self.metadata_from_producer]
        self.arys = conn.init_windows(shapes, dtypes)
        conn.lock_windows()
        conn.Barrier()
        self.engine = None
        self.cuda = self.device.type == "cuda"
        depth = self.cfg.ring_depth
        self._consumed = [None] * depth
        if not self.cuda:
            self.main = self.fetch = NullStream()
            self.fetch_streams = [self.fetch]
            return
        self.main = torch.cuda.current_stream(self.device)
        self.fetch_streams = [torch.cuda.Stream(self.device, priority=self.cfg.fetch_priority)
                              for _ in range(self.cfg.fetch_streams)]
        self.fetch = self.fetch_streams[0]
        lib = N.lib()
        self._win_dev, self._win_dev_alt, self._registered = [], [], []
        for primary, alt in zip(conn.wins, conn.wins_alt):  # pin + map every borrowed window once
            for win, out in ((primary, self._win_dev), (alt, self._win_dev_alt)):
                if win is None:
                    out.append(0)
                    continue
                d = C.c_void_p()
                N.check(lib.bddl_host_register(C.c_void_p(win[0]), win[1], C.byref(d)))
                self._registered.append(win[0])
                out.append(int(d.value))
        B = self.batch_size
        stage_row = max(s.sample_bytes for s in self.specs)
        out_row = max(s.dst_row_bytes for s in self.specs)
        self._stage = [torch.empty(B * stage_row, dtype=torch.uint8, device=self.device) for _ in range(depth)]
        self._out = [torch.empty((B, out_row), dtype=torch.uint8, device=self.device) for _ in range(depth)]
        self._plans = []
        for k, s in enumerate(self.specs):
            kernel_ok = s.sample_bytes % 16 == 0
            if s.convert.mode == N.CVT_COPY and not self.cfg.host_pull:
                self._plans.append(None)  # the staged bytes are the batch
                continue
            if not kernel_ok:
                raise ValueError("converting windows need rows that are a multiple of 16 bytes")
            if self.cfg.host_pull:
                # source 0 = primary window, source 1 = the producer's second window (if any)
                table = SourceTable([self._win_dev[k], self._win_dev_alt[k] or self._win_dev[k]], self.device)
                slots = torch.stack([torch.arange(s.nData, dtype=torch.int64) + make_slot(w, 0)
                                     for w in (0, 1)]).to(self.device)
            else:
                table = SourceTable([t.data_ptr() for t in self._stage], self.device)
                slots = torch.stack([torch.arange(B, dtype=torch.int64) + make_slot(i, 0) for i in range(depth)]).to(self.device)
            plan = GatherPlan(table, s.sample_bytes, s.sample_bytes, s.convert, dst_row_bytes=out_row,
                              variant=self.cfg.gather_variant, max_ctas=self.cfg.gather_max_ctas,
                              tune=self.cfg.gather_tune)
            self._plans.append((plan, slots))

    # -- reference protocol ---------------------------------------------------------------------------------
    def __len__(self) -> int:
        return self._len

    def __iter__(self):
        i, self._iter_start = getattr(self, "_iter_start", 0), 0
        while i < self._len:
            yield self[i]
            i += 1

    def __getitem__(self, idx: int) -> Tuple[torch.Tensor, ...]:
        if idx >= self._len:
            raise IndexError
        if idx < 0:
            raise ValueError
        if self._finalized:
            raise ShutdownError("the loader was closed (all epochs consumed, or close() was called)")
        k = self.target_rank - 1
        b = idx % self.batches_per_window[k]
        if self.transport == "process" and not self.cuda:
            # zero-copy views of the shared window (reference ddl/mpi_dataloader.py:190-198)
            start = self.batch_size * b
            win = self.connection.arys_alt[k] if self.connection.current_window(k + 1) else self.arys[k]
            data = torch.from_numpy(win[start:start + self.batch_size])
            self.metrics.batch(self.batch_size, data.numel() * data.element_size(), 0)
            return torch.split(data, list(self.splits[k]), dim=1)
        hit = self._pref.pop((self._lend_seq, b), None)
        slot, ev = hit if hit is not None else self._launch(k, b)
        if self.metrics.enabled and self.metrics.timing:
            t0 = time.perf_counter()
            ev.synchronize()
            self.metrics.stall(time.perf_counter() - t0)
        stream_wait(self._train_stream(), ev)  # whichever stream the training step runs on
        self._cur_slot, self._cur_batch = slot, b
        if self._canary is not None:
            ev.synchronize()
            self._check_canary("handed out")
        if self.cfg.prefetch:
            for d in range(1, self.cfg.prefetch_depth + 1):
                if b + d < self.batches_per_window[k]:
                    self._prefetch(b + d)
        return self._batch_views(k, slot)

    def _train_stream(self):
        """The stream the user's step runs on *now* (``torch.cuda.stream(...)`` blocks, graph side
        streams): the batch hand-over and the ring-slot release are ordered on it."""
        return torch.cuda.current_stream(self.device) if self.cuda else self.main

    def _prefetch(self, b: int) -> None:
        if not self.cfg.prefetch or (self.transport == "process" and not self.cuda) or self._finalized:
            return
        key = (self._lend_seq, b)
        if key in self._pref or b >= self.batches_per_window[self.target_rank - 1]:
            return
        if len(self._pref) > self.cfg.prefetch_depth:  # stale entries of an abandoned window
            self._pref = {kk: v for kk, v in self._pref.items() if kk[0] == self._lend_seq}
        self._pref[key] = self._launch(self.target_rank - 1, b)

    def _launch(self, k: int, b: int):
        """Enqueue the assembly of batch ``b`` of window ``k`` on the fetch stream."""
        slot = self._next_slot
        self._next_slot = (slot + 1) % len(self._consumed)
        stream = self.fetch_streams[slot % len(self.fetch_streams)]
        B = self.batch_size
        if self._fast:
            # hot path: wait for the slot, assemble, record done -- one native call
            eng = self.engine
            done = self._done[slot]
            eng.plans[k].step_ptr(eng.table_ptr(k, self.connection.pushers[k].lent_buf, b * B), self._ring_ptr[slot],
                                  B, stream.cuda_stream, self._consumed[slot].handle, done.handle)
            eng.launches += 1
            if self.metrics.enabled:
                s = eng.specs[k]
                self.metrics.batch(B, B * s.sample_bytes, B * s.out_sample_bytes)
            return slot, done
        done = self._consumed[slot]
        if done is not None:
            stream_wait(stream, done)  # the training stream has finished with this ring slot
        if self.transport == "device":
            eng = self.engine
            if eng.tracer.enabled:
                with eng.tracer.span("gather", stream):
                    eng.gather(k, self.connection.pushers[k].lent_buf, b * B, slot, stream=stream)
            elif eng.push:
                pusher = self.connection.pushers[k]
                eng.push_step(k, pusher.lent_buf, b * B, slot, stream, pusher.lent_seq)
            else:
                eng.gather(k, self.connection.pushers[k].lent_buf, b * B, slot, stream=stream)
            if self.cfg.verify and eng.cuda and not eng.push and eng.stage is None:
                self._verify_pending.append((b, eng.verify_batch(k, self.connection.pushers[k].lent_buf, b * B,
                                                                 slot, stream)))
            if self._canary is not None:
                self._stamp(slot, b, stream)
            ev = self._done[slot]
            if eng.cuda:
                ev.record(stream)
            s = eng.specs[k]
        else:
            s = self.specs[k]
            lib = N.lib()
            fs = C.c_void_p(stream.cuda_stream)
            plan = self._plans[k]
            which = self.connection.current_window(k + 1)
            if self.cfg.host_pull:
                p, slots = plan
                p.run_ptr(slots[which].data_ptr() + b * B * 8, self._out[slot].data_ptr(), B, stream.cuda_stream)
            else:
                base = self.connection.wins_alt[k][0] if which else self.connection.wins[k][0]
                src = base + b * B * s.sample_bytes
                N.check(lib.bddl_memcpy_async(C.c_void_p(self._stage[slot].data_ptr()), C.c_void_p(src),
                                              B * s.sample_bytes, 1, fs))
                if plan is not None:
                    p, slots = plan
                    p.run_ptr(slots[slot].data_ptr(), self._out[slot].data_ptr(), B, stream.cuda_stream)
            ev = torch.cuda.Event()
            ev.record(stream)
        self.metrics.batch(B, B * s.sample_bytes, B * s.out_sample_bytes)
        return slot, ev

    def _batch_views(self, k: int, slot: int):
        if self.transport == "device":
            v = self._views.get((k, slot))
            if v is None:
                v = self._views[(k, slot)] = self.engine.batch_views(k, slot)
            return v
        s = self.specs[k]
        if self._plans[k] is None:
            typed = self._stage[slot][: self.batch_size * s.sample_bytes].view(s.out_dtype).view(self.batch_size, s.nValues)
        else:
            raw = self._out[slot]
            typed = (raw if s.out_dtype == torch.uint8 else raw.view(s.out_dtype))[:, : s.nValues]
        return torch.split(typed, list(s.splits), dim=1)

    def _start_access_epoch(self, target_rank: int = 0) -> None:
        self.connection.start_access_epoch(target_rank or self.target_rank)
        self._lend_seq += 1

    def _end_access_epoch(self, target_rank: int = 0) -> None:
        self.connection.end_access_epoch(target_rank or self.target_rank)

    def _can_continue(self) -> bool:
        return self.epoch < self.n_epochs

    def _advance_to_next_producer(self) -> None:
        self.target_rank += 1
        if self.target_rank > self.connection.n_producers:
            self.target_rank = 1
        self._len = self.batches_per_window[self.target_rank - 1]

    def _check_verified(self) -> None:
        pending, self._verify_pending = self._verify_pending, []
        for b, pair in pending:
            x, y = pair.tolist()  # synchronises: integrity mode is a debugging aid
            if x != y:
                raise RuntimeError(f"integrity check failed for batch {b}: TMA and LDG forms of "
                                   f"exchange_gather_collate disagree ({x:#x} != {y:#x})")

    # -- debug canaries -------------------------------------------------------------------------------------
    def _stamp(self, slot: int, b: int, stream) -> None:
        """(lend number, batch index) into the slot's canary words, ordered behind its assembly."""
        want = (self._lend_seq & 0x7FFFFFFF, b)
        if self.cuda:
            base = self._canary[slot].data_ptr()
            for i, v in enumerate(want):
                N.check(N.lib().bddl_stream_write_value32(C.c_void_p(stream.cuda_stream), C.c_void_p(base + 4 * i), v))
        else:
            self._canary[slot, 0], self._canary[slot, 1] = want

    def _check_canary(self, when: str) -> None:
        if self.cuda:
            self._train_stream().synchronize()
        got = tuple(int(v) for v in self._canary[self._cur_slot].tolist())
        want = (self._lend_seq & 0x7FFFFFFF, self._cur_batch)
        if got != want:
            raise RuntimeError(f"canary: ring slot {self._cur_slot} holds (lend, batch) = {got} but {want} was "
                               f"{when}: the slot was refilled while in use, or handed out before it was assembled")

    def _on_batch_end(self) -> None:
        if self._verify_pending:
            self._check_verified()
        if self._canary is not None and self._cur_slot >= 0:
            self._check_canary("being marked")
        if self._cur_slot >= 0 and not isinstance(self.main, NullStream):
            ev = self._consumed[self._cur_slot]
            if ev is None:
                ev = self._consumed[self._cur_slot] = torch.cuda.Event()
            ts = self._train_stream()
            ev.record(ts)  # the "free" flag of this ring slot, in stream-event form
            if self.engine is not None and self.engine.push:
                self.engine.release_slot(self._cur_slot, ev)
        self.batch += 1
        if self.batch % self.batches_per_window[self.target_rank - 1] == 0:
            self._window_done = True  # every batch of this epoch's window was delivered
            if self.transport == "process" and self.cuda:
                # the producer may overwrite the window the moment it is returned: every copy or
                # zero-copy pull that still reads it must have finished on the device first
                for fs in self.fetch_streams:
                    fs.synchronize()
            self._end_access_epoch()
            self._advance_to_next_producer()
            self._start_access_epoch()
            for d in range(self.cfg.prefetch_depth):
                self._prefetch(d)

    def _on_epoch_end(self) -> None:
        self.batch = 0
        self._window_done = False
        self.epoch += 1
        self.metrics.epoch()
        if not self._can_continue():
            self.connection.shutdown_operation()
            self._finalize()

    def mark(self, mark: Marker) -> None:
        if mark is Marker.END_OF_BATCH:
            self._on_batch_end()
        elif mark is Marker.END_OF_EPOCH:
            self._on_epoch_end()
        else:
            raise ValueError(f"Unknown mark {mark}")

    def _finalize(self) -> None:
        if self._finalized:
            return
        self._finalized = True
        self._pref = {}
        if self.transport == "device":
            self._views = {}
            self.connection.close()
            self.engine = None
            return
        if self.cuda:
            torch.cuda.synchronize(self.device)
            for ptr in self._registered:
                N.lib().bddl_host_unregister(C.c_void_p(ptr))
            self._stage, self._out, self._plans = [], [], []
        self.connection.Barrier()
        self.connection.unlock_windows()
        self.connection.finalize(self.arys)

    def batch_lengths(self) -> torch.Tensor:
        """Valid elements per sample of the batch handed out last (variable_length windows)."""
        if self.transport != "device" or not self.engine.specs[self.target_rank - 1].var:
            raise RuntimeError("batch_lengths is defined for variable_length windows of the device transport")
        k = self.target_rank - 1
        b = self._cur_batch
        return self.engine.batch_lengths(k, self.connection.pushers[k].lent_buf, b * self.batch_size, self.batch_size)

    def remote_fraction(self) -> float:
        """Share of the current window's samples that live on another GPU (what crosses NVLink)."""
        if self.transport != "device":
            return 0.0
        eng, k = self.engine, self.target_rank - 1
        tbl = eng.table(k, self.connection.pushers[k].lent_buf)
        owner = (tbl >> 48) // eng.P
        where = owner if not eng.remote_storage else owner * 0 - 1
        return float((where != eng.me).float().mean().item())

    def batch_checksum(self) -> torch.Tensor:
        """Enqueue (on the current stream) the 64-bit fingerprint of the batch handed out last and
        return the one-element device tensor that will hold it: a cheap per-step result to read
        back, and the integrity check the reference lacks (csrc/checksum.cu)."""
        if self._cur_slot < 0 or not self.cuda:
            raise RuntimeError("batch_checksum needs a batch delivered on a CUDA device")
        from .ops.checksum import BatchChecksum

        k = self.target_rank - 1
        if self.transport == "device":
            s, raw = self.engine.specs[k], self.engine.ring[self._cur_slot]
        else:
            s = self.specs[k]
            raw = self._out[self._cur_slot] if self._plans[k] is not None else \
                self._stage[self._cur_slot][: self.batch_size * s.sample_bytes].view(self.batch_size, s.sample_bytes)
        payload = s.out_sample_bytes // 8 * 8
        if getattr(self, "_checksum", None) is None:
            self._checksum = BatchChecksum(self.device, self.batch_size, max(8, payload))
        return self._checksum.run(raw, payload)

    def reference_batch(self) -> Tuple[torch.Tensor, ...]:
        """The batch handed out last, rebuilt on an independent path: rows fetched one
        ``cudaMemcpyAsync`` at a time from their owners' shards (no kernel of this library), then
        converted by a plain PyTorch fp32 expression.  Must equal the delivered batch bit for bit;
        benchmarks and tests use it as the numerics oracle on real multi-GPU runs (the reference
        has no check of what a consumer received, reference tests/test_ddl.py:8-28)."""
        if self._cur_slot < 0 or self.transport != "device":
            raise RuntimeError("reference_batch needs a batch delivered by the device transport")
        from .ops.gather import convert_reference

        eng, k, B = self.engine, self.target_rank - 1, self.batch_size
        buf = self.connection.pushers[k].lent_buf
        first = self._cur_batch * B
        if eng.cuda:
            eng.side.synchronize()
        s = eng.specs[k]
        if s.replicate and eng.n_instances > 1 and eng.cuda:
            # a replicated window delivers instance 0's table: read that slice through its peer mapping
            mine = torch.empty(B, dtype=torch.int64, device=self.device)
            N.check(N.lib().bddl_memcpy_async(C.c_void_p(mine.data_ptr()),
                                              C.c_void_p(eng.heap.peer_ptr(0, "tables") + ((k * 2 + buf) * eng.n_max + first) * 8),
                                              B * 8, 4, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
            slots = mine.cpu().numpy()
        else:
            slots = eng.tables[k, buf, first:first + B].cpu().numpy()
        raw = eng.reference_rows(k, slots)
        lens = eng.batch_lengths(k, buf, first, B) if s.var else None
        y = convert_reference(raw, s.convert, lens, copy_elem_size=s.np_dtype.itemsize)
        if y.dtype != s.out_dtype:  # plain copies keep the window's own element type
            y = y.view(s.out_dtype)
        return torch.split(y[:, :s.nValues], list(s.splits), dim=1)

    def trace_summary(self) -> dict:
        """Device-time statistics per loader stage (needs ``LoaderConfig(trace=True)``)."""
        return self.engine.tracer.summary() if self.engine is not None else {}

    def __enter__(self) -> "DistributedDataLoader":
        return self

    def __exit__(self, exc_type, exc, tb) -> bool:
        self.close()  # also on an exception in the training loop, so producers and peers are released
        return False

    def close(self) -> None:
        """Early shutdown (the reference can only stop by running out of epochs)."""
        if not self._finalized:
            self.connection.shutdown_operation()
            self._finalize()

    # -- checkpoint (absent in the reference) ---------------------------------------------------
    def state_dict(self) -> dict:
        """Everything needed to continue the exact sample sequence.  Take it between batches
        (after ``mark(END_OF_BATCH)``), at the same step on every rank.

        device transport: position, slot tables, RNG streams, partner-RNG streams -- cheap because
        rows never move (a table is 8 bytes per sample).  process transport: the position only; the
        windows and generators are state of the producer processes, which are deterministic
        functions of their seeds, so ``load_state_dict`` replays the hand-overs instead."""
        if self.transport != "device":
            return {"transport": "process", "epoch": self.epoch, "batch": self.batch, "target_rank": self.target_rank,
                    "window_done": getattr(self, "_window_done", False)}
        return {"epoch": self.epoch, "batch": self.batch, "target_rank": self.target_rank, "len": self._len,
                "window_done": getattr(self, "_window_done", False),
                "producers": [p.state_dict() for p in self.connection.pushers]}

    def _replay_to(self, d: dict) -> None:
        """Process transport: bring a freshly constructed loader (and, through the ordinary lend /
        return tokens, its producer processes) to the checkpointed position.  Every skipped epoch
        borrows its window and hands it straight back, so each producer runs exactly the exchanges
        and shuffles it ran before the checkpoint; no batch is fetched."""
        if self.epoch != 0 or self.batch != 0:
            raise RuntimeError("process-transport checkpoints are loaded into a freshly constructed loader")
        self._pref = {}
        windows = d["epoch"] + (1 if d.get("window_done") else 0)
        for _ in range(windows):
            if self.cuda:
                for fs in self.fetch_streams:
                    fs.synchronize()  # prefetches of the window being skipped
            self._end_access_epoch()
            self._advance_to_next_producer()
            self._start_access_epoch()
        self.epoch, self.batch = d["epoch"], d["batch"]
        self._window_done = d.get("window_done", False)
        self._iter_start = self._len if self._window_done else self.batch
        for i in range(self.cfg.prefetch_depth):
            self._prefetch(self._iter_start + i)

    def load_state_dict(self, d: dict) -> None:
        """Resume on a freshly constructed loader (same producer function, batch size, world)."""
        if self.transport != "device":
            self._replay_to(d)
            return
        for p, pd in zip(self.connection.pushers, d["producers"]):
            p.load_state_dict(pd)
        self.epoch, self.batch, self.target_rank, self._len = d["epoch"], d["batch"], d["target_rank"], d["len"]
        self._pref = {}
        self._lend_seq += 1
        self._window_done = d.get("window_done", False)
        # an epoch whose window is finished but not yet marked yields nothing more
        self._iter_start = self._len if self._window_done else self.batch
        for i in range(self.cfg.prefetch_depth):
            self._prefetch(self._iter_start + i)
