"""Producer role (reference ddl/datapusher.py:14-170).

``DataPusher`` is the reference's producer *process*: it receives the user's producer function
from the consumer, asks it about its data (``on_init``), creates its shared window, lets the
user fill it (``post_init``) and then loops ``global_shuffle -> execute_function -> offer ->
(consumer reads) -> take back`` until shutdown.

``DevicePusher`` is the same state machine as a role inside the consumer process of the device
transport: "producing" means enqueueing ``index_exchange`` + ``perm_gen`` on the engine's side
stream, "offering" means recording a CUDA event, and nothing ever blocks the host.
"""
from __future__ import annotations

import logging
import pickle
from abc import ABC, abstractmethod
from typing import Any, List, Optional

import numpy as np

from .connection import Connection
from .ops.rng import Pcg64
from .shuffle import DeviceGlobalShuffler, make_global_shuffler
from .types import (DataProducerOnInitReturn, MetaData_Consumer_To_Producer, MetaData_Producer_To_Consumer,
                    WorkerInfo)
from .utils import execute_callbacks, for_all_methods, with_logging

log = logging.getLogger("ddl_b200")

__all__ = ["DataPusher", "DevicePusher", "DataProducerOnInitReturn"]


def _meta_from_init(ret: DataProducerOnInitReturn, batch_size: int) -> MetaData_Producer_To_Consumer:
    bpw = ret.nData // batch_size
    assert bpw > 0, "a window must hold at least one batch"
    return MetaData_Producer_To_Consumer(
        nData=ret.nData, nValues=ret.nValues, shape=tuple(ret.shape), splits=tuple(ret.splits),
        batches_per_window=bpw, dtype=getattr(ret, "dtype", np.float32), convert=getattr(ret, "convert", None),
        shuffle_seed=getattr(ret, "shuffle_seed", None), variable_length=getattr(ret, "variable_length", False),
        replicate=getattr(ret, "replicate", False))


def _private_copy(obj):
    """What pickling through MPI does to the producer function in the reference
    (reference ddl/types.py:9): each producer works on its own copy."""
    try:
        return pickle.loads(pickle.dumps(obj))
    except Exception:
        try:
            import cloudpickle

            return pickle.loads(cloudpickle.dumps(obj))
        except Exception:
            import copy

            return copy.deepcopy(obj)


class DataPusherABC(ABC):
    @abstractmethod
    def push_data(self) -> None: ...

    @abstractmethod
    def sync(self) -> None: ...

    @abstractmethod
    def _finalize(self) -> None: ...


@for_all_methods(with_logging)
class DataPusher(DataPusherABC):
    """Producer process of the ``process`` transport (reference ddl/datapusher.py:44-170)."""

    def __init__(self, connection: Connection, comm_per_gpu_shm, comm_nth_pusher, comm_global) -> None:
        self.callbacks: List[Any] = []
        self.connection = connection
        self.compat = connection.config.compat == "as_wired"
        self.comm_per_gpu_shm = comm_per_gpu_shm
        self.rank_per_gpu_shm: int = comm_per_gpu_shm.Get_rank()
        self.rank_global: int = comm_global.Get_rank()
        self.n_instance: int = comm_nth_pusher.Get_size()
        self.iterations = 0

        self.metadata_from_consumer: MetaData_Consumer_To_Producer = connection.recv_metadata_as_producer()
        self.callbacks.append(self.metadata_from_consumer.producer_function)
        on_init_return: DataProducerOnInitReturn = execute_callbacks(
            "on_init", callbacks=self.callbacks, compat=self.compat, rank_global=self.rank_global)
        self.metadata_to_consumer = _meta_from_init(on_init_return, self.metadata_from_consumer.batch_size)
        connection.send_metadata(self.metadata_to_consumer, "producer")

        # A declared shuffle (shuffle_seed) runs natively, out of place, between two windows: the
        # producer prepares iteration t+1 while the consumer still reads iteration t.
        cfg = connection.config
        self.declared = (on_init_return.shuffle_seed is not None and cfg.shuffle_mode in ("auto", "device")
                         and not getattr(on_init_return, "variable_length", False))
        self.arys = connection.init_windows(shapes=self.metadata_to_consumer.shape,
                                            dtypes=self.metadata_to_consumer.dtype, double=self.declared)
        self.my_ary = self.arys[self.rank_per_gpu_shm - 1]  # rank 0 is the consumer and owns no window
        self.windows = [self.my_ary, connection.arys_alt[self.rank_per_gpu_shm - 1]]
        self.cur = 0
        self.shuffle_rng = Pcg64(on_init_return.shuffle_seed) if self.declared else None
        connection.lock_windows()

        num_exchange = int(on_init_return.nData * self.metadata_from_consumer.global_shuffle_fraction_exchange)
        self.global_shuffler = None
        if self.n_instance > 1 and num_exchange > 0:
            self.global_shuffler = make_global_shuffler(
                self.metadata_from_consumer.global_shuffle_exchange_method, self.my_ary, num_exchange,
                comm_nth_pusher, comm_global)
            self.callbacks.append(self.global_shuffler)

        connection.sync(self.rank_per_gpu_shm)
        execute_callbacks("post_init", callbacks=self.callbacks, compat=self.compat, rank_global=self.rank_global,
                          my_ary=self.my_ary, n_instance=self.n_instance)
        connection.sync(self.rank_per_gpu_shm)
        connection.Barrier()

    def sync(self) -> None:
        self.connection.sync(self.rank_per_gpu_shm)

    def _istart_access_epoch(self, target_rank: int = 0) -> WorkerInfo:
        return self.connection.Istart_access_epoch(target_rank)

    def _iend_access_epoch(self, target_rank: int = 0) -> WorkerInfo:
        return self.connection.Iend_access_epoch(target_rank)

    def _finalize(self) -> None:
        self.connection.Barrier()
        self.connection.unlock_windows()
        self.connection.finalize(self.arys)

    def _prepare_declared(self) -> None:
        """One producer iteration in declared mode: (exchange +) shuffle from the current window
        into the other one, natively and threaded; same rows in the same order as
        global_shuffle() followed by rng.shuffle(my_ary) (reference ddl/datapusher.py:152-154)."""
        import ctypes as C

        from . import _native as N

        src, dst = self.windows[self.cur], self.windows[self.cur ^ 1]
        n = src.shape[0]
        recv, ne = None, 0
        gs = self.global_shuffler
        if gs is not None and not self.compat:
            send_to, recv_from = gs._calculate_comm_partner()
            ne, half = gs.num_exchange, gs.num_exchange // 2
            recv = np.ascontiguousarray(src[:ne]).copy()
            gs.comm_nth_pusher.sendrecv_replace(recv[:half], dest=send_to, source=recv_from, sendtag=50, recvtag=50)
            gs.comm_nth_pusher.sendrecv_replace(recv[half:ne], dest=recv_from, source=send_to, sendtag=51, recvtag=51)
        perm = self.shuffle_rng.permutation(n)
        N.check(N.lib().bddl_host_permute_rows(
            dst.ctypes.data, src.ctypes.data, recv.ctypes.data if recv is not None else None, ne,
            perm.ctypes.data, n, src.strides[0], self.connection.config.host_threads))
        self.cur ^= 1
        self.iterations += 1
        self.connection.heartbeat(self.rank_per_gpu_shm)

    def _push_declared(self) -> None:
        cbs, compat = self.callbacks, self.compat
        self._prepare_declared()
        while True:
            if self._iend_access_epoch() is WorkerInfo.STOP:
                break
            # the consumer reads windows[cur]; it returned the other window before borrowing this
            # one, so the next iteration can be prepared right now
            self._prepare_declared()
            if self._istart_access_epoch() is WorkerInfo.STOP:
                break
            execute_callbacks("on_shuffle_end", callbacks=cbs, compat=compat)

    def push_data(self) -> None:
        cbs, compat = self.callbacks, self.compat
        execute_callbacks("on_push_begin", callbacks=cbs, compat=compat)
        if self.declared:
            self._push_declared()
            execute_callbacks("on_push_end", callbacks=cbs, compat=compat)
            self._finalize()
            return
        while True:
            execute_callbacks("global_shuffle", callbacks=cbs, compat=compat)
            self.sync()
            execute_callbacks("execute_function", callbacks=cbs, compat=compat)
            self.iterations += 1
            self.connection.heartbeat(self.rank_per_gpu_shm)
            if self._iend_access_epoch() is WorkerInfo.STOP:
                break
            # the consumer reads the window here
            if self._istart_access_epoch() is WorkerInfo.STOP:
                break
            execute_callbacks("on_shuffle_end", callbacks=cbs, compat=compat)
        execute_callbacks("on_push_end", callbacks=cbs, compat=compat)
        self._finalize()


class DevicePusher:
    """Producer role of the ``device`` transport: the reference's producer loop, unrolled into
    stream-ordered steps (reference ddl/datapusher.py:46-170).

    ``produce`` = global_shuffle -> sync -> execute_function -> offer (event on the side stream);
    ``lend`` = the consumer's start_access_epoch; ``take_back`` = its end_access_epoch followed by
    on_shuffle_end.  With a declared shuffle seed (or in trace mode) the next iteration is
    produced while the consumer is still reading the current one (tables are double-buffered),
    so even a single producer never stalls the consumer.
    """

    def __init__(self, connection: Connection, env, producer_idx: int, remote_meta=None):
        self.connection, self.env = connection, env
        self.remote = remote_meta is not None  # the data lives on a dedicated storage GPU
        self.cfg = connection.config
        self.compat = self.cfg.compat == "as_wired"
        self.producer_idx = int(producer_idx)  # 1-based, like the reference's shm rank
        self.k = self.producer_idx - 1
        self.rank_global = env.virtual_rank_global(self.producer_idx)
        self.n_instance = env.n_instances
        self.metadata_from_consumer: MetaData_Consumer_To_Producer = connection.recv_metadata_as_producer()
        # every producer gets its own copy, as it would after pickling through MPI
        pf = _private_copy(self.metadata_from_consumer.producer_function)
        self.producer_function = pf
        self.callbacks: List[Any] = [pf]
        if self.remote:
            # on_init / post_init ran on the storage rank that holds this producer's shard
            self.on_init_return = remote_meta
            self.metadata_to_consumer = remote_meta
        else:
            self.on_init_return = execute_callbacks(
                "on_init", callbacks=self.callbacks, compat=self.compat, rank_global=self.rank_global)
            self.metadata_to_consumer = _meta_from_init(self.on_init_return, self.metadata_from_consumer.batch_size)
        connection.send_metadata(self.metadata_to_consumer, "producer")
        self.engine = None
        self.my_ary: Optional[np.ndarray] = None
        self.global_shuffler = None
        self.mode = "device"
        self.iterations = 0
        self._ready = None        # (table buffer, event) of the newest produced iteration
        self._produce_seq = 0     # tables produced so far (never rewound: the push form's `table` word)
        self._ready_seq = 0
        self.lent_seq = 0
        self._returned_ev = None  # consumer finished reading the previous iteration
        self.lent_buf = 0
        self._stopped = False

    # -- set-up -----------------------------------------------------------------------------------------
    def bind(self, engine) -> None:
        """Second half of the reference constructor: window, shuffler, post_init, barrier."""
        self.engine = engine
        k, meta = self.k, self.metadata_from_consumer
        mode = self.cfg.shuffle_mode
        seed = self.metadata_to_consumer.shuffle_seed
        if mode == "auto":
            mode = "device" if seed is not None else "host"
        if mode == "device" and seed is None:
            raise ValueError("shuffle_mode='device' needs DataProducerOnInitReturn.shuffle_seed")
        self.mode = mode
        self.run_ahead = mode in ("device", "trace")
        if self.remote and mode != "device":
            raise ValueError("dedicated storage GPUs need a declared shuffle (shuffle_seed): the window is not local")

        self.my_ary = None if self.remote else engine.alloc_window(k)
        num_exchange = int(self.on_init_return.nData * meta.global_shuffle_fraction_exchange)
        if self.n_instance > 1 and num_exchange > 0:
            if meta.global_shuffle_exchange_method != "sendrecv_replace":
                raise NotImplementedError(f"{meta.global_shuffle_exchange_method} is not implemented")
            if mode == "host":
                from .shuffle import SendRecvReplaceGlobalShuffler

                self.global_shuffler = SendRecvReplaceGlobalShuffler(
                    self.my_ary, num_exchange, _TaggedComm(self.env.comm_nth_pusher, self.k), None,
                    seed=self.producer_idx, n_instance=self.n_instance, my_index=self.env.color)
            else:
                self.global_shuffler = DeviceGlobalShuffler(engine, k, num_exchange, self.n_instance,
                                                            self.env.color, seed=self.producer_idx)
            self.callbacks.append(self.global_shuffler)
        if not self.remote:
            execute_callbacks("post_init", callbacks=self.callbacks, compat=self.compat,
                              rank_global=self.rank_global, my_ary=self.my_ary, n_instance=self.n_instance)
            engine.load_shard(k, self.my_ary)
        if engine.specs[k].var:
            if self.remote:  # post_init ran on the storage rank, which reports the row lengths
                lens = self.env.comm_global.recv_obj(self.env.storage_rank_of(self.env.color), tag=40 + k)
            else:
                lens = getattr(self.producer_function, "lengths", None)
            if lens is None:
                raise ValueError("variable_length windows need producer_function.lengths after post_init")
            engine.set_lengths(k, np.asarray(lens))
        engine.init_table(k, Pcg64(seed) if mode == "device" else None)
        if mode == "trace":
            self._proxy = np.arange(self.on_init_return.nData, dtype=np.int64).reshape(-1, 1)

    def start(self) -> None:
        execute_callbacks("on_push_begin", callbacks=self.callbacks, compat=self.compat)
        self.produce()

    # -- one producer iteration ---------------------------------------------------------------------------
    def produce(self) -> None:
        eng, k = self.engine, self.k
        for ev in self._returned_ev or ():
            eng.wait(eng.side, ev)  # the buffer about to be rewritten is no longer read
        if eng.tracer.enabled:
            with eng.tracer.span(f"produce.window{k + 1}", eng.side):
                self._produce_body(eng, k)
        else:
            self._produce_body(eng, k)
        self._produce_seq += 1
        if eng.push:
            # owners on other GPUs read this table from inside their push kernels: publish its number
            eng.signal_table_ready(k, self._produce_seq)
        ev = eng.event()
        ev.record(eng.side) if eng.cuda else None
        self._ready = (eng.cur[k], ev)
        self._ready_seq = self._produce_seq
        self.iterations += 1

    def _produce_body(self, eng, k) -> None:
        if self.mode == "host":
            if eng.cuda:
                eng.side.synchronize()  # the previous upload has left the pinned window
            execute_callbacks("global_shuffle", callbacks=self.callbacks, compat=self.compat)
            execute_callbacks("execute_function", callbacks=self.callbacks, compat=self.compat)
            eng.load_shard(k, self.my_ary)
        else:
            before = eng.cur[k]
            execute_callbacks("global_shuffle", callbacks=self.callbacks, compat=self.compat)
            if eng.cur[k] == before:
                if eng.push and eng.specs[k].replicate:
                    # instance 0's table is read by every GPU for every consumer: nobody may rewrite a
                    # table buffer before ALL ranks have finished the window that used it (the
                    # exchange, when active, brings this barrier along)
                    eng.stream_barrier()
                eng.copy_forward(k)  # no exchange this iteration: still move to the other buffer
            if self.mode == "device":
                eng.local_shuffle(k)
            else:
                self._trace_shuffle()

    def _trace_shuffle(self) -> None:
        """Run the user's execute_function on a row-id proxy and apply the permutation it made."""
        import torch

        eng, k, pf = self.engine, self.k, self.producer_function
        n = self.on_init_return.nData
        self._proxy[:, 0] = np.arange(n, dtype=np.int64)
        real, pf.my_ary = getattr(pf, "my_ary", None), self._proxy
        try:
            execute_callbacks("execute_function", callbacks=[pf], compat=self.compat)
        finally:
            pf.my_ary = real
        perm = torch.from_numpy(self._proxy[:, 0].copy())
        cur = eng.table(k)
        with torch.cuda.stream(eng.side) if eng.cuda else _null():
            perm_d = perm.to(eng.device, non_blocking=True)
            cur.copy_(cur.index_select(0, perm_d))

    # -- hand-over -------------------------------------------------------------------------------------------
    def lend(self) -> None:
        buf, ev = self._ready
        eng = self.engine
        for st in eng.fetch_streams:
            eng.wait(st, ev)
        self.lent_buf = buf
        self.lent_seq = self._ready_seq
        eng.on_lend(self.k, buf)
        if self.run_ahead and not self._stopped:
            self.produce()

    def take_back(self) -> None:
        eng = self.engine
        evs = []
        for st in eng.fetch_streams:
            ev = eng.event()
            if eng.cuda:
                ev.record(st)
            evs.append(ev)
        self._returned_ev = evs
        execute_callbacks("on_shuffle_end", callbacks=self.callbacks, compat=self.compat)
        if not self.run_ahead and not self._stopped:
            self.produce()

    def stop(self) -> None:
        if not self._stopped:
            self._stopped = True
            execute_callbacks("on_push_end", callbacks=self.callbacks, compat=self.compat)

    # -- checkpoint --------------------------------------------------------------------------------------------
    def state_dict(self) -> dict:
        eng, k = self.engine, self.k
        if eng.cuda:
            eng.side.synchronize()
        d = {"iterations": self.iterations, "mode": self.mode,
             "table": eng.table(k, self._ready[0]).cpu().numpy().copy(),
             "lent_table": eng.table(k, self.lent_buf).cpu().numpy().copy(),
             "lent_is_ready": self.lent_buf == self._ready[0],
             "rng_words": eng.states[k].cpu().numpy().copy()}
        if self.global_shuffler is not None:
            d["shuffler"] = self.global_shuffler.state_dict()
        return d

    def load_state_dict(self, d: dict) -> None:
        """Restore tables, RNG stream and iteration count (device / trace shuffle modes)."""
        import torch

        if self.mode == "host":
            raise NotImplementedError("host-shuffle windows are user state; checkpoint them in the producer function")
        eng, k = self.engine, self.k
        if eng.cuda:
            eng.side.synchronize()
            for st in eng.fetch_streams:
                st.synchronize()
        ready_buf = eng.cur[k]
        eng.table(k, ready_buf).copy_(torch.from_numpy(d["table"]))
        if d["lent_is_ready"]:
            self.lent_buf = ready_buf
        else:
            self.lent_buf = 1 - ready_buf
            eng.table(k, self.lent_buf).copy_(torch.from_numpy(d["lent_table"]))
        eng.states[k].copy_(torch.from_numpy(d["rng_words"]))
        self.iterations = d["iterations"]
        if self.global_shuffler is not None and "shuffler" in d:
            self.global_shuffler.load_state_dict(d["shuffler"])
        ev = eng.event()
        if eng.cuda:
            torch.cuda.current_stream(eng.device).synchronize()
            ev.record(eng.side)
        self._ready = (ready_buf, ev)
        self._returned_ev = None


class _TaggedComm:
    """comm_nth_pusher as seen by the k-th in-process producer: tags are offset per producer so
    that several producers of one process can exchange concurrently without cross-talk."""

    def __init__(self, comm, k: int):
        self._c, self._off = comm, 1000 * (k + 1)

    def Get_size(self):
        return self._c.Get_size()

    def Get_rank(self):
        return self._c.Get_rank()

    def sendrecv_replace(self, array, dest, source, sendtag=0, recvtag=0):
        return self._c.sendrecv_replace(array, dest, source, sendtag + self._off, recvtag + self._off)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
