"""Connection / protocol layer between the consumer and its producers
(reference ddl/connection.py:17-187).

Same public surface as the reference -- metadata handshake, window creation, the lend/return
token protocol in blocking and cancellable form, shutdown, teardown -- over two transports:

``process``  Producers are OS processes (the reference layout).  Windows are POSIX shared
    memory instead of ``MPI.Win.Allocate_shared``; the zero-byte tag-7 tokens become three
    monotonically increasing counters per producer in a shared flag page (``offered``,
    ``borrowed``, ``returned``); the Ibarrier shutdown becomes a ``stop`` word that every wait
    polls.  Waits are bounded: a dead peer raises ``LoaderTimeoutError`` instead of hanging
    (the reference's only watchdog is the pytest timeout).

``device``   Producers are roles inside the consumer process (one process per GPU).  The same
    calls map onto CUDA events and streams of the ``DeviceEngine``; nothing blocks the host.

Per-window ownership state machine, identical in both::

    FILLING --offer--> OFFERED --borrow--> LENT --return--> FILLING
"""
from __future__ import annotations

import ctypes as C
import logging
import math
import uuid
from typing import Any, List, Optional

import numpy as np

from . import _native as N
from .exceptions import LoaderTimeoutError
from .types import MetaData_Consumer_To_Producer, MetaData_Producer_To_Consumer, WorkerInfo
from .utils import for_all_methods, with_logging

log = logging.getLogger("ddl_b200")

_FLAG_BYTES = 4096
_STOP, _HEART0 = 0, 512  # word indices in the flag page
_PER_PRODUCER = 4        # offered, borrowed, returned, spare


def _word(base: int, idx: int) -> C.c_void_p:
    return C.c_void_p(base + 4 * idx)


@for_all_methods(with_logging)
class Connection:
    def __init__(self, comm_per_gpu_shm, env=None, config=None):
        from .config import LoaderConfig

        self.comm_per_gpu_shm = comm_per_gpu_shm
        self.env = env
        self.config = config or LoaderConfig.from_env()
        self.transport = env.transport if env is not None else ("process" if comm_per_gpu_shm.Get_size() > 1 else "device")
        self.n_producers = env.n_producers if env is not None else max(1, comm_per_gpu_shm.Get_size() - 1)
        self.rank = comm_per_gpu_shm.Get_rank()
        self.wins: List[Any] = []
        self._stopped = False
        # device transport: attached by the loader once the windows are known
        self.engine = None
        self.pushers: List[Any] = []
        self._mailbox: dict = {}
        self._flag_base = 0
        self._flag_name = ""
        self._counts = [0] * (self.n_producers + 1)  # per producer: completed lend/return cycles
        if self.transport == "process":
            assert comm_per_gpu_shm.Get_size() > 1
            self._init_flag_page()

    # -- process transport: flag page --------------------------------------------------------------
    def _init_flag_page(self) -> None:
        lib = N.lib()
        p = C.c_void_p()
        if self.rank == 0:
            name = f"/bddl_flags_{uuid.uuid4().hex[:12]}"
            N.check(lib.bddl_shm_create(name.encode(), _FLAG_BYTES, C.byref(p)))
            N.shm_created(name)
            C.memset(p.value, 0, _FLAG_BYTES)
            self.comm_per_gpu_shm.bcast_obj(name, 0)
        else:
            name = self.comm_per_gpu_shm.bcast_obj(None, 0)
            N.check(lib.bddl_shm_attach(name.encode(), _FLAG_BYTES, C.byref(p)))
        self._flag_base, self._flag_name = p.value, name
        self.comm_per_gpu_shm.barrier()
        if self.rank == 0:
            lib.bddl_shm_unlink(name.encode())
            N.shm_unlinked(name)

    def _flag(self, producer: int, which: int) -> C.c_void_p:
        return _word(self._flag_base, 16 + producer * _PER_PRODUCER + which)

    def _wait(self, flag: C.c_void_p, value: int, cancellable: bool) -> WorkerInfo:
        stop = _word(self._flag_base, _STOP) if cancellable else None
        rc = N.lib().bddl_hflag_wait_geq(flag, value, int(self.config.timeout_s * 1e6), stop)
        if rc == 0:
            return WorkerInfo.CONTINUE
        if rc == -5:  # BDDL_ERR_STOPPED
            return WorkerInfo.STOP
        msg = N.lib().bddl_last_error().decode("utf-8", "replace")
        raise LoaderTimeoutError(f"peer did not answer within {self.config.timeout_s}s: {msg}")

    def heartbeat(self, producer: int) -> None:
        """Producers bump a counter every iteration; consumers can read it (failure detection)."""
        if self._flag_base:
            N.lib().bddl_hflag_add(_word(self._flag_base, _HEART0 + producer), 1)

    def heartbeats(self) -> List[int]:
        if not self._flag_base:
            return []
        return [int(N.lib().bddl_hflag_load(_word(self._flag_base, _HEART0 + k))) for k in range(1, self.n_producers + 1)]

    # -- metadata handshake (reference ddl/connection.py:65-86) ------------------------------------
    def send_metadata(self, metadata, called_from: str) -> None:
        if self.transport == "device":
            self._mailbox[called_from] = self._mailbox.get(called_from, []) + [metadata]
            return
        if called_from == "consumer":
            for t in range(1, self.comm_per_gpu_shm.Get_size()):
                self.comm_per_gpu_shm.ssend(metadata, dest=t, tag=0)
        elif called_from == "producer":
            self.comm_per_gpu_shm.ssend(metadata, dest=0, tag=0)
        else:
            raise ValueError(f"{called_from=} is not valid")

    def recv_metadata_as_producer(self) -> MetaData_Consumer_To_Producer:
        if self.transport == "device":
            return self._mailbox["consumer"][-1]
        return self.comm_per_gpu_shm.recv(source=0, tag=0)

    def recv_metadata_as_consumer(self) -> List[MetaData_Producer_To_Consumer]:
        if self.transport == "device":
            out, self._mailbox["producer"] = self._mailbox.get("producer", []), []
            return out
        return [self.comm_per_gpu_shm.recv(source=t, tag=0) for t in range(1, self.comm_per_gpu_shm.Get_size())]

    # -- windows (reference ddl/connection.py:88-139) ---------------------------------------------------
    def init_windows(self, shapes, dtypes=None, double=False) -> List[Optional[np.ndarray]]:
        """Process transport: one shared-memory array per producer, mapped by every rank of the
        instance.  ``shapes`` is one shape (producer side) or one per producer (consumer side).

        ``double`` (producer side: this producer's choice) allocates a second window per producer --
        the scheme the reference sketches but never built (reference ddl/mpi_dataloader.py:19-28):
        the producer fills one window while the consumer reads the other.  Which one is current
        follows from the hand-over count (``current_window``)."""
        if self.transport == "device":
            raise RuntimeError("device transport keeps windows as slot tables; see DeviceEngine")
        comm = self.comm_per_gpu_shm
        size = comm.Get_size()
        if not isinstance(shapes, list):
            shapes = [shapes] * (size - 1)
        if dtypes is None:
            dtypes = [np.float32] * (size - 1)
        elif not isinstance(dtypes, list):
            dtypes = [dtypes] * (size - 1)
        # every rank learns every producer's geometry (Shared_query in the reference)
        mine = None if self.rank == 0 else (tuple(int(x) for x in shapes[self.rank - 1]),
                                            np.dtype(dtypes[self.rank - 1]).str, bool(double))
        info = comm.allgather_obj(mine)
        shapes = [info[t][0] for t in range(1, size)]
        dtypes = [np.dtype(info[t][1]) for t in range(1, size)]
        doubles = [info[t][2] for t in range(1, size)]
        lib = N.lib()
        wins, arys = [], []
        self.wins_alt, self.arys_alt = [None] * (size - 1), [None] * (size - 1)
        session = comm.bcast_obj(uuid.uuid4().hex[:10] if self.rank == 0 else None, 0)
        jobs = [(t, 0) for t in range(1, size)] + [(t, 1) for t in range(1, size) if doubles[t - 1]]
        for target, which in jobs:
            shape = tuple(int(x) for x in shapes[target - 1])
            dt = np.dtype(dtypes[target - 1])
            nbytes = max(16, math.prod(shape) * dt.itemsize)
            name = f"/bddl_win_{session}_{target}_{which}"
            p = C.c_void_p()
            if target == self.rank:
                N.check(lib.bddl_shm_create(name.encode(), nbytes, C.byref(p)))
                N.shm_created(name)
            comm.barrier()
            if target != self.rank:
                N.check(lib.bddl_shm_attach(name.encode(), nbytes, C.byref(p)))
            comm.barrier()
            if target == self.rank:
                lib.bddl_shm_unlink(name.encode())
                N.shm_unlinked(name)
            raw = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p.value))
            ary = raw[: math.prod(shape) * dt.itemsize].view(dt).reshape(shape)
            if which == 0:
                wins.append((p.value, nbytes))
                arys.append(ary)
            else:
                self.wins_alt[target - 1] = (p.value, nbytes)
                self.arys_alt[target - 1] = ary
        self.wins = wins
        return arys

    def current_window(self, target_rank: int) -> int:
        """0 = primary, 1 = second window of producer ``target_rank`` for the hand-over in progress:
        iteration t (0-based) of a double-window producer lands in window (t + 1) & 1."""
        if not getattr(self, "arys_alt", None) or self.arys_alt[target_rank - 1] is None:
            return 0
        return (self._counts[target_rank] + 1) & 1

    def lock_windows(self) -> None:
        """MPI passive-target epochs have no counterpart: ordering comes from the flags."""

    def unlock_windows(self) -> None:
        pass

    def sync(self, target: int = 0) -> None:
        """Win.Sync: a full memory fence (reference ddl/connection.py:61-63)."""
        if self._flag_base:
            N.lib().bddl_hflag_add(_word(self._flag_base, 8), 0)

    _sync = sync

    def finalize(self, arys: List[Optional[np.ndarray]]) -> None:
        lib = N.lib()
        for i in range(len(arys)):
            arys[i] = None
        for p, n in list(self.wins) + [w for w in getattr(self, "wins_alt", []) if w is not None]:
            lib.bddl_shm_detach(C.c_void_p(p), n)
        self.wins, self.wins_alt, self.arys_alt = [], [], []
        if self._flag_base:
            lib.bddl_shm_detach(C.c_void_p(self._flag_base), _FLAG_BYTES)
            self._flag_base = 0

    def Barrier(self) -> None:
        self.comm_per_gpu_shm.barrier()

    # -- lend / return tokens (reference ddl/connection.py:153-182) -------------------------------------
    # consumer side, blocking
    def start_access_epoch(self, target_rank: int) -> None:
        """Borrow producer ``target_rank``'s window: wait until it is offered."""
        if self.transport == "device":
            self.pushers[target_rank - 1].lend()
            return
        n = self._counts[target_rank] + 1
        self._wait(self._flag(target_rank, 0), n, cancellable=False)
        N.lib().bddl_hflag_store(self._flag(target_rank, 1), n)
        self.sync(target_rank)

    def end_access_epoch(self, target_rank: int) -> None:
        """Return the window; its producer may reshuffle it."""
        if self.transport == "device":
            self.pushers[target_rank - 1].take_back()
            return
        self.sync(target_rank)
        self._counts[target_rank] += 1
        N.lib().bddl_hflag_store(self._flag(target_rank, 2), self._counts[target_rank])

    # producer side, cancellable by shutdown
    def Iend_access_epoch(self, target_rank: int = 0) -> WorkerInfo:
        """Offer my window and wait until the consumer has borrowed it (or shutdown)."""
        me = self.rank
        self.sync(me)
        n = self._counts[me] + 1
        N.lib().bddl_hflag_store(self._flag(me, 0), n)
        return self._wait(self._flag(me, 1), n, cancellable=True)

    def Istart_access_epoch(self, target_rank: int = 0) -> WorkerInfo:
        """Wait until the consumer has returned my window (or shutdown)."""
        me = self.rank
        n = self._counts[me] + 1
        info = self._wait(self._flag(me, 2), n, cancellable=True)
        if info is WorkerInfo.CONTINUE:
            self._counts[me] = n
        self.sync(me)
        return info

    # -- shutdown (reference ddl/connection.py:184-187) ---------------------------------------------------
    def shutdown_operation(self) -> None:
        """Consumer only: cancel every producer wait."""
        self._stopped = True
        if self.transport == "device":
            for p in self.pushers:
                p.stop()
            return
        assert self.rank == 0
        N.lib().bddl_hflag_store(_word(self._flag_base, _STOP), 1)

    @property
    def stopped(self) -> bool:
        if self._flag_base:
            return bool(N.lib().bddl_hflag_load(_word(self._flag_base, _STOP)))
        return self._stopped

    # -- device transport wiring -----------------------------------------------------------------------------
    def attach(self, engine, pushers) -> None:
        self.engine, self.pushers = engine, list(pushers)

    def close(self) -> None:
        if self.engine is not None:
            self.engine.close()
            self.engine = None
        self.pushers = []
