"""BASELINE-ONLY stand-in for the subset of mpi4py the reference uses.

There is no MPI library, mpirun or mpi4py in this image or its wheelhouse, so the reference's
unmodified Python cannot run on a real MPI here.  This module provides exactly the calls it makes
-- pickled ssend/recv, zero-byte Ssend/Recv/Issend/Irecv, Request.Waitany/Cancel/Wait,
Barrier/Ibarrier, Split/Split_type/Dup, shared windows (Allocate_shared/Shared_query/tomemory/
Fence/Lock_all/Unlock_all/Sync/Free), Sendrecv_replace, Info, FLOAT/INT, Abort -- over a c10d
TCPStore (messages) and multiprocessing.shared_memory (windows).  Ranks are started by
baseline/mpirun.py.  It is never imported by the ddl_b200 package; every number produced with it
is labelled "reference code over in-box MPI stand-in".
"""
from __future__ import annotations

import atexit
import os
import pickle
import time
from datetime import timedelta
from multiprocessing import resource_tracker, shared_memory

import numpy as np
from torch.distributed import TCPStore

_RANK = int(os.environ.get("BDDL_MPI_RANK", "0"))
_SIZE = int(os.environ.get("BDDL_MPI_SIZE", "1"))
_HOST = os.environ.get("BDDL_MPI_ADDR", "127.0.0.1")
_PORT = int(os.environ.get("BDDL_MPI_PORT", "29610"))
_TIMEOUT = float(os.environ.get("BDDL_MPI_TIMEOUT", "1800"))

_store = TCPStore(_HOST, _PORT, world_size=_SIZE, is_master=(_RANK == 0), timeout=timedelta(seconds=_TIMEOUT),
                  wait_for_workers=True) if _SIZE > 1 else None
_local: dict = {}

COMM_TYPE_SHARED = 1
INFO_NULL = None
BSEND_OVERHEAD = 0


class _Datatype:
    def __init__(self, name, np_dtype):
        self.name, self.np_dtype = name, np.dtype(np_dtype)

    def Get_size(self):
        return self.np_dtype.itemsize


FLOAT = _Datatype("MPI_FLOAT", np.float32)
INT = _Datatype("MPI_INT", np.int32)
DOUBLE = _Datatype("MPI_DOUBLE", np.float64)


def _set(key, val):
    if _store is None:
        _local[key] = val
    else:
        _store.set(key, val)


def _get(key):
    if _store is None:
        return _local[key]
    return _store.get(key)


def _has(key):
    if _store is None:
        return key in _local
    return _store.check([key])


def _del(key):
    try:
        if _store is None:
            _local.pop(key, None)
        else:
            _store.delete_key(key)
    except Exception:
        pass


def _add(key, n):
    if _store is None:
        _local[key] = _local.get(key, 0) + n
        return _local[key]
    return _store.add(key, n)


class Info:
    def __init__(self):
        self.d = {}

    @staticmethod
    def Create():
        return Info()

    def Set(self, k, v):
        self.d[k] = v


class Request:
    """kind: None (null request), 'send', 'recv', 'barrier'."""

    def __init__(self, kind=None, key=None, done_key=None):
        self.kind, self.key, self.done_key = kind, key, done_key
        self.complete = kind is None
        self.cancelled = False

    def _test(self):
        if self.complete or self.cancelled:
            return self.complete
        if self.kind == "send":
            if _has(self.done_key):
                _del(self.done_key)
                self.complete = True
        elif self.kind == "recv":
            if _has(self.key):
                _del(self.key)
                _set(self.done_key, b"1")
                self.complete = True
        elif self.kind == "barrier":
            if _has(self.done_key):
                self.complete = True
        return self.complete

    def Wait(self):
        delay = 20e-6
        while not self._test():
            time.sleep(delay)
            delay = min(delay * 1.5, 1e-3)
        return True

    def Test(self):
        return self._test()

    def Cancel(self):
        if not self.complete:
            self.cancelled = True
            if self.kind == "send":
                _del(self.key)  # withdraw the posted message

    @staticmethod
    def Waitany(requests):
        delay = 20e-6
        while True:
            for i, r in enumerate(requests):
                if r.kind is not None and r._test():
                    return i
            time.sleep(delay)
            delay = min(delay * 1.5, 1e-3)

    @staticmethod
    def Waitall(requests):
        for r in requests:
            r.Wait()


_ALL_COMMS = []


class Intracomm:
    _ids = 0

    def __init__(self, members, cid):
        self.members = list(members)
        self.cid = cid
        self.rank = self.members.index(_RANK)
        self.size = len(self.members)
        self.name = cid
        self._seq = {}
        self._coll = 0
        self._segments = {}
        self._attached = {}
        _ALL_COMMS.append(self)

    # -- bookkeeping ------------------------------------------------------------------------------
    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size

    def Set_name(self, n):
        self.name = n

    def Get_name(self):
        return self.name

    def _next(self, *k):
        v = self._seq.get(k, 0)
        self._seq[k] = v + 1
        return v

    def _allgather(self, obj):
        c = self._coll
        self._coll += 1
        base = f"{self.cid}/ag{c}"
        _set(f"{base}/{self.rank}", pickle.dumps(obj))
        return [pickle.loads(_get(f"{base}/{r}")) for r in range(self.size)]

    # -- pickled point to point -------------------------------------------------------------------------
    def _post(self, payload, dest, tag):
        seq = self._next("s", dest, tag)
        key = f"{self.cid}/{self.rank}>{dest}/{tag}/{seq}"
        _set(key, payload)
        return key

    def _take(self, source, tag):
        seq = self._next("r", source, tag)
        key = f"{self.cid}/{source}>{self.rank}/{tag}/{seq}"
        data = _get(key)
        _del(key)
        _set(key + "/ack", b"1")
        return data

    def ssend(self, obj, dest, tag=0):
        key = self._post(pickle.dumps(obj), dest, tag)
        _get(key + "/ack")
        _del(key + "/ack")

    send = ssend

    def recv(self, source=0, tag=0):
        return pickle.loads(self._take(source, tag))

    # -- buffer point to point (the reference only sends zero bytes: [None, 0, MPI.INT]) ------------------
    @staticmethod
    def _bytes_of(buf):
        if isinstance(buf, (list, tuple)):
            if buf[0] is None:
                return b""
            return np.ascontiguousarray(buf[0]).tobytes()
        return np.ascontiguousarray(buf).tobytes()

    def Ssend(self, buf, dest, tag=0):
        key = self._post(self._bytes_of(buf), dest, tag)
        _get(key + "/ack")
        _del(key + "/ack")

    def Recv(self, buf, source=0, tag=0):
        data = self._take(source, tag)
        if not isinstance(buf, (list, tuple)) and len(data):
            np.asarray(buf).reshape(-1).view(np.uint8)[:] = np.frombuffer(data, dtype=np.uint8)

    def Issend(self, buf, dest, tag=0):
        key = self._post(self._bytes_of(buf), dest, tag)
        return Request("send", key, key + "/ack")

    def Irecv(self, buf, source=0, tag=0):
        seq = self._next("r", source, tag)
        key = f"{self.cid}/{source}>{self.rank}/{tag}/{seq}"
        return Request("recv", key, key + "/ack")

    def Sendrecv_replace(self, buf, dest, sendtag=0, source=0, recvtag=0):
        """Bulk payloads travel through POSIX shared-memory segments (all ranks share the node),
        one persistent segment per (destination, tag) like a shared-memory BTL's eager buffers;
        only the segment's name goes through the store."""
        arr = np.asarray(buf)
        flat = arr.reshape(-1).view(np.uint8)
        seq = self._next("bs", dest, sendtag)
        key_s = f"{self.cid}/b/{self.rank}>{dest}/{sendtag}/{seq}"
        if flat.size:
            seg = self._segments.get((dest, sendtag))
            if seg is None or seg.size < flat.size:
                if seg is not None:
                    seg.close()
                    seg.unlink()
                name = (f"bddlmpi_{os.environ.get('BDDL_MPI_SESSION', 'x')}_sr_{self.cid.replace('/', '_')}"
                        f"_{self.rank}_{dest}_{sendtag}_{seq}")
                seg = shared_memory.SharedMemory(name=name, create=True, size=flat.size)
                self._segments[(dest, sendtag)] = seg
            np.frombuffer(seg.buf, dtype=np.uint8, count=flat.size)[:] = flat
            _set(key_s, seg.name.encode())
        else:
            _set(key_s, b"")
        seq = self._next("br", source, recvtag)
        key_r = f"{self.cid}/b/{source}>{self.rank}/{recvtag}/{seq}"
        name = _get(key_r).decode()
        if name:
            src = self._attached.get(name)
            if src is None:
                src = shared_memory.SharedMemory(name=name)
                try:
                    resource_tracker.unregister(src._name, "shared_memory")
                except Exception:
                    pass
                self._attached[name] = src
            flat[:] = np.frombuffer(src.buf, dtype=np.uint8, count=flat.size)
        _del(key_r)
        _set(key_r + "/ack", b"1")
        _get(key_s + "/ack")  # the receiver has copied my payload: the segment may be reused
        _del(key_s + "/ack")

    # -- collectives --------------------------------------------------------------------------------------
    def Ibarrier(self):
        c = self._next("ibar")
        key = f"{self.cid}/ibar{c}"
        if _add(key, 1) == self.size:
            _set(key + "/go", b"1")
        return Request("barrier", key, key + "/go")

    def Barrier(self):
        self.Ibarrier().Wait()

    def Split(self, color=0, key=0):
        info = self._allgather((color, key, _RANK))
        mine = sorted((k, w) for c, k, w in info if c == color)
        return Intracomm([w for _, w in mine], f"{self.cid}.s{self._coll}c{color}")

    def Split_type(self, split_type=COMM_TYPE_SHARED, key=0, info=INFO_NULL):
        return self.Split(0, key)  # single node: everyone shares memory

    def Dup(self):
        c = self._coll
        self._coll += 1
        return Intracomm(self.members, f"{self.cid}.d{c}")

    def Create_group(self, members):
        return Intracomm([self.members[m] for m in members], f"{self.cid}.g{'_'.join(map(str, members))}")

    def Abort(self, code=1):
        os._exit(code)


class Win:
    _count = 0

    def __init__(self, comm, shm, names, sizes, disp_unit):
        self.comm, self.shm, self.names, self.sizes, self.disp_unit = comm, shm, names, sizes, disp_unit
        self._attached = {}

    @staticmethod
    def Allocate_shared(size=0, disp_unit=1, comm=None, info=INFO_NULL):
        n = comm._next("win")
        shm, name = None, ""
        if size > 0:
            name = f"bddlmpi_{os.environ.get('BDDL_MPI_SESSION', 'x')}_{comm.cid.replace('/', '_')}_{n}_{comm.rank}"
            shm = shared_memory.SharedMemory(name=name, create=True, size=size)
        table = comm._allgather((name, size))
        return Win(comm, shm, [t[0] for t in table], [t[1] for t in table], disp_unit)

    def tomemory(self):
        return self.shm.buf if self.shm is not None else memoryview(b"")

    def Shared_query(self, rank):
        if rank == self.comm.rank:
            return self.tomemory(), self.disp_unit
        if rank not in self._attached:
            s = shared_memory.SharedMemory(name=self.names[rank])
            try:  # attaching must not make this process responsible for unlinking the segment
                resource_tracker.unregister(s._name, "shared_memory")
            except Exception:
                pass
            self._attached[rank] = s
        return self._attached[rank].buf, self.disp_unit

    def Fence(self, assertion=0):
        self.comm.Barrier()

    def Lock_all(self):
        pass

    def Unlock_all(self):
        pass

    def Sync(self):
        pass

    def Free(self):
        for s in self._attached.values():
            try:
                s.close()
            except BufferError:
                pass
        self._attached = {}
        if self.shm is not None:
            try:
                self.shm.close()
            except BufferError:
                pass
            try:
                self.shm.unlink()
            except FileNotFoundError:
                pass
            self.shm = None


COMM_WORLD = Intracomm(list(range(_SIZE)), "w")
COMM_SELF = Intracomm([_RANK], f"self{_RANK}")


def _finalize():
    # keep the store's server (rank 0) alive until every rank is done, then drop bulk segments
    try:
        if _store is not None:
            COMM_WORLD.Barrier()
    except Exception:
        pass
    for c in _ALL_COMMS:
        for seg in c._segments.values():
            try:
                seg.close()
                seg.unlink()
            except Exception:
                pass


atexit.register(_finalize)
