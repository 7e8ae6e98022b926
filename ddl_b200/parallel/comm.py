"""Control-plane communicator: the subset of MPI the reference uses, over a c10d Store.

The reference's control traffic is pickled ``ssend``/``recv`` of small dataclasses, barriers,
communicator splits and one ``Sendrecv_replace`` of row blocks between producer processes
(SURVEY.md 2.2, rows M1-M7, M15).  None of it is hot on B200 -- sample bytes move inside the
fused kernel -- so the control plane is a mailbox on the key-value store that
``torch.distributed`` already runs (TCPStore under torchrun, HashStore for in-process ranks).
No process groups are created here, which keeps sub-communicators cheap and lets a group be
split without the participation of ranks outside it (as MPI's Split allows and
``dist.new_group`` does not).

Matching is FIFO per (source, destination, tag): sender and receiver each count their
messages, so no wildcard receives are needed (the reference uses none).
"""
from __future__ import annotations

import os
import pickle
import socket
import time
from datetime import timedelta
from typing import Any, List, Optional, Sequence

import numpy as np

from ..exceptions import LoaderTimeoutError

_CHUNK = 32 << 20  # bytes per store value for bulk payloads


class Comm:
    """An ordered group of world ranks sharing a store namespace."""

    def __init__(self, store, members: Sequence[int], world_rank: int, name: str, timeout_s: float = 600.0):
        if world_rank not in members:
            raise ValueError(f"rank {world_rank} is not a member of {name}")
        self.store = store
        self.members = list(members)
        self.world_rank = int(world_rank)
        self.rank = self.members.index(self.world_rank)
        self.size = len(self.members)
        self.name = name
        self.timeout_s = float(timeout_s)
        self._seq: dict = {}
        self._coll = 0
        try:
            store.set_timeout(timedelta(seconds=self.timeout_s))
        except Exception:
            pass

    # -- MPI-like accessors ---------------------------------------------------------------------
    def Get_rank(self) -> int:
        return self.rank

    def Get_size(self) -> int:
        return self.size

    def Get_name(self) -> str:
        return self.name

    def Set_name(self, name: str) -> None:  # names are part of the key space: keep the original
        self._display = name

    # -- helpers --------------------------------------------------------------------------------
    def _next(self, kind: str, a: int, b: int, tag: int) -> int:
        k = (kind, a, b, tag)
        v = self._seq.get(k, 0)
        self._seq[k] = v + 1
        return v

    def _get(self, key: str) -> bytes:
        try:
            return self.store.get(key)
        except Exception as e:  # store timeouts surface as RuntimeError / DistStoreError
            raise LoaderTimeoutError(f"{self.name}: no message under {key!r} within {self.timeout_s}s ({e})") from e

    def _delete(self, key: str) -> None:
        try:
            self.store.delete_key(key)
        except Exception:
            pass

    # -- point to point -------------------------------------------------------------------------
    def send_obj(self, obj: Any, dest: int, tag: int = 0, _sync: bool = False) -> None:
        seq = self._next("s", self.rank, dest, tag)
        self.store.set(f"{self.name}/m/{self.rank}>{dest}/{tag}/{seq}",
                       pickle.dumps((_sync, obj), protocol=pickle.HIGHEST_PROTOCOL))

    def recv_obj(self, source: int, tag: int = 0) -> Any:
        seq = self._next("r", source, self.rank, tag)
        key = f"{self.name}/m/{source}>{self.rank}/{tag}/{seq}"
        data = self._get(key)
        self._delete(key)
        sync, obj = pickle.loads(data)
        if sync:
            self.store.set(key + "/ack", b"1")
        return obj

    def ssend_obj(self, obj: Any, dest: int, tag: int = 0) -> None:
        """Synchronous send: returns once the receiver has taken the message (MPI ssend)."""
        seq = self._seq.get(("s", self.rank, dest, tag), 0)
        self.send_obj(obj, dest, tag, _sync=True)
        key = f"{self.name}/m/{self.rank}>{dest}/{tag}/{seq}/ack"
        self._get(key)
        self._delete(key)

    # lower-case MPI spellings used by reference-style code
    ssend = ssend_obj
    send = send_obj

    def recv(self, source: int, tag: int = 0) -> Any:
        return self.recv_obj(source, tag)

    # -- collectives ----------------------------------------------------------------------------
    def allgather_obj(self, obj: Any) -> List[Any]:
        c = self._coll
        self._coll += 1
        base = f"{self.name}/ag/{c}"
        self.store.set(f"{base}/{self.rank}", pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))
        out = [pickle.loads(self._get(f"{base}/{r}")) for r in range(self.size)]
        # last reader cleans up
        if self.store.add(f"{base}/read", 1) == self.size:
            for r in range(self.size):
                self._delete(f"{base}/{r}")
            self._delete(f"{base}/read")
        return out

    def bcast_obj(self, obj: Any, root: int = 0) -> Any:
        c = self._coll
        self._coll += 1
        key = f"{self.name}/bc/{c}"
        if self.rank == root:
            self.store.set(key, pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))
            res = obj
        else:
            res = pickle.loads(self._get(key))
        if self.store.add(key + "/read", 1) == self.size:
            self._delete(key)
            self._delete(key + "/read")
        return res

    def barrier(self) -> None:
        if self.size == 1:
            return
        c = self._coll
        self._coll += 1
        key = f"{self.name}/bar/{c}"
        if self.store.add(key, 1) == self.size:
            self.store.set(key + "/go", b"1")
        self._get(key + "/go")
        if self.store.add(key + "/left", 1) == self.size:
            for k in (key, key + "/go", key + "/left"):
                self._delete(k)

    Barrier = barrier

    def split(self, color: int, key: Optional[int] = None, name: Optional[str] = None) -> "Comm":
        """MPI_Comm_split: members with equal ``color`` form a new communicator ordered by key."""
        key = self.rank if key is None else key
        info = self.allgather_obj((int(color), int(key), self.world_rank))
        mine = sorted((k, w) for c, k, w in info if c == color)
        members = [w for _, w in mine]
        c = self._coll  # same on every member: collective counters advance in lock-step
        return Comm(self.store, members, self.world_rank, f"{self.name}/{name or 'split'}{c}c{color}", self.timeout_s)

    Split = split

    def dup(self, name: str = "dup") -> "Comm":
        c = self._coll
        self._coll += 1
        return Comm(self.store, self.members, self.world_rank, f"{self.name}/{name}{c}", self.timeout_s)

    Dup = dup

    def hostnames(self) -> List[str]:
        return self.allgather_obj(socket.gethostname())

    # -- bulk (the reference's Sendrecv_replace, ddl/shuffle.py:95-108) -----------------------------
    def sendrecv_replace(self, array: np.ndarray, dest: int, source: int, sendtag: int = 0, recvtag: int = 0) -> None:
        """Send ``array`` to ``dest`` and overwrite it with what ``source`` sends (in place).

        Payloads of peers on the same node travel through a POSIX shared-memory segment (one
        memcpy in, one out); only its name goes through the store.  Peers on other nodes, and
        tiny payloads, use store values in chunks."""
        if not array.flags.c_contiguous:
            raise ValueError("sendrecv_replace needs a C-contiguous array")
        buf = array.reshape(-1).view(np.uint8)
        nbytes = int(buf.size)
        seq_s = self._next("bs", self.rank, dest, sendtag)
        base_s = f"{self.name}/b/{self.rank}>{dest}/{sendtag}/{seq_s}"
        seq_r = self._next("br", source, self.rank, recvtag)
        base_r = f"{self.name}/b/{source}>{self.rank}/{recvtag}/{seq_r}"
        via_shm = nbytes >= (1 << 16) and self._same_host(dest)
        seg = None
        if via_shm:
            import ctypes as C

            from .. import _native as N

            name = "/bddl_sr_%x" % (abs(hash((base_s, os.getpid(), time.monotonic_ns()))) % (1 << 60))
            p = C.c_void_p()
            N.check(N.lib().bddl_shm_create(name.encode(), nbytes, C.byref(p)))
            N.shm_created(name)  # unlinked at exit should this process die before the ack arrives
            C.memmove(p.value, buf.ctypes.data, nbytes)
            seg = (name, p.value)
            self.store.set(base_s, pickle.dumps(("shm", name, nbytes)))
        else:
            n_chunks = max(1, (nbytes + _CHUNK - 1) // _CHUNK)
            self.store.set(base_s, pickle.dumps(("store", n_chunks, nbytes)))
            for i in range(n_chunks):
                self.store.set(f"{base_s}/{i}", buf[i * _CHUNK:(i + 1) * _CHUNK].tobytes())
        kind, a, peer_bytes = pickle.loads(self._get(base_r))
        self._delete(base_r)
        if peer_bytes != nbytes:
            raise ValueError(f"{self.name}: sendrecv_replace size mismatch with rank {source} "
                             f"({peer_bytes} != {nbytes} bytes)")
        if kind == "shm":
            import ctypes as C

            from .. import _native as N

            q = C.c_void_p()
            N.check(N.lib().bddl_shm_attach(a.encode(), nbytes, C.byref(q)))
            C.memmove(buf.ctypes.data, q.value, nbytes)
            N.lib().bddl_shm_detach(q, nbytes)
        else:
            for i in range(a):
                data = self._get(f"{base_r}/{i}")
                buf[i * _CHUNK:(i + 1) * _CHUNK] = np.frombuffer(data, dtype=np.uint8)
                self._delete(f"{base_r}/{i}")
        self.store.set(base_r + "/ack", b"1")
        self._get(base_s + "/ack")  # my payload has been copied out
        self._delete(base_s + "/ack")
        if seg is not None:
            import ctypes as C

            from .. import _native as N

            N.lib().bddl_shm_detach(C.c_void_p(seg[1]), nbytes)
            N.lib().bddl_shm_unlink(seg[0].encode())
            N.shm_unlinked(seg[0])

    def _same_host(self, peer: int) -> bool:
        if not hasattr(self, "_hosts"):
            self._hosts = None
        if self._hosts is None:
            # filled lazily by the first collective use; sendrecv_replace is always called by the
            # whole group in the reference's exchange, so a collective here is safe
            self._hosts = self.hostnames()
        return self._hosts[peer] == self._hosts[self.rank]

    Sendrecv_replace = sendrecv_replace

    def abort(self, code: int = 1) -> None:
        """MPI_Abort: tell everyone, then leave (reference ddl/ddl_env.py:30)."""
        try:
            self.store.set(f"{self.name}/abort", str(code).encode())
        finally:
            os._exit(code)

    Abort = abort

    def __repr__(self) -> str:
        return f"Comm({self.name!r}, rank={self.rank}/{self.size})"


class World:
    """Process-wide bootstrap: the store, this process's rank and the world size."""

    def __init__(self, store, rank: int, size: int, name: str = "ddl"):
        self.store, self.rank, self.size = store, rank, size
        self.comm = Comm(store, list(range(size)), rank, name)

    @staticmethod
    def from_environment(timeout_s: float = 600.0) -> "World":
        """torchrun / torch.distributed if present, else a one-rank world."""
        import torch
        import torch.distributed as dist

        world_size = int(os.environ.get("WORLD_SIZE", "1"))
        if not dist.is_initialized() and world_size > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            use_cuda = torch.cuda.is_available()
            if use_cuda:
                torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())
            backend = "cpu:gloo,cuda:nccl" if use_cuda else "gloo"
            dist.init_process_group(backend=backend, timeout=timedelta(seconds=timeout_s))
        if dist.is_initialized():
            store = dist.PrefixStore("ddl_b200", dist.distributed_c10d._get_default_store())
            return World(store, dist.get_rank(), dist.get_world_size())
        return World(dist.HashStore(), 0, 1)

    @staticmethod
    def local(size: int):
        """``size`` in-process worlds sharing one HashStore (one per thread), for tests."""
        import torch.distributed as dist

        store = dist.HashStore()
        uid = f"ddl{time.monotonic_ns()}"
        return [World(store, r, size, name=uid) for r in range(size)]
