"""Symmetric heap: named buffers every rank allocates and every peer can address.

The reference maps every producer's float32 array into all ranks of an instance with MPI-3
shared-memory windows (``Win.Allocate_shared`` + ``Shared_query``,
reference ddl/connection.py:88-139).  Here:

* on a GPU each rank ``cudaMalloc``s its buffers through the native library and publishes CUDA
  IPC handles over the control plane; peers map them, after which a kernel reaches remote
  shards, slot tables and batch buffers through NVLink with ordinary pointers;
* on the CPU (tests, the CPU execution path) the buffers are POSIX shared memory, so the same
  multi-rank logic runs without a GPU;
* ranks that share a process (threads in tests, emulated ranks on one GPU) skip the handle
  export and use each other's pointers directly.
"""
from __future__ import annotations

import ctypes as C
import os
import uuid
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import _native as N
from .comm import Comm


class _CudaView:
    """Minimal ``__cuda_array_interface__`` carrier so torch can alias native memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class Buffer:
    """One heap allocation: raw address + typed views."""

    def __init__(self, name: str, ptr: int, nbytes: int, device: torch.device, keep=None, shm_name: str = ""):
        self.name, self.ptr, self.nbytes, self.device = name, int(ptr), int(nbytes), device
        self._keep = keep
        self.shm_name = shm_name
        self._u8: Optional[torch.Tensor] = None

    def bytes(self) -> torch.Tensor:
        """The whole buffer as a uint8 tensor on its device (aliases the memory)."""
        if self._u8 is None:
            if self.device.type == "cuda":
                self._u8 = torch.as_tensor(_CudaView(self.ptr, self.nbytes, self), device=self.device)
            else:
                arr = np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr))
                self._u8 = torch.from_numpy(arr)
        return self._u8

    def tensor(self, dtype: torch.dtype, shape, offset_bytes: int = 0) -> torch.Tensor:
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        return self.bytes()[offset_bytes:offset_bytes + n].view(dtype).view(*shape)

    def numpy(self, dtype=np.uint8, shape=None, offset_bytes: int = 0) -> np.ndarray:
        if self.device.type != "cpu":
            raise ValueError("numpy views exist for host buffers only")
        dt = np.dtype(dtype)
        count = (self.nbytes - offset_bytes) // dt.itemsize if shape is None else int(np.prod(shape))
        arr = np.ctypeslib.as_array((C.c_uint8 * (count * dt.itemsize)).from_address(self.ptr + offset_bytes))
        arr = arr.view(dt)
        return arr if shape is None else arr.reshape(shape)


class SymmetricHeap:
    """Per-rank allocator + peer address table.  ``alloc`` then ``publish`` (collective)."""

    def __init__(self, comm: Comm, device: torch.device):
        self.comm = comm
        self.device = torch.device(device)
        self.local: Dict[str, Buffer] = {}
        self.peers: Dict[Tuple[int, str], Tuple[int, int]] = {}  # (rank, name) -> (ptr, nbytes)
        self._opened: List[int] = []
        self._host_dev: Dict[Tuple[int, str], int] = {}
        self._registered: List[int] = []
        self._attached: List[Tuple[int, int]] = []
        self._vmm_imports: List[Tuple[int, int, int]] = []
        self._session = comm.bcast_obj(uuid.uuid4().hex[:12] if comm.rank == 0 else None, 0)
        self._closed = False
        if self.device.type == "cuda":
            N.check(N.lib().bddl_set_device(self.device.index or 0))

    # -- allocation -----------------------------------------------------------------------------
    def alloc(self, name: str, nbytes: int, zero: bool = True) -> Buffer:
        if name in self.local:
            raise ValueError(f"heap buffer {name!r} exists")
        nbytes = max(16, (int(nbytes) + 255) // 256 * 256)
        lib = N.lib()
        p = C.c_void_p()
        if self.device.type == "cuda":
            N.check(lib.bddl_set_device(self.device.index or 0))
            N.check(lib.bddl_malloc(C.byref(p), nbytes))
            buf = Buffer(name, p.value, nbytes, self.device)
            if zero:
                buf.bytes().zero_()
        else:
            shm = f"/bddl_{self._session}_{self.comm.world_rank}_{len(self.local)}"
            N.check(lib.bddl_shm_create(shm.encode(), nbytes, C.byref(p)))
            N.shm_created(shm)
            buf = Buffer(name, p.value, nbytes, self.device, shm_name=shm)
        self.local[name] = buf
        return buf

    def alloc_pinned(self, name: str, nbytes: int) -> Tuple[Buffer, int]:
        """Pinned host memory mapped into the device: (host buffer, device-visible address)."""
        nbytes = max(16, int(nbytes))
        h, d = C.c_void_p(), C.c_void_p()
        N.check(N.lib().bddl_host_alloc(C.byref(h), C.byref(d), nbytes))
        buf = Buffer(name, h.value, nbytes, torch.device("cpu"), keep=("pinned", h.value))
        self.local[name] = buf
        return buf, int(d.value)

    def alloc_host_shared(self, name: str, nbytes: int) -> Buffer:
        """Pinned host memory every process of the node can map: POSIX shm + cudaHostRegister.
        The buffer is published like device buffers; peers register it too, so any GPU can copy
        from it or read it from a kernel (host-resident datasets)."""
        nbytes = max(4096, (int(nbytes) + 4095) // 4096 * 4096)
        lib = N.lib()
        p = C.c_void_p()
        shm = f"/bddl_{self._session}_{self.comm.world_rank}_h{len(self.local)}"
        N.check(lib.bddl_shm_create(shm.encode(), nbytes, C.byref(p)))
        N.shm_created(shm)
        buf = Buffer(name, p.value, nbytes, torch.device("cpu"), shm_name=shm)
        buf.host_shared = True
        buf.dev_ptr = 0
        if self.device.type == "cuda":
            d = C.c_void_p()
            N.check(lib.bddl_host_register(p, nbytes, C.byref(d)))
            buf.dev_ptr = int(d.value)
        self.local[name] = buf
        return buf

    def alloc_vmm(self, name: str, nbytes: int, multicast: bool = True) -> Buffer:
        """Collective.  A buffer made with the virtual-memory-management API instead of cudaMalloc:
        every rank creates physical memory of the same size, hands its POSIX file descriptor to all
        peers (who map it: ``peer_ptr`` works as for IPC buffers), and -- where the platform has NVLS
        multicast -- binds it to one multicast object; ``buf.mc_ptr`` is then an address through
        which one ``multimem.st`` writes every rank's copy (0 when unavailable).  The reference's
        windows are per instance (ddl/connection.py:88-139); this is what lets a batch ring be the
        target of replicated delivery."""
        from .multicast import _share_fd, query

        if self.device.type != "cuda":
            raise ValueError("VMM buffers live on CUDA devices")
        lib, comm, dev = N.lib(), self.comm, self.device.index or 0
        caps = comm.allgather_obj(query(self.device))
        if not all(c["vmm"] and c["posix_fd"] for c in caps):
            raise N.NativeError(-2, "VMM allocations with POSIX file-descriptor handles are not supported here")
        use_mc = multicast and comm.size > 1 and all(c["multicast"] for c in caps)
        gran = C.c_size_t(0)
        if use_mc:
            try:
                N.check(lib.bddl_mc_granularity(int(nbytes), comm.size, C.byref(gran)))
            except N.NativeError:
                use_mc = False
            use_mc = all(comm.allgather_obj(use_mc))
        g = max(int(gran.value), 2 << 20)
        size = (int(nbytes) + g - 1) // g * g
        ptr, handle, fd, rounded = C.c_void_p(), C.c_uint64(), C.c_int32(), C.c_size_t()
        N.check(lib.bddl_vmm_alloc(size, dev, g, C.byref(ptr), C.byref(handle), C.byref(fd), C.byref(rounded)))
        size = int(rounded.value)
        buf = Buffer(name, ptr.value, size, self.device, keep=("vmm", handle.value))
        buf.mc_ptr = 0
        buf.bytes().zero_()
        torch.cuda.synchronize(self.device)
        self.local[name] = buf
        self.peers[(comm.rank, name)] = (ptr.value, size)
        for root in range(comm.size):  # every rank's memory, mapped by every peer
            got = _share_fd(comm, fd.value if comm.rank == root else None, root)
            if comm.rank != root:
                p, h = C.c_void_p(), C.c_uint64()
                N.check(lib.bddl_vmm_import(got, size, dev, C.byref(p), C.byref(h)))
                os.close(got)
                self._vmm_imports.append((p.value, size, h.value))
                self.peers[(root, name)] = (p.value, size)
        os.close(fd.value)
        if use_mc:
            ok = True
            try:
                mc, mfd = C.c_uint64(), C.c_int32(-1)
                if comm.rank == 0:
                    N.check(lib.bddl_mc_create(size, comm.size, C.byref(mc), C.byref(mfd)))
                my_fd = _share_fd(comm, mfd.value if comm.rank == 0 else None, 0)
                if comm.rank != 0:
                    N.check(lib.bddl_mc_import(my_fd, C.byref(mc)))
                os.close(my_fd)
                N.check(lib.bddl_mc_add_device(mc.value, dev))
                comm.barrier()  # every device has joined before anyone binds memory
                mc_ptr = C.c_void_p()
                N.check(lib.bddl_mc_bind_and_map(mc.value, handle.value, size, dev, C.byref(mc_ptr)))
                buf.mc_ptr, buf.mc_handle = int(mc_ptr.value), mc.value
            except N.NativeError as e:
                ok, buf.mc_error = False, str(e)
            if not all(comm.allgather_obj(ok)):
                buf.mc_ptr = 0  # one rank could not bind: everybody falls back to unicast stores
        comm.barrier()
        return buf

    def peer_dev_ptr(self, rank: int, name: str) -> int:
        """Device-visible address of a host-shared buffer (this process's mapping of it)."""
        return self._host_dev[(rank, name)]

    # -- exchange -------------------------------------------------------------------------------
    def publish(self) -> None:
        """Collective: make every local buffer addressable by every peer."""
        lib = N.lib()
        mine = {}
        for name, b in self.local.items():
            if b._keep is not None:
                continue  # pinned host staging is private; VMM buffers were exchanged when they were made
            if getattr(b, "host_shared", False):
                mine[name] = ("hshm", b.shm_name, b.nbytes, b.ptr, os.getpid(), -1)
                self._host_dev[(self.comm.rank, name)] = b.dev_ptr
            elif self.device.type == "cuda":
                h = (C.c_uint8 * 64)()
                N.check(lib.bddl_ipc_get_handle(C.c_void_p(b.ptr), h))
                mine[name] = ("ipc", bytes(h), b.nbytes, b.ptr, os.getpid(), self.device.index or 0)
            else:
                mine[name] = ("shm", b.shm_name, b.nbytes, b.ptr, os.getpid(), -1)
        table = self.comm.allgather_obj(mine)
        for r, entries in enumerate(table):
            for name, (kind, handle, nbytes, ptr, pid, dev) in entries.items():
                if (r, name) in self.peers:
                    continue
                if kind == "hshm" and r != self.comm.rank:
                    if pid == os.getpid():
                        self.peers[(r, name)] = (ptr, nbytes)
                        self._host_dev[(r, name)] = self._host_dev.get((r, name), 0)
                        continue
                    p = C.c_void_p()
                    N.check(lib.bddl_shm_attach(handle.encode(), nbytes, C.byref(p)))
                    self._attached.append((p.value, nbytes))
                    self.peers[(r, name)] = (p.value, nbytes)
                    if self.device.type == "cuda":
                        d = C.c_void_p()
                        N.check(lib.bddl_host_register(p, nbytes, C.byref(d)))
                        self._registered.append(p.value)
                        self._host_dev[(r, name)] = int(d.value)
                elif r == self.comm.rank or pid == os.getpid() and (kind == "shm" or dev == (self.device.index or 0)):
                    self.peers[(r, name)] = (ptr, nbytes)
                elif kind == "ipc":
                    if pid == os.getpid():
                        # another device driven by this process: direct peer access
                        N.check(lib.bddl_enable_peer_access(dev))
                        self.peers[(r, name)] = (ptr, nbytes)
                        continue
                    p = C.c_void_p()
                    hb = (C.c_uint8 * 64).from_buffer_copy(handle)
                    N.check(lib.bddl_ipc_open_handle(hb, C.byref(p)))
                    self._opened.append(p.value)
                    self.peers[(r, name)] = (p.value, nbytes)
                else:
                    p = C.c_void_p()
                    N.check(lib.bddl_shm_attach(handle.encode(), nbytes, C.byref(p)))
                    self._attached.append((p.value, nbytes))
                    self.peers[(r, name)] = (p.value, nbytes)
        self.comm.barrier()
        # names can go once everybody is attached; the memory lives until the last unmap
        for b in self.local.values():
            if self.device.type == "cpu" or getattr(b, "host_shared", False):
                if b.shm_name:
                    lib.bddl_shm_unlink(b.shm_name.encode())
                    N.shm_unlinked(b.shm_name)
                    b.shm_name = ""

    def peer_ptr(self, rank: int, name: str) -> int:
        return self.peers[(rank, name)][0]

    def peer_buffer(self, rank: int, name: str) -> Buffer:
        """A peer's buffer as a local object (CPU heaps and same-process peers only)."""
        ptr, nbytes = self.peers[(rank, name)]
        return Buffer(name, ptr, nbytes, self.device, keep=("peer", rank))

    # -- teardown -------------------------------------------------------------------------------
    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        lib = N.lib()
        if self.device.type == "cuda":
            try:
                torch.cuda.synchronize(self.device)
            except Exception:
                pass
        for p in self._registered:
            lib.bddl_host_unregister(C.c_void_p(p))
        for p in self._opened:
            lib.bddl_ipc_close_handle(C.c_void_p(p))
        for p, n in self._attached:
            lib.bddl_shm_detach(C.c_void_p(p), n)
        for p, n, h in self._vmm_imports:
            lib.bddl_vmm_free(C.c_void_p(p), n, h)
        # freeing exported memory while a peer still holds it open is undefined (CUDA IPC): every
        # rank closes its imports first, then all meet, then each frees what it exported
        try:
            self.comm.barrier()
        except Exception:
            pass
        for b in self.local.values():
            b._u8 = None
            if b._keep is not None and b._keep[0] == "pinned":
                lib.bddl_host_free(C.c_void_p(b.ptr))
            elif b._keep is not None and b._keep[0] == "vmm":
                if getattr(b, "mc_ptr", 0):
                    lib.bddl_vmm_free(C.c_void_p(b.mc_ptr), b.nbytes, getattr(b, "mc_handle", 0))
                lib.bddl_vmm_free(C.c_void_p(b.ptr), b.nbytes, b._keep[1])
            elif getattr(b, "host_shared", False):
                if self.device.type == "cuda":
                    lib.bddl_host_unregister(C.c_void_p(b.ptr))
                if b.shm_name:
                    lib.bddl_shm_unlink(b.shm_name.encode())
                lib.bddl_shm_detach(C.c_void_p(b.ptr), b.nbytes)
            elif self.device.type == "cuda":
                lib.bddl_free(C.c_void_p(b.ptr))
            else:
                if b.shm_name:
                    lib.bddl_shm_unlink(b.shm_name.encode())
                lib.bddl_shm_detach(C.c_void_p(b.ptr), b.nbytes)
        self.local.clear()
        self.peers.clear()
