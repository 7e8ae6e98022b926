"""Replicated delivery: one buffer with a copy on every GPU, written once by its owner.

Not in the reference.  Where the platform has NVLS multicast
(``CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED``, NVSwitch + fabric manager) the copies are bound to
one multicast object and the owner's kernel stores through the multicast mapping with
``multimem.st``: NVSwitch replicates the write into every GPU's HBM, so a sample every consumer
needs (a shared validation batch, class weights, a broadcast of labels) crosses each link once
instead of once per consumer from the owner's side.  Elsewhere the same API falls back to one
peer copy per consumer over the CUDA-IPC heap.

Set-up is collective: rank 0 creates the multicast object and passes its POSIX file descriptor to
the other processes over an abstract Unix socket (SCM_RIGHTS); every rank adds its device, creates
physical memory with the VMM API, binds it and maps the object.
"""
from __future__ import annotations

import array
import ctypes as C
import os
import socket
import uuid
from typing import Optional

import torch

from .. import _native as N
from .comm import Comm
from .heap import SymmetricHeap, _CudaView


def query(device: torch.device) -> dict:
    v, f, m = C.c_int32(), C.c_int32(), C.c_int32()
    N.check(N.lib().bddl_vmm_query(torch.device(device).index or 0, C.byref(v), C.byref(f), C.byref(m)))
    return {"vmm": bool(v.value), "posix_fd": bool(f.value), "multicast": bool(m.value)}


def _share_fd(comm: Comm, fd: Optional[int], root: int = 0) -> int:
    """Rank `root` hands `fd` to every other rank (SCM_RIGHTS over an abstract Unix socket);
    returns each rank's own descriptor of the same object."""
    if comm.size == 1:
        return fd
    name = comm.bcast_obj("\0bddl_fd_" + uuid.uuid4().hex[:12] if comm.rank == root else None, root)
    if comm.rank == root:
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name)
        srv.listen(comm.size)
        comm.barrier()
        for _ in range(comm.size - 1):
            c, _ = srv.accept()
            c.sendmsg([b"f"], [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))])
            c.recv(1)
            c.close()
        srv.close()
        return fd
    comm.barrier()
    s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    s.connect(name)
    msg, anc, _, _ = s.recvmsg(1, socket.CMSG_LEN(4))
    got = array.array("i")
    for level, typ, data in anc:
        if level == socket.SOL_SOCKET and typ == socket.SCM_RIGHTS:
            got.frombytes(data[:4])
    s.send(b"k")
    s.close()
    if not got:
        raise RuntimeError("no file descriptor received for the multicast object")
    return int(got[0])


class MulticastBuffer:
    """``nbytes`` replicated on every rank of ``comm``.  ``local`` is this rank's copy (uint8
    tensor); ``put(src)`` on any rank writes ``src`` into *all* copies."""

    def __init__(self, comm: Comm, device: torch.device, nbytes: int, allow_fallback: bool = True):
        self.comm, self.device = comm, torch.device(device)
        self.nbytes = (int(nbytes) + 15) // 16 * 16
        lib, dev = N.lib(), self.device.index or 0
        caps = comm.allgather_obj(query(self.device))
        self.mode = "multicast" if all(c["vmm"] and c["posix_fd"] and c["multicast"] for c in caps) else "unicast"
        self._heap = None
        if self.mode == "multicast":
            try:
                self._init_multicast(lib, dev)
            except N.NativeError as e:
                ok = False
                self._mc_error = str(e)
            else:
                ok = True
            if not all(comm.allgather_obj(ok)):
                if not allow_fallback:
                    raise RuntimeError(f"multicast set-up failed: {getattr(self, '_mc_error', 'on a peer')}")
                self.mode = "unicast"
        if self.mode == "unicast":
            if not allow_fallback:
                raise RuntimeError(f"NVLS multicast is not available here: {caps}")
            self._heap = SymmetricHeap(comm, self.device)
            buf = self._heap.alloc("replica", self.nbytes)
            self._heap.publish()
            self.local = buf.bytes()[: self.nbytes]
            self._peer_ptrs = [self._heap.peer_ptr(r, "replica") for r in range(comm.size)]

    def _init_multicast(self, lib, dev) -> None:
        comm = self.comm
        gran = C.c_size_t()
        N.check(lib.bddl_mc_granularity(self.nbytes, comm.size, C.byref(gran)))
        self.size = (self.nbytes + gran.value - 1) // gran.value * gran.value
        mc, fd = C.c_uint64(), C.c_int32(-1)
        if comm.rank == 0:
            N.check(lib.bddl_mc_create(self.size, comm.size, C.byref(mc), C.byref(fd)))
        my_fd = _share_fd(comm, fd.value if comm.rank == 0 else None)
        if comm.rank != 0:
            N.check(lib.bddl_mc_import(my_fd, C.byref(mc)))
        os.close(my_fd)
        N.check(lib.bddl_mc_add_device(mc.value, dev))
        comm.barrier()  # every device has joined before anyone binds memory
        ptr, handle, mfd, rounded = C.c_void_p(), C.c_uint64(), C.c_int32(), C.c_size_t()
        N.check(lib.bddl_vmm_alloc(self.size, dev, gran.value, C.byref(ptr), C.byref(handle), C.byref(mfd),
                                   C.byref(rounded)))
        os.close(mfd.value)
        mc_ptr = C.c_void_p()
        N.check(lib.bddl_mc_bind_and_map(mc.value, handle.value, self.size, dev, C.byref(mc_ptr)))
        comm.barrier()
        self._ptr, self._handle, self._mc, self._mc_ptr = ptr.value, handle.value, mc.value, mc_ptr.value
        self.local = torch.as_tensor(_CudaView(self._ptr, self.size, self), device=self.device)[: self.nbytes]
        self.local.zero_()
        torch.cuda.synchronize(self.device)
        comm.barrier()

    def put(self, src: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Write ``src`` (contiguous, on this device, multiple of 16 bytes) into every copy."""
        raw = src.contiguous().view(torch.uint8).reshape(-1)
        if raw.numel() > self.nbytes or raw.numel() % 16:
            raise ValueError("multicast put needs a multiple of 16 bytes that fits the buffer")
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        sp = C.c_void_p(s.cuda_stream)
        if self.mode == "multicast":
            N.check(N.lib().bddl_multicast_put(C.c_void_p(self._mc_ptr), C.c_void_p(raw.data_ptr()), raw.numel(), sp))
        else:
            for p in self._peer_ptrs:  # one NVLink copy per consumer
                N.check(N.lib().bddl_memcpy_async(C.c_void_p(p), C.c_void_p(raw.data_ptr()), raw.numel(), 4, sp))

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        self.comm.barrier()
        self.local = None
        if self._heap is not None:
            self._heap.close()
            self._heap = None
