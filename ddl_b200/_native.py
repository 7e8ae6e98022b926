"""ctypes binding of the native library (``ddl_b200/lib/libddl_b200.so``, C ABI in
``include/ddl_b200.h``).

The reference (maximilian-tech/ddl) is pure Python over mpi4py and has no native layer; this
module is where the framework's hand-written sm_100a kernels and C++ runtime are reached from
Python.  The library is built in-tree by ``__graft_entry__.build()`` / ``python -m
ddl_b200.build`` and loads on machines without a GPU (the host RNG, partner selection and
shuffles need none).  On a machine with a GPU a missing library is an error, never a silent
fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libddl_b200.so")
_lib: Optional[C.CDLL] = None


class NativeError(RuntimeError):
    """A native entry point returned a non-zero code."""

    def __init__(self, code: int, message: str):
        super().__init__(f"ddl_b200 native error {code}: {message}")
        self.code = code


class RngState(C.Structure):
    """Mirror of ``bddl_rng_state``: numpy PCG64 state + the buffered 32-bit half."""

    _fields_ = [
        ("state_lo", C.c_uint64),
        ("state_hi", C.c_uint64),
        ("inc_lo", C.c_uint64),
        ("inc_hi", C.c_uint64),
        ("has_uint32", C.c_uint32),
        ("uinteger", C.c_uint32),
        ("draws32", C.c_uint64),
    ]


MAX_CHANNELS = 8
SLOT_ROW_BITS = 48

CVT_COPY, CVT_U8_BF16, CVT_F32_BF16, CVT_U8_F32, CVT_I32_I64 = range(5)
VARIANT_AUTO, VARIANT_LDG, VARIANT_TMA = range(3)


class GatherDesc(C.Structure):
    """Mirror of ``bddl_gather_desc``."""

    _fields_ = [
        ("d_src_base", C.c_void_p),
        ("d_src_offsets", C.c_void_p),
        ("d_src_lengths", C.c_void_p),
        ("d_slots", C.c_void_p),
        ("d_dst", C.c_void_p),
        ("row_bytes", C.c_int64),
        ("sample_bytes", C.c_int64),
        ("dst_row_bytes", C.c_int64),
        ("chan_stride", C.c_int64),
        ("n_samples", C.c_int32),
        ("n_sources", C.c_int32),
        ("mode", C.c_int32),
        ("n_channels", C.c_int32),
        ("scale", C.c_float * MAX_CHANNELS),
        ("bias", C.c_float * MAX_CHANNELS),
        ("pad_bits", C.c_uint32),
        ("variant", C.c_int32),
        ("max_ctas", C.c_int32),
        ("reserved0", C.c_int32),
        ("reserved1", C.c_int32),
        ("col_offset_bytes", C.c_int32),
        ("tune", C.c_int32),
    ]


# flag page layout of the device transport (mirror of the BDDL_FLAG_* constants)
FLAG_PAGE_BYTES, FLAG_MAX_RANKS, FLAG_MAX_SLOTS, FLAG_MAX_WINDOWS = 16384, 64, 16, 24
FLAG_BARRIER0, FLAG_READY0, FLAG_STOP, FLAG_FREE0, FLAG_DONE0, FLAG_TABLE0 = 0, 256, 272, 512, 1536, 2560
FLAG_DONE_MC0 = 288


class PushDesc(C.Structure):
    """Mirror of ``bddl_push_desc``."""

    _fields_ = [
        ("d_tables", C.c_void_p),
        ("d_rings", C.c_void_p),
        ("d_flags", C.c_void_p),
        ("table_offset_bytes", C.c_int64),
        ("ring_offset_bytes", C.c_int64),
        ("timeout_ns", C.c_uint64),
        ("d_status", C.c_void_p),
        ("n_ranks", C.c_int32),
        ("me", C.c_int32),
        ("ring_slot", C.c_int32),
        ("window", C.c_int32),
        ("owner_lo", C.c_int32),
        ("owner_hi", C.c_int32),
        ("ctas_per_consumer", C.c_int32),
        ("use", C.c_uint32),
        ("table_seq", C.c_uint32),
        ("d_mc_ring", C.c_void_p),
        ("table_rank", C.c_int32),
        ("fills", C.c_uint32),
    ]


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _declare(lib: C.CDLL) -> None:
    P = C.POINTER
    vp, i32, i64, u32, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64
    sigs = {
        "bddl_last_error": (C.c_char_p, []),
        "bddl_abi_version": (C.c_int, []),
        "bddl_rng_seed": (None, [P(RngState), u64]),
        "bddl_rng_next32": (u32, [P(RngState)]),
        "bddl_rng_next64": (u64, [P(RngState)]),
        "bddl_rng_interval": (u64, [P(RngState), u64]),
        "bddl_rng_advance": (None, [P(RngState), u64]),
        "bddl_host_shuffle_u64": (C.c_int, [P(RngState), vp, i64]),
        "bddl_host_shuffle_i32": (C.c_int, [P(RngState), vp, i64]),
        "bddl_host_shuffle_rows": (C.c_int, [P(RngState), vp, i64, i64]),
        "bddl_partner_select": (C.c_int, [P(RngState), i32, i32, P(i32), P(i32), vp, P(i32)]),
        "bddl_perm_gen": (C.c_int, [vp, i64, i64, vp, i32, vp, i64, vp]),
        "bddl_perm_gen_scratch_bytes": (i64, [i64]),
        "bddl_exchange_gather_collate": (C.c_int, [P(GatherDesc), vp]),
        "bddl_launch_count": (i64, []),
        "bddl_push_ctas_per_consumer": (C.c_int, [i32, i32, P(i32)]),
        "bddl_exchange_push_collate": (C.c_int, [P(GatherDesc), P(PushDesc), vp]),
    }
    sz = C.c_size_t
    sigs.update({
        "bddl_index_exchange": (C.c_int, [vp, vp, vp, vp, i64, i64, vp]),
        "bddl_slot_iota": (C.c_int, [vp, i64, i32, i64, vp]),
        "bddl_device_count": (C.c_int, [P(i32)]),
        "bddl_set_device": (C.c_int, [i32]),
        "bddl_get_device": (C.c_int, [P(i32)]),
        "bddl_malloc": (C.c_int, [P(vp), sz]),
        "bddl_free": (C.c_int, [vp]),
        "bddl_memset_async": (C.c_int, [vp, C.c_int, sz, vp]),
        "bddl_memcpy_async": (C.c_int, [vp, vp, sz, i32, vp]),
        "bddl_memcpy2d_async": (C.c_int, [vp, sz, vp, sz, sz, sz, i32, vp]),
        "bddl_stream_synchronize": (C.c_int, [vp]),
        "bddl_event_create": (C.c_int, [P(vp)]),
        "bddl_event_destroy": (C.c_int, [vp]),
        "bddl_event_record": (C.c_int, [vp, vp]),
        "bddl_stream_wait_event": (C.c_int, [vp, vp]),
        "bddl_event_synchronize": (C.c_int, [vp]),
        "bddl_gather_step": (C.c_int, [P(GatherDesc), vp, vp, vp]),
        "bddl_ipc_get_handle": (C.c_int, [vp, vp]),
        "bddl_ipc_open_handle": (C.c_int, [vp, P(vp)]),
        "bddl_ipc_close_handle": (C.c_int, [vp]),
        "bddl_can_access_peer": (C.c_int, [i32, i32, P(i32)]),
        "bddl_enable_peer_access": (C.c_int, [i32]),
        "bddl_host_alloc": (C.c_int, [P(vp), P(vp), sz]),
        "bddl_host_free": (C.c_int, [vp]),
        "bddl_host_register": (C.c_int, [vp, sz, P(vp)]),
        "bddl_host_unregister": (C.c_int, [vp]),
        "bddl_stage_rows": (C.c_int, [vp, i64, vp, i64, vp, i64, i64, vp, P(i64)]),
        "bddl_stage_slots": (C.c_int, [vp, i64, vp, i32, i64, vp, i64, i64, vp, P(i64)]),
        "bddl_batch_checksum_scratch_words": (i64, [i64, i64]),
        "bddl_batch_checksum": (C.c_int, [vp, i64, i64, i64, vp, i64, vp, vp]),
        "bddl_batch_checksum_static_max_bytes": (i64, []),
        "bddl_batch_checksum_static": (C.c_int, [vp, i64, i64, i64, vp, vp]),
        "bddl_vmm_query": (C.c_int, [i32, P(i32), P(i32), P(i32)]),
        "bddl_vmm_alloc": (C.c_int, [sz, i32, sz, P(vp), P(u64), P(i32), P(sz)]),
        "bddl_vmm_import": (C.c_int, [i32, sz, i32, P(vp), P(u64)]),
        "bddl_vmm_free": (C.c_int, [vp, sz, u64]),
        "bddl_mc_granularity": (C.c_int, [sz, i32, P(sz)]),
        "bddl_mc_create": (C.c_int, [sz, i32, P(u64), P(i32)]),
        "bddl_mc_import": (C.c_int, [i32, P(u64)]),
        "bddl_mc_add_device": (C.c_int, [u64, i32]),
        "bddl_mc_bind_and_map": (C.c_int, [u64, u64, sz, i32, P(vp)]),
        "bddl_multicast_put": (C.c_int, [vp, vp, i64, vp]),
        "bddl_stage_slots_mt": (C.c_int, [vp, i64, vp, i32, i64, vp, i64, i64, vp, i32]),
        "bddl_host_gather_slots": (C.c_int, [vp, i64, vp, i32, i64, vp, i64, i64, i32]),
        "bddl_host_permute_rows": (C.c_int, [vp, vp, vp, i64, vp, i64, i64, i32]),
        "bddl_stream_wait_value32": (C.c_int, [vp, vp, u32]),
        "bddl_stream_write_value32": (C.c_int, [vp, vp, u32]),
        "bddl_flag_signal": (C.c_int, [vp, i32, u32, i32, vp]),
        "bddl_flag_wait": (C.c_int, [vp, u32, u64, vp, vp, vp]),
        "bddl_flag_barrier": (C.c_int, [vp, i32, i32, u32, u64, vp, vp, vp]),
        "bddl_shm_create": (C.c_int, [C.c_char_p, sz, P(vp)]),
        "bddl_shm_attach": (C.c_int, [C.c_char_p, sz, P(vp)]),
        "bddl_shm_detach": (C.c_int, [vp, sz]),
        "bddl_shm_unlink": (C.c_int, [C.c_char_p]),
        "bddl_hflag_store": (None, [vp, u32]),
        "bddl_hflag_load": (u32, [vp]),
        "bddl_hflag_add": (u32, [vp, u32]),
        "bddl_hflag_wait_geq": (C.c_int, [vp, u32, i64, vp]),
    })
    optional = {}
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    for name, (res, args) in optional.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args


def lib() -> C.CDLL:
    """Load (once) and return the native library; build it first if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            from . import build as _build

            _build.build_native(verbose=False)
        _lib = C.CDLL(_LIB_PATH)
        _declare(_lib)
    return _lib


def check(code: int) -> None:
    if code != 0:
        msg = lib().bddl_last_error()
        raise NativeError(code, msg.decode("utf-8", "replace") if msg else "")


def launch_count() -> int:
    return int(lib().bddl_launch_count())


# POSIX shared-memory names created by this process and not yet unlinked: removed at exit so that a
# crash between create and unlink does not leave segments behind in /dev/shm.
_shm_live: set = set()


def shm_created(name: str) -> None:
    _shm_live.add(name)


def shm_unlinked(name: str) -> None:
    _shm_live.discard(name)


def _unlink_leftovers() -> None:
    if _lib is None:
        return
    for name in list(_shm_live):
        try:
            _lib.bddl_shm_unlink(name.encode())
        except Exception:
            pass
    _shm_live.clear()


import atexit as _atexit  # noqa: E402

_atexit.register(_unlink_leftovers)
