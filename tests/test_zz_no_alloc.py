"""The fingerprint entry point for callers that cannot allocate device memory.

`bddl_batch_checksum_static` is meant for processes whose HBM is already fully owned by a
framework allocator, for steps that cannot afford the device-wide synchronisation a cudaMalloc or
cudaFree may cause, and for hosts that forbid allocation after start-up: its reduction scratch is
an array inside the library image instead of caller memory.  Here it is called through the C ABI,
twice in a row on different batches (the array keeps the first call's contents), and must give
exactly what the allocating entry point gives -- the fingerprint is an integer sum, so the
tolerance is zero, as in tests/test_checksum_gpu.py.  The file sorts last on purpose: a run that
stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from ddl_b200 import _native as N

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _with_scratch(lib, batch, rows, stride, payload, stream):
    words = int(lib.bddl_batch_checksum_scratch_words(rows, payload))
    scratch = torch.zeros(words, dtype=torch.int64, device=DEV)
    out = torch.zeros(1, dtype=torch.int64, device=DEV)
    N.check(lib.bddl_batch_checksum(C.c_void_p(batch.data_ptr()), rows, stride, payload, C.c_void_p(scratch.data_ptr()),
                                    words, C.c_void_p(out.data_ptr()), stream))
    return out


def test_static_scratch_matches_allocating_entry_point_twice_in_a_row():
    lib = N.lib()
    stream = C.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)
    g = np.random.default_rng(2024)
    # second batch is smaller than the first: stale partials of call one must not leak into call two
    for rows, payload, stride in ((512, 150528, 150528), (37, 1000, 1024)):
        host = g.integers(0, 256, size=(rows, stride), dtype=np.uint8)
        batch = torch.from_numpy(host).to(DEV)
        expect = _with_scratch(lib, batch, rows, stride, payload, stream)
        out = torch.full((1,), -1, dtype=torch.int64, device=DEV)  # caller-owned output word
        N.check(lib.bddl_batch_checksum_static(C.c_void_p(batch.data_ptr()), rows, stride, payload,
                                               C.c_void_p(out.data_ptr()), stream))
        torch.cuda.synchronize(DEV)
        assert out.item() == expect.item()


def test_static_scratch_refuses_batches_beyond_its_limit():
    lib = N.lib()
    limit = int(lib.bddl_batch_checksum_static_max_bytes())
    assert limit == 512 << 20
    small = torch.zeros(64, dtype=torch.uint8, device=DEV)
    out = torch.zeros(1, dtype=torch.int64, device=DEV)
    rows, payload = limit // 4096 + 1, 4096  # one row too many; rejected before anything is launched
    rc = lib.bddl_batch_checksum_static(C.c_void_p(small.data_ptr()), rows, payload, payload,
                                        C.c_void_p(out.data_ptr()), C.c_void_p(0))
    assert rc != 0 and b"exceeds the built-in scratch" in lib.bddl_last_error()
    torch.cuda.synchronize(DEV)
