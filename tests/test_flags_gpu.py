"""N8 flags on a GPU: stream memory ops, release/acquire signal + bounded wait kernels and the
flag barrier (single rank here; the multi-rank form runs in the multi-GPU loader test).
Waits are only ever issued after their signal, so no kernel spins on another launch
(B200_PROFILING.md: no co-resident spinning kernels on one GPU)."""
import ctypes as C

import pytest
import torch

from ddl_b200 import _native as N

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _s():
    return C.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)


def test_stream_write_then_wait_value32():
    flag = torch.zeros(4, dtype=torch.int32, device=DEV)
    L = N.lib()
    N.check(L.bddl_stream_write_value32(_s(), C.c_void_p(flag.data_ptr() + 4), 7))
    N.check(L.bddl_stream_wait_value32(_s(), C.c_void_p(flag.data_ptr() + 4), 5))  # 7 >= 5: passes
    torch.cuda.synchronize()
    assert flag.tolist() == [0, 7, 0, 0]


def test_signal_store_add_and_bounded_wait():
    L = N.lib()
    flags = torch.zeros(8, dtype=torch.int32, device=DEV)
    targets = torch.tensor([flags.data_ptr() + 4 * i for i in (1, 3, 5)], dtype=torch.int64, device=DEV)
    status = torch.full((1,), 9, dtype=torch.int32, device=DEV)
    N.check(L.bddl_flag_signal(C.c_void_p(targets.data_ptr()), 3, 4, 0, _s()))
    N.check(L.bddl_flag_signal(C.c_void_p(targets.data_ptr()), 3, 2, 1, _s()))
    N.check(L.bddl_flag_wait(C.c_void_p(flags.data_ptr() + 4), 6, 1000, None, C.c_void_p(status.data_ptr()), _s()))
    torch.cuda.synchronize()
    assert flags.tolist() == [0, 6, 0, 6, 0, 6, 0, 0] and status.item() == 0
    # a wait that cannot be satisfied times out instead of hanging
    N.check(L.bddl_flag_wait(C.c_void_p(flags.data_ptr()), 1, 50, None, C.c_void_p(status.data_ptr()), _s()))
    torch.cuda.synchronize()
    assert status.item() == 1
    # ... and is cancelled by the stop word
    stop = torch.ones(1, dtype=torch.int32, device=DEV)
    N.check(L.bddl_flag_wait(C.c_void_p(flags.data_ptr()), 1, 1 << 40, C.c_void_p(stop.data_ptr()),
                             C.c_void_p(status.data_ptr()), _s()))
    torch.cuda.synchronize()
    assert status.item() == 2


def test_flag_barrier_single_rank_epochs():
    L = N.lib()
    page = torch.zeros(64, dtype=torch.int32, device=DEV)
    pages = torch.tensor([page.data_ptr()], dtype=torch.int64, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    for epoch in (1, 2, 3):
        N.check(L.bddl_flag_barrier(C.c_void_p(pages.data_ptr()), 1, 0, epoch, 1000, None,
                                    C.c_void_p(status.data_ptr()), _s()))
    torch.cuda.synchronize()
    assert page[0].item() == 3 and status.item() == 0
