"""Launches every kernel of the library a few times on ONE GPU so that `ncu` can capture each of them
(profiles/ncu_*_r2.txt).  Run it plainly first (exit code 0), then under ncu, e.g.

    ncu --set full --clock-control none --import-source on \
        -k regex:'gather_|push_tma|index_exchange|slot_iota|perm_gen|flag_|checksum' -o gpurun_out/all python benchmarks/ncu_targets.py

`--steady N` only launches the uint8->bf16 batch assembly N times back to back (for a capture with
--cache-control none: L2 stays dirty between launches, so DRAM writes per launch are the steady-state ones).
"""
import argparse
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200 import _native as N  # noqa: E402
from ddl_b200.ops import Convert, GatherPlan, Pcg64, SourceTable, perm_gen, rng_states_to_tensor  # noqa: E402
from ddl_b200.ops.checksum import BatchChecksum  # noqa: E402
from ddl_b200.ops.tables import index_exchange, slot_iota  # noqa: E402

DEV = torch.device("cuda", 0)
ROW = 3 * 224 * 224


def gather_case(convert, sample_bytes, batch, rows, variant, launches):
    shard = torch.randint(0, 255, (rows, sample_bytes), dtype=torch.uint8, device=DEV)
    plan = GatherPlan(SourceTable([shard.data_ptr()], DEV), sample_bytes, sample_bytes, convert, variant=variant)
    outs = [plan.alloc_batch(batch) for _ in range(3)]
    lists = [torch.randperm(rows, device=DEV)[:batch].to(torch.int64).contiguous() for _ in range(launches)]
    torch.cuda.synchronize()
    for i in range(launches):
        plan.run(lists[i], outs[i % 3])
    torch.cuda.synchronize()
    return plan, shard


def push_case(launches=4):
    """The push kernel with one rank: this GPU is owner and consumer of every sample."""
    B, rows, k, slot = 256, 4096, 0, 0
    shard = torch.randint(0, 255, (rows, ROW), dtype=torch.uint8, device=DEV)
    cv = Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), 224 * 224)
    plan = GatherPlan(SourceTable([shard.data_ptr()], DEV), ROW, ROW, cv)
    table = torch.randperm(rows, device=DEV)[:B].to(torch.int64).contiguous()
    ring = plan.alloc_batch(B)
    flags = torch.zeros(N.FLAG_PAGE_BYTES // 4, dtype=torch.int32, device=DEV)
    status = torch.zeros(4, dtype=torch.int32, device=DEV)
    ptrs = lambda t: torch.tensor([t.data_ptr()], dtype=torch.int64, device=DEV)  # noqa: E731
    pt, pr, pf = ptrs(table), ptrs(ring), ptrs(flags)
    g = C.c_int32()
    N.check(N.lib().bddl_push_ctas_per_consumer(1, 0, C.byref(g)))
    d = N.PushDesc()
    d.d_tables, d.d_rings, d.d_flags, d.d_status = pt.data_ptr(), pr.data_ptr(), pf.data_ptr(), status.data_ptr()
    d.timeout_ns = int(5e9)
    d.n_ranks, d.me, d.ring_slot, d.window, d.owner_lo, d.owner_hi = 1, 0, slot, k, 0, 1
    d.ctas_per_consumer, d.table_rank, d.d_mc_ring = g.value, -1, None
    flags[N.FLAG_TABLE0 + k * N.FLAG_MAX_RANKS] = 1
    s = torch.cuda.current_stream(DEV).cuda_stream
    for use in range(launches):
        flags[N.FLAG_FREE0 + slot * N.FLAG_MAX_RANKS] = use  # "the consumer released fill number `use`"
        d.use, d.fills, d.table_seq = use, use, 1
        plan.push_ptr(d, B, s)
    torch.cuda.synchronize()
    assert int(status[0]) == 0, "a bounded wait of the push kernel gave up"
    ref = torch.empty_like(ring)
    plan.run(table, ref)
    torch.cuda.synchronize()
    assert torch.equal(ring, ref), "push and pull forms disagree"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steady", type=int, default=0)
    args = ap.parse_args()
    img = Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), 224 * 224)
    if args.steady:
        gather_case(img, ROW, 256, 8192, N.VARIANT_TMA, args.steady)
        print("steady ok")
        return
    gather_case(img, ROW, 256, 4096, N.VARIANT_TMA, 3)                       # gather_tma_cvt_kernel
    gather_case(Convert.copy(), ROW, 256, 4096, N.VARIANT_TMA, 3)            # gather_tma_copy_kernel
    gather_case(img, ROW, 256, 4096, N.VARIANT_LDG, 2)                       # gather_ldg_kernel (u8->bf16)
    gather_case(Convert.copy(), 48, 4096, 1 << 16, N.VARIANT_LDG, 2)         # gather_ldg_kernel (tiny rows)
    gather_case(Convert.widen_tokens(), 4096, 4096, 1 << 15, N.VARIANT_TMA, 2)  # tokens i32 -> i64
    push_case()                                                              # push_tma_cvt_kernel
    n = 1 << 20
    t = torch.empty((3, n), dtype=torch.int64, device=DEV)
    slot_iota(t[0], 0)
    slot_iota(t[1], 1)
    dst = torch.empty(n, dtype=torch.int64, device=DEV)
    for _ in range(2):
        index_exchange(dst, t[0], t[1], t[1], n // 2)                        # index_exchange_kernel
    for m in (16384, 100520):
        tb = torch.arange(m, dtype=torch.int64, device=DEV).unsqueeze(0).contiguous()
        st = rng_states_to_tensor([Pcg64(7)], DEV)
        for _ in range(2):
            perm_gen(tb, st)                                                 # perm_gen_kernel
    L, s = N.lib(), C.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)
    flags = torch.zeros(64, dtype=torch.int32, device=DEV)
    targets = torch.tensor([flags.data_ptr() + 4 * i for i in range(8)], dtype=torch.int64, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    pages = torch.tensor([flags.data_ptr() + 128], dtype=torch.int64, device=DEV)
    for e in (1, 2):
        N.check(L.bddl_flag_signal(C.c_void_p(targets.data_ptr()), 8, e, 0, s))                       # flag_signal_kernel
        N.check(L.bddl_flag_wait(C.c_void_p(flags.data_ptr()), e, int(1e9), None, C.c_void_p(status.data_ptr()), s))
        N.check(L.bddl_flag_barrier(C.c_void_p(pages.data_ptr()), 1, 0, e, int(1e9), None, C.c_void_p(status.data_ptr()), s))
    batch = torch.randint(0, 255, (256, 2 * ROW), dtype=torch.uint8, device=DEV)
    ck = BatchChecksum(DEV, 256, 2 * ROW)
    for _ in range(2):
        ck.run(batch, 2 * ROW)                                               # checksum kernels
    torch.cuda.synchronize()
    print("targets ok")


if __name__ == "__main__":
    main()
