// N2 perm_gen: bit-exact numpy Generator.shuffle on slot tables, on the GPU.
//
// Replaces the reference's per-iteration in-place row shuffle (user execute_function:
// rng.shuffle(my_ary), reference tests/run_ddl.py:163-167, driven from ddl/datapusher.py:154).
// Rows never move here: the permutation is applied to a table of 64-bit slot ids.
//
// Fisher-Yates with masked rejection is a sequential chain, so the design
// is one warp per table that walks the 32-bit draw stream 32 draws at a time:
//   * each lane owns one draw of the batch (PCG64 is an LCG: lane l keeps its own state and
//     jumps 16 outputs per batch, so the raw stream is produced in parallel);
//   * whether a draw is accepted depends only on how many earlier draws of the batch were, so
//     accept/reject of all 32 draws is a fixed point found with a few ballots (two in practice);
//   * the accepted swaps of a batch touch disjoint entries unless two draws collide
//     (__match_any_sync) or a target falls into the batch's own index window; then they are
//     applied as up to 32 parallel swaps, otherwise one at a time in order;
//   * the last 64 entries replay their draws one at a time (exact end-of-stream accounting).
// Table accesses go through L2 (ld/st .cg) so that lanes always see each other's swaps.  Once
// the running index drops below kSmemEntries the live prefix of the table is pulled into shared
// memory and finished there: small indices are where batches fall back to the sequential replay
// (the ambiguous band is a large share of the range), and a replay step then costs a
// shared-memory round trip instead of an L2 one.
// The kernel is latency-bound by construction and is meant to run one iteration ahead on a
// side stream, off the batch-delivery critical path.
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"
#include "rng_pcg64.h"

namespace {

constexpr int kSmemEntries = 24576;  // 192 KiB of slot ids

__device__ __forceinline__ U128 shfl_u128(U128 v, int src) {
    U128 r;
    r.lo = __shfl_sync(0xffffffffu, v.lo, src);
    r.hi = __shfl_sync(0xffffffffu, v.hi, src);
    return r;
}

__global__ void __launch_bounds__(32) perm_gen_kernel(uint64_t* tables, int64_t table_stride, int64_t n,
                                                      BddlRng* states) {
    extern __shared__ uint64_t smem_tbl[];
    const int lane = threadIdx.x;
    uint64_t* const gtbl = tables + (int64_t)blockIdx.x * table_stride;
    uint64_t* tbl = gtbl;
    bool in_smem = false;
    int64_t cached = 0;
    auto ld = [&](int64_t x) -> uint64_t { return in_smem ? tbl[x] : __ldcg(tbl + x); };
    auto st_ = [&](int64_t x, uint64_t v) {
        if (in_smem)
            tbl[x] = v;
        else
            __stcg(tbl + x, v);
    };
    BddlRng* gst = states + blockIdx.x;

    // Warp-uniform copy of the generator.
    U128 base = {gst->state_lo, gst->state_hi};
    const U128 inc = {gst->inc_lo, gst->inc_hi};
    uint32_t has32 = gst->has_uint32;
    uint32_t pending = gst->uinteger;
    uint64_t draws = gst->draws32;

    int64_t i = n - 1;

    // Prologue: a buffered upper half from an earlier call is the first draw of the stream.
    while (has32 && i >= 1) {
        uint64_t mask = bddl_mask_for((uint64_t)i);
        uint64_t v = (uint64_t)pending & mask;
        has32 = 0;
        draws++;
        if (v <= (uint64_t)i) {
            if (lane == 0) {
                uint64_t a = ld(i), b = ld((int64_t)v);
                st_(i, b);
                st_((int64_t)v, a);
            }
            __syncwarp();
            --i;
        }
    }

    if (i >= 1) {
        // 16-output jump, and lane l positioned (l/2 + 1) outputs after `base`.
        U128 j16m, j16p;
        pcg_jump_coeffs(inc, 16, &j16m, &j16p);
        U128 lm, lp;
        pcg_jump_coeffs(inc, (uint64_t)(lane / 2 + 1), &lm, &lp);
        U128 st = u128_add(u128_mul(lm, base), lp);

        while (i >= 1) {
            if (!in_smem && i < kSmemEntries) {
                cached = i + 1;  // entries [0, i] are all that can still change
                for (int64_t x = lane; x < cached; x += 32) smem_tbl[x] = __ldcg(gtbl + x);
                __syncwarp();
                tbl = smem_tbl;
                in_smem = true;
            }
            const uint64_t out = pcg_output(st);
            const uint32_t d = (lane & 1) ? (uint32_t)(out >> 32) : (uint32_t)out;
            if (i >= 64) {
                // All 32 draws of the batch are consumed.  Which of them are accepted depends only
                // on how many earlier lanes were accepted, so iterate that count to its (unique)
                // fixed point: lane t is exact after t+1 rounds, in practice two rounds suffice.
                const unsigned lt = (1u << lane) - 1u;
                unsigned acc = 0;
                int64_t idx;
                uint64_t v;
                for (;;) {
                    idx = i - (int64_t)__popc(acc & lt);
                    v = (uint64_t)d & (~0ULL >> __clzll((long long)idx));  // numpy's mask for this index
                    const unsigned now = __ballot_sync(0xffffffffu, v <= (uint64_t)idx);
                    if (now == acc) break;
                    acc = now;
                }
                const bool mine = (acc >> lane) & 1u;
                // The accepted swaps commute unless two of them share a target or a target lies
                // inside the batch's own index window.
                bool clash = false;
                if (mine)
                    clash = __match_any_sync(acc, v) != (1u << lane) || (v + 31 >= (uint64_t)i && v != (uint64_t)idx);
                if (!__any_sync(0xffffffffu, clash)) {
                    if (mine && v != (uint64_t)idx) {
                        const uint64_t a = ld(idx), b = ld((int64_t)v);
                        st_(idx, b);
                        st_((int64_t)v, a);
                    }
                    __syncwarp();
                } else {
                    for (unsigned rem = acc; rem; rem &= rem - 1) {  // in order, one swap at a time
                        const int t = __ffs(rem) - 1;
                        const int64_t it = __shfl_sync(0xffffffffu, idx, t);
                        const int64_t jt = (int64_t)__shfl_sync(0xffffffffu, v, t);
                        if (lane == 0) {
                            const uint64_t a = ld(it), b = ld(jt);
                            st_(it, b);
                            st_(jt, a);
                        }
                        __syncwarp();
                    }
                }
                i -= (int64_t)__popc(acc);
                draws += 32;
                base = shfl_u128(st, 31);  // state after 16 outputs
                st = u128_add(u128_mul(j16m, st), j16p);
                continue;
            }
            // Replay: consume this batch's draws in order with the exact sequential rule.
            int used = 0;
            for (int t = 0; t < 32 && i >= 1; ++t) {
                const uint32_t dt = __shfl_sync(0xffffffffu, d, t);
                const uint64_t m = bddl_mask_for((uint64_t)i);
                const uint64_t vt = (uint64_t)dt & m;
                used = t + 1;
                if (vt <= (uint64_t)i) {
                    if (lane == 0) {
                        const uint64_t a = ld(i), b = ld((int64_t)vt);
                        st_(i, b);
                        st_((int64_t)vt, a);
                    }
                    __syncwarp();
                    --i;
                }
            }
            draws += (uint64_t)used;
            if (used == 32) {
                base = shfl_u128(st, 31);
                st = u128_add(u128_mul(j16m, st), j16p);
            } else {
                // Table finished mid-batch: leave the generator exactly where numpy would.
                const int outs = (used + 1) / 2;  // 64-bit outputs consumed from this batch
                base = shfl_u128(st, 2 * (outs - 1));
                if (used & 1) {
                    has32 = 1;
                    pending = __shfl_sync(0xffffffffu, d, used);  // upper half of the last output
                }
            }
        }
    }

    if (in_smem) {
        __syncwarp();
        for (int64_t x = lane; x < cached; x += 32) gtbl[x] = smem_tbl[x];
    }
    if (lane == 0) {
        gst->state_lo = base.lo;
        gst->state_hi = base.hi;
        gst->has_uint32 = has32;
        gst->uinteger = pending;
        gst->draws32 = draws;
    }
}

}  // namespace

extern "C" int bddl_perm_gen(uint64_t* d_tables, int64_t table_stride, int64_t n, bddl_rng_state* d_states,
                             int32_t n_tables, void* stream) {
    if (!d_tables || !d_states || n < 0 || n_tables < 1 || table_stride < n)
        return bddl_fail(BDDL_ERR_INVALID, "perm_gen: bad arguments (n=%lld tables=%d stride=%lld)",
                         (long long)n, n_tables, (long long)table_stride);
    if (n >= (1LL << 32)) return bddl_fail(BDDL_ERR_UNSUPPORTED, "perm_gen: tables beyond 2^32 entries");
    constexpr size_t smem_max = (size_t)kSmemEntries * sizeof(uint64_t);
    // only as much shared memory as the table can use, so that small tables co-reside with other kernels
    const size_t smem = (size_t)(n < kSmemEntries ? (n > 0 ? n : 1) : kSmemEntries) * sizeof(uint64_t);
    static bool configured[64];
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        BDDL_CUDA(cudaFuncSetAttribute(perm_gen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
        configured[dev & 63] = true;
    }
    perm_gen_kernel<<<n_tables, 32, smem, (cudaStream_t)stream>>>(d_tables, table_stride, n, (BddlRng*)d_states);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}
