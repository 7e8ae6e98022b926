// N2 perm_gen: bit-exact numpy Generator.shuffle on slot tables, on the GPU, in parallel.
//
// Replaces the reference's per-iteration in-place row shuffle (user execute_function:
// rng.shuffle(my_ary), reference tests/run_ddl.py:163-167, driven from ddl/datapusher.py:154).
// Rows never move here: the permutation is applied to a table of 64-bit slot ids.
//
// numpy's shuffle is "for i = n-1 .. 1: j = bounded_draw(i); swap(t[i], t[j])" with masked
// rejection on a stream of 32-bit draws -- sequential as written.  One CTA of
// 512 threads per table runs it in two parallel phases that give the same result bit for bit:
//
//   A. targets.  PCG64 is an LCG, so thread t jumps straight to output t of the current batch
//      (1024 draws per batch).  Whether draw k is accepted depends only on how many earlier draws
//      of the batch were (that fixes the index it is drawn for, hence the mask and the bound), so
//      accept/reject of the whole batch is the unique fixed point of "scan the accept bits,
//      re-evaluate" -- a few block-wide scans (2-6 rounds; more only in the last few thousand
//      entries).  Accepted draws are written as J[i] = j: the target of the swap for index i.
//      The batch in which the table ends leaves the generator exactly where numpy would (number
//      of outputs consumed, buffered upper half).
//   B. swaps.  Step i touches positions i and J[i]; steps commute unless they share a position.
//      Deterministic reservations: every pending step writes its priority (earlier step = larger
//      i) to both of its positions with a 64-bit atomicMax tagged by the round number; a step that
//      holds both positions has no pending predecessor and swaps; the rest retry.  The dependence
//      depth of a random Fisher-Yates is O(log n): 33 / 35 / 49 rounds and 3.5 n step attempts in
//      total for 16 k / 100 k / 1 M entries.
//
// Scratch (caller-owned, bddl_perm_gen_scratch_bytes): J, the reservation words and two work
// lists; everything else lives in registers and 200 bytes of shared memory, so the kernel
// co-resides with the batch-assembly kernel on one SM and runs one iteration ahead on a side
// stream.
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"
#include "rng_pcg64.h"

namespace {

constexpr int kThreads = 512;  // + <= 64 registers: co-resides with a 512-thread batch-assembly CTA on one SM
constexpr int kWarps = kThreads / 32;

struct Ctl {
    uint32_t warp_sum[kWarps];
    uint32_t end_thread;  // thread whose draw completed the table
    uint32_t end_half;    // 0: its low half was the last draw consumed, 1: its high half
    uint32_t count[2];    // work-list lengths (ping-pong)
};

// Exclusive scan of `x` over the CTA; *total = sum over all threads.  Two barriers.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t x, Ctl& c, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    __syncthreads();  // the previous scan's warp_sum has been read by everyone
    if (lane == 31) c.warp_sum[warp] = incl;
    __syncthreads();
    uint32_t w = lane < kWarps ? c.warp_sum[lane] : 0u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
    }
    *total = __shfl_sync(0xffffffffu, w, 31);
    const uint32_t before = warp ? __shfl_sync(0xffffffffu, w, warp - 1) : 0u;
    return before + incl - x;
}

__global__ void __launch_bounds__(kThreads) perm_gen_kernel(uint64_t* tables, int64_t table_stride, int64_t n,
                                                            BddlRng* states, uint8_t* scratch, int64_t scratch_stride) {
    __shared__ Ctl ctl;
    const int tid = threadIdx.x;
    uint64_t* const tbl = tables + (int64_t)blockIdx.x * table_stride;
    BddlRng* const gst = states + blockIdx.x;
    uint8_t* const my = scratch + (int64_t)blockIdx.x * scratch_stride;
    const int64_t n_pad = (n + 1) & ~1LL;
    unsigned long long* const R = reinterpret_cast<unsigned long long*>(my);  // [n] round << 32 | priority
    uint32_t* const J = reinterpret_cast<uint32_t*>(R + n);                   // [n] swap target of index i
    uint32_t* list[2] = {J + n_pad, J + 2 * n_pad};                           // pending steps (ping-pong)

    // CTA-uniform copy of the generator
    const U128 base = {gst->state_lo, gst->state_hi};
    const U128 inc = {gst->inc_lo, gst->inc_hi};
    uint32_t has32 = gst->has_uint32;
    uint32_t pending = gst->uinteger;
    uint64_t draws = gst->draws32;
    int64_t i = n - 1;

    for (int64_t x = tid; x < n; x += kThreads) R[x] = 0ULL;
    if (tid < 2) ctl.count[tid] = 0;

    // ---- phase A: J[i] for i = n-1 .. 1 ----------------------------------------------------
    if (has32 && i >= 1) {  // a buffered upper half from an earlier call is the first draw
        const uint64_t v = (uint64_t)pending & bddl_mask_for((uint64_t)i);
        has32 = 0;
        draws++;
        if (v <= (uint64_t)i) {
            if (tid == 0) J[i] = (uint32_t)v;
            --i;
        }
    }
    if (i >= 1) {
        U128 jm, jp, lm, lp;
        pcg_jump_coeffs(inc, (uint64_t)kThreads, &jm, &jp);    // one batch ahead
        pcg_jump_coeffs(inc, (uint64_t)(tid + 1), &lm, &lp);   // thread t owns output t+1 after `base`
        U128 st = u128_add(u128_mul(lm, base), lp);
        for (;;) {
            const uint64_t out = pcg_output(st);
            const uint32_t d0 = (uint32_t)out, d1 = (uint32_t)(out >> 32);  // numpy hands out the low half first
            uint32_t a0 = 0, a1 = 0, excl = 0, total = 0;
            uint64_t v0 = 0, v1 = 0;
            int64_t i0 = 0, i1 = 0;
            for (;;) {
                excl = block_excl_scan(a0 + a1, ctl, &total);
                i0 = i - (int64_t)excl;
                i1 = i0 - (int64_t)a0;
                v0 = i0 >= 1 ? ((uint64_t)d0 & bddl_mask_for((uint64_t)i0)) : 0;
                v1 = i1 >= 1 ? ((uint64_t)d1 & bddl_mask_for((uint64_t)i1)) : 0;
                const uint32_t n0 = i0 >= 1 && v0 <= (uint64_t)i0;
                const uint32_t n1 = i1 >= 1 && v1 <= (uint64_t)i1;
                const int changed = (n0 != a0) | (n1 != a1);
                a0 = n0;
                a1 = n1;
                if (!__syncthreads_or(changed)) break;  // fixed point: (a, idx, v) are numpy's
            }
            if (a0) J[i0] = (uint32_t)v0;
            if (a1) J[i1] = (uint32_t)v1;
            if ((int64_t)total >= i) {
                // the table ends inside this batch: the draw that was accepted for index 1 is the
                // last one consumed; leave the generator exactly where numpy would
                if (a0 && i0 == 1) {
                    ctl.end_thread = tid;
                    ctl.end_half = 0;
                } else if (a1 && i1 == 1) {
                    ctl.end_thread = tid;
                    ctl.end_half = 1;
                }
                __syncthreads();
                const uint32_t et = ctl.end_thread, eh = ctl.end_half;
                if (tid == (int)et) {
                    gst->state_lo = st.lo;
                    gst->state_hi = st.hi;
                    gst->has_uint32 = eh == 0 ? 1u : 0u;  // low half used, high half buffered
                    gst->uinteger = eh == 0 ? d1 : pending;
                    gst->draws32 = draws + 2ULL * et + eh + 1ULL;
                }
                i = 0;
                break;
            }
            i -= (int64_t)total;
            draws += 2ULL * kThreads;
            st = u128_add(u128_mul(jm, st), jp);
        }
    } else if (tid == 0) {
        // nothing drawn from fresh outputs: only the buffered half may have been used up
        gst->has_uint32 = has32;
        gst->draws32 = draws;
    }
    __syncthreads();  // J and the zeroed reservation words are complete

    // ---- phase B: apply the swaps --------------------------------------------------------------
    // pending steps: every i with J[i] != i (a step that swaps an entry with itself is a no-op)
    for (int64_t b = 1; b < n; b += kThreads) {
        const int64_t x = b + tid;
        const bool live = x < n && J[x] != (uint32_t)x;
        const unsigned m = __ballot_sync(0xffffffffu, live);
        uint32_t pos = 0;
        if ((tid & 31) == 0 && m) pos = atomicAdd(&ctl.count[0], (uint32_t)__popc(m));
        pos = __shfl_sync(0xffffffffu, pos, 0);
        if (live) list[0][pos + __popc(m & ((1u << (tid & 31)) - 1u))] = (uint32_t)x;
    }
    __syncthreads();
    int cur = 0;
    for (unsigned long long round = 1;; ++round) {
        const uint32_t cnt = ctl.count[cur];
        if (cnt == 0) break;
        const uint32_t* in = list[cur];
        uint32_t* outl = list[cur ^ 1];
        const unsigned long long tag = round << 32;
        if (tid == 0) ctl.count[cur ^ 1] = 0;
        for (uint32_t q = tid; q < cnt; q += kThreads) {  // reserve both positions
            const uint32_t x = __ldcg(in + q);
            const uint32_t j = __ldcg(J + x);
            atomicMax(R + x, tag | x);
            atomicMax(R + j, tag | x);
        }
        __syncthreads();
        for (uint32_t q0 = 0; q0 < cnt; q0 += kThreads) {  // commit or retry (warp-uniform trip count)
            const uint32_t q = q0 + tid;
            bool retry = false;
            uint32_t x = 0;
            if (q < cnt) {
                x = __ldcg(in + q);
                const uint32_t j = __ldcg(J + x);
                const unsigned long long want = tag | x;
                if (__ldcg(R + x) == want && __ldcg(R + j) == want) {
                    const uint64_t a = __ldcg(tbl + x), b = __ldcg(tbl + j);
                    __stcg(tbl + x, b);
                    __stcg(tbl + j, a);
                } else {
                    retry = true;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, retry);
            uint32_t pos = 0;
            if ((tid & 31) == 0 && m) pos = atomicAdd(&ctl.count[cur ^ 1], (uint32_t)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (retry) __stcg(outl + pos + __popc(m & ((1u << (tid & 31)) - 1u)), x);
        }
        __syncthreads();
        cur ^= 1;
    }
}

}  // namespace

extern "C" int64_t bddl_perm_gen_scratch_bytes(int64_t n) {
    if (n < 0) return 0;
    const int64_t n_pad = (n + 1) & ~1LL;
    return n * 8 + 3 * n_pad * 4 + 64;  // reservations, targets, two work lists
}

extern "C" int bddl_perm_gen(uint64_t* d_tables, int64_t table_stride, int64_t n, bddl_rng_state* d_states,
                             int32_t n_tables, void* d_scratch, int64_t scratch_bytes, void* stream) {
    if (!d_tables || !d_states || n < 0 || n_tables < 1 || table_stride < n)
        return bddl_fail(BDDL_ERR_INVALID, "perm_gen: bad arguments (n=%lld tables=%d stride=%lld)",
                         (long long)n, n_tables, (long long)table_stride);
    if (n >= (1LL << 32)) return bddl_fail(BDDL_ERR_UNSUPPORTED, "perm_gen: tables beyond 2^32 entries");
    const int64_t per_table = (bddl_perm_gen_scratch_bytes(n) + 63) & ~63LL;
    if (!d_scratch || ((uintptr_t)d_scratch & 15) || scratch_bytes < per_table * n_tables)
        return bddl_fail(BDDL_ERR_INVALID, "perm_gen: scratch of %lld bytes needed (16-byte aligned), got %lld",
                         (long long)(per_table * n_tables), (long long)scratch_bytes);
    perm_gen_kernel<<<n_tables, kThreads, 0, (cudaStream_t)stream>>>(d_tables, table_stride, n, (BddlRng*)d_states,
                                                                      (uint8_t*)d_scratch, per_table);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}
