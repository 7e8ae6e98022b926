// N2 perm_gen: bit-exact numpy Generator.shuffle on slot tables, on the GPU, in parallel.
//
// Replaces the reference's per-iteration in-place row shuffle (user execute_function:
// rng.shuffle(my_ary), reference tests/run_ddl.py:163-167, driven from ddl/datapusher.py:154).
// Rows never move here: the permutation is applied to a table of 64-bit slot ids.
//
// numpy's shuffle is "for i = n-1 .. 1: j = bounded_draw(i); swap(t[i], t[j])" with masked
// rejection on a stream of 32-bit draws -- sequential as written.  One thread-block
// cluster (8 CTAs x 512 threads) per table runs it in two parallel phases that give the same
// result bit for bit:
//
//   A. targets (CTA 0 of the cluster; the others clear the reservation words meanwhile).  PCG64 is
//      an LCG, so thread t jumps straight to its two outputs of the current batch (2048 draws per
//      batch).  Whether draw k is accepted depends only on how many earlier draws of the batch
//      were (that fixes the index it is drawn for, hence the mask and the bound), so accept/reject
//      of the whole batch is the unique fixed point of "scan the accept bits, re-evaluate": one
//      block-wide scan and ONE barrier per round (double-buffered warp sums, the "nothing
//      changed" flag is read behind the next round's barrier), 3-6 rounds, more only in the last
//      few thousand entries.  Accepted draws are written as J[i] = j: the target of the swap for
//      index i.  The batch in which the table ends leaves the generator exactly where numpy would
//      (number of outputs consumed, buffered upper half).
//   B. swaps (the whole cluster).  Step i touches positions i and J[i]; steps commute unless they
//      share a position.  Deterministic reservations: every pending step writes its priority
//      (earlier step = larger i) to both of its positions with a 64-bit atomicMax tagged by the
//      round number; a step that holds both positions has no pending predecessor and swaps; the
//      rest retry.  The dependence depth of a random Fisher-Yates is O(log n): 33 / 35 / 49 rounds
//      and 3.5 n step attempts in total for 16 k / 100 k / 1 M entries; a round is two cluster
//      barriers and two levels of L2 latency.
//
// Scratch (caller-owned, bddl_perm_gen_scratch_bytes): J, the reservation words, two work lists
// and two counters; everything else lives in registers and 300 bytes of shared memory, so the
// CTAs co-reside with the batch-assembly kernel and the shuffle runs one iteration ahead on a
// high-priority side stream.
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"
#include "rng_pcg64.h"

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 512;  // + <= 64 registers: co-resides with a 512-thread batch-assembly CTA on one SM
constexpr int kWarps = kThreads / 32;
constexpr int kCluster = 8;
constexpr int kDraws = 4;  // per thread and batch: two 64-bit outputs

struct Ctl {
    uint32_t warp_sum[2][kWarps];
    uint32_t changed[2];
    uint32_t end_thread;  // thread whose draw completed the table
    uint32_t end_draw;    // which of its four draws
};

// Exclusive scan of `x` over the CTA through buffer `par` (callers alternate it); *total = sum over
// all threads.  One barrier.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t x, Ctl& c, int par, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) c.warp_sum[par][warp] = incl;
    __syncthreads();
    uint32_t w = lane < kWarps ? c.warp_sum[par][lane] : 0u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
    }
    *total = __shfl_sync(0xffffffffu, w, 31);
    const uint32_t before = warp ? __shfl_sync(0xffffffffu, w, warp - 1) : 0u;
    return before + incl - x;
}

// Phase A, one CTA: J[i] = swap target of index i for i = n-1 .. 1; advances *gst like numpy.
__device__ void draw_targets(Ctl& ctl, uint32_t* __restrict__ J, int64_t n, BddlRng* gst) {
    const int tid = threadIdx.x;
    const U128 base = {gst->state_lo, gst->state_hi};  // CTA-uniform copy of the generator
    const U128 inc = {gst->inc_lo, gst->inc_hi};
    uint32_t has32 = gst->has_uint32;
    const uint32_t pending = gst->uinteger;
    uint64_t draws = gst->draws32;
    int64_t i = n - 1;
    if (tid < 2) ctl.changed[tid] = 0;
    __syncthreads();  // every thread has read the generator before the finishing thread rewrites it

    if (has32 && i >= 1) {  // a buffered upper half from an earlier call is the first draw
        const uint64_t v = (uint64_t)pending & bddl_mask_for((uint64_t)i);
        has32 = 0;
        draws++;
        if (v <= (uint64_t)i) {
            if (tid == 0) J[i] = (uint32_t)v;
            --i;
        }
    }
    if (i < 1) {  // nothing drawn from fresh outputs: only the buffered half may have been used up
        if (tid == 0) {
            gst->has_uint32 = has32;
            gst->draws32 = draws;
        }
        return;
    }
    U128 jm, jp, lm, lp;
    pcg_jump_coeffs(inc, (uint64_t)(2 * kThreads), &jm, &jp);  // one batch ahead
    pcg_jump_coeffs(inc, (uint64_t)(2 * tid + 1), &lm, &lp);   // thread t owns outputs 2t+1 and 2t+2 after `base`
    U128 st0 = u128_add(u128_mul(lm, base), lp);
    U128 st1 = pcg_step(st0, inc);
    uint32_t rr = 0;  // scans so far: alternates the scan / flag buffers, also across batches
    for (;;) {
        const uint64_t o0 = pcg_output(st0), o1 = pcg_output(st1);
        // numpy hands out the low half of an output first
        const uint32_t d[kDraws] = {(uint32_t)o0, (uint32_t)(o0 >> 32), (uint32_t)o1, (uint32_t)(o1 >> 32)};
        uint32_t a[kDraws] = {0, 0, 0, 0};  // accept bits
        int64_t idx[kDraws] = {0, 0, 0, 0};
        uint32_t v[kDraws] = {0, 0, 0, 0};
        uint32_t total = 0;
        for (uint32_t r = 0;; ++r, ++rr) {
            const uint32_t excl = block_excl_scan(a[0] + a[1] + a[2] + a[3], ctl, (int)(rr & 1u), &total);
            // did the previous evaluation change anything anywhere?  It stamped its flag with its
            // round number + 1 if so; a flag left from two rounds ago carries an older stamp.
            if (r > 0 && ctl.changed[(rr - 1u) & 1u] != rr)
                break;  // fixed point: (a, idx, v) are numpy's, `total` counts the accepted draws
            int64_t run = i - (int64_t)excl;
            uint32_t changed = 0;
#pragma unroll
            for (int k = 0; k < kDraws; ++k) {
                idx[k] = run;
                const bool live = run >= 1;
                const uint64_t m = live ? ((uint64_t)d[k] & bddl_mask_for((uint64_t)run)) : 0ULL;
                const uint32_t nk = live && m <= (uint64_t)run;
                v[k] = (uint32_t)m;
                changed |= nk ^ a[k];
                a[k] = nk;
                run -= (int64_t)nk;  // later draws of this thread already see the new bit
            }
            if (changed) ctl.changed[rr & 1u] = rr + 1u;
        }
#pragma unroll
        for (int k = 0; k < kDraws; ++k)
            if (a[k]) J[idx[k]] = v[k];
        if ((int64_t)total >= i) {
            // the table ends inside this batch: the draw accepted for index 1 is the last one
            // consumed; leave the generator exactly where numpy would
#pragma unroll
            for (int k = 0; k < kDraws; ++k)
                if (a[k] && idx[k] == 1) {
                    ctl.end_thread = tid;
                    ctl.end_draw = k;
                }
            __syncthreads();
            const uint32_t et = ctl.end_thread, ek = ctl.end_draw;
            if (tid == (int)et) {
                const U128 fin = ek < 2 ? st0 : st1;
                gst->state_lo = fin.lo;
                gst->state_hi = fin.hi;
                gst->has_uint32 = (ek & 1u) ? 0u : 1u;  // low half used last: its high half is buffered
                gst->uinteger = (ek & 1u) ? pending : d[ek + 1 < kDraws ? ek + 1 : ek];
                gst->draws32 = draws + (uint64_t)kDraws * et + ek + 1ULL;
            }
            return;
        }
        i -= (int64_t)total;
        draws += (uint64_t)kDraws * kThreads;
        st0 = u128_add(u128_mul(jm, st0), jp);
        st1 = u128_add(u128_mul(jm, st1), jp);
    }
}

struct Hdr {
    uint32_t count[2];  // lengths of the two work lists
};

__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kThreads)
    perm_gen_kernel(uint64_t* tables, int64_t table_stride, int64_t n, BddlRng* states, uint8_t* scratch,
                    int64_t scratch_stride) {
    __shared__ Ctl ctl;
    cg::cluster_group cluster = cg::this_cluster();
    const int tid = threadIdx.x;
    const int crank = (int)cluster.block_rank();
    const int table = blockIdx.x / kCluster;
    const uint32_t gtid = (uint32_t)crank * kThreads + tid;
    constexpr uint32_t gthreads = kCluster * kThreads;
    uint64_t* const tbl = tables + (int64_t)table * table_stride;
    uint8_t* const my = scratch + (int64_t)table * scratch_stride;
    Hdr* const hdr = reinterpret_cast<Hdr*>(my);
    unsigned long long* const R = reinterpret_cast<unsigned long long*>(my + 64);  // [n] round << 32 | priority
    unsigned long long* const list0 = R + n;                                       // pending steps: j << 32 | i
    unsigned long long* const list1 = list0 + n;
    uint32_t* const J = reinterpret_cast<uint32_t*>(list1 + n);                    // [n_pad] swap target of index i

    if (crank == 0) {
        draw_targets(ctl, J, n, states + table);
    } else {
        for (int64_t x = (int64_t)gtid - kThreads; x < n; x += gthreads - kThreads) R[x] = 0ULL;
        if (gtid == kThreads) hdr->count[0] = hdr->count[1] = 0;
    }
    if (kCluster == 1) {
        for (int64_t x = tid; x < n; x += kThreads) R[x] = 0ULL;
        if (tid == 0) hdr->count[0] = hdr->count[1] = 0;
    }
    __threadfence();
    cluster.sync();  // J, the cleared reservation words and the counters are visible to the cluster

    // ---- phase B: apply the swaps ----------------------------------------------------------------
    // pending steps: every i with J[i] != i (a step that swaps an entry with itself is a no-op)
    for (int64_t b = 1; b < n; b += gthreads) {
        const int64_t x = b + gtid;
        uint32_t j = 0;
        bool live = false;
        if (x < n) {
            j = __ldcg(J + x);
            live = j != (uint32_t)x;
        }
        const unsigned m = __ballot_sync(0xffffffffu, live);
        uint32_t pos = 0;
        if ((tid & 31) == 0 && m) pos = atomicAdd(&hdr->count[0], (uint32_t)__popc(m));
        pos = __shfl_sync(0xffffffffu, pos, 0);
        if (live) __stcg(list0 + pos + __popc(m & ((1u << (tid & 31)) - 1u)), ((unsigned long long)j << 32) | (uint32_t)x);
    }
    __threadfence();
    cluster.sync();
    int cur = 0;
    for (unsigned long long round = 1;; ++round) {
        const uint32_t cnt = __ldcg(&hdr->count[cur]);
        if (cnt == 0) break;
        const unsigned long long* in = cur ? list1 : list0;
        unsigned long long* outl = cur ? list0 : list1;
        const unsigned long long tag = round << 32;
        // the other counter was last read at the top of the previous round (two barriers ago)
        if (gtid == 0) hdr->count[cur ^ 1] = 0;
        for (uint32_t q = gtid; q < cnt; q += gthreads) {  // reserve both positions
            const unsigned long long e = __ldcg(in + q);
            const uint32_t x = (uint32_t)e, j = (uint32_t)(e >> 32);
            atomicMax(R + x, tag | x);
            atomicMax(R + j, tag | x);
        }
        __threadfence();
        cluster.sync();
        for (uint32_t q0 = 0; q0 < cnt; q0 += gthreads) {  // commit or retry (warp-uniform trip count)
            const uint32_t q = q0 + gtid;
            bool retry = false;
            unsigned long long e = 0;
            if (q < cnt) {
                e = __ldcg(in + q);
                const uint32_t x = (uint32_t)e, j = (uint32_t)(e >> 32);
                const unsigned long long want = tag | x;
                // the four loads are independent: one level of L2 latency for the whole decision
                const unsigned long long rx = __ldcg(R + x), rj = __ldcg(R + j);
                const uint64_t a = __ldcg(tbl + x), b = __ldcg(tbl + j);
                if (rx == want && rj == want) {
                    __stcg(tbl + x, b);
                    __stcg(tbl + j, a);
                } else {
                    retry = true;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, retry);
            uint32_t pos = 0;
            if ((tid & 31) == 0 && m) pos = atomicAdd(&hdr->count[cur ^ 1], (uint32_t)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (retry) __stcg(outl + pos + __popc(m & ((1u << (tid & 31)) - 1u)), e);
        }
        __threadfence();
        cluster.sync();
        cur ^= 1;
    }
}

}  // namespace

extern "C" int64_t bddl_perm_gen_scratch_bytes(int64_t n) {
    if (n < 0) return 0;
    const int64_t n_pad = (n + 1) & ~1LL;
    return 64 + n * 8 * 3 + n_pad * 4;  // counters, reservation words, two work lists, targets
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
    perm_gen_kernel<<<n_tables * kCluster, kThreads, 0, (cudaStream_t)stream>>>(
        d_tables, table_stride, n, (BddlRng*)d_states, (uint8_t*)d_scratch, per_table);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}
