// N5 exchange_gather_collate, pull form: the hot kernel of the loader.
//
// One launch assembles one batch: for every slot id (source << 48 | row) of the consumer's
// window slice it pulls the sample from the owning shard -- local HBM, an NVLink peer mapping
// or mapped pinned host memory, all just addresses in d_src_base -- converts it
// (uint8 -> bf16 normalise, fp32 -> bf16, ...) and writes row i of the batch tensor.  This is
// the reference's batch extraction (ddl/mpi_dataloader.py:179-198), the data half of its
// global shuffle (ddl/shuffle.py:92-108: rows cross the fabric once, here) and the H2D copy it
// leaves to the user (tests/run_ddl.py:233-235), fused.
//
// Work item = (sample, chunk).  Every warp runs its own pipeline over a private ring of
// shared-memory stages, so there is no inter-warp synchronisation at all:
//   * TMA/copy   : lane 0 alone issues cp.async.bulk global->shared (mbarrier complete_tx) and
//                  cp.async.bulk shared->global; no thread ever touches the bytes.
//   * TMA/convert: lane 0 issues the bulk loads; the warp converts a landed stage with
//                  LDS -> ALU -> STG.128 and re-arms the stage.
//   * LDG        : no shared memory, ld.global.nc.v4 -> convert -> st.global.v4 (tiny rows,
//                  and the comparison baseline for the TMA forms).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"
#include "ptx.cuh"

namespace {

using namespace bddl;

constexpr int kWarps = 8;  // warps per CTA, each with a private stage ring
constexpr uint64_t kRowMask = (1ULL << BDDL_SLOT_ROW_BITS) - 1;

struct KParams {
    const void* const* src_base;
    const int64_t* const* src_offsets;
    const int32_t* const* src_lengths;
    const uint64_t* slots;
    uint8_t* dst;
    int64_t row_bytes;
    int64_t sample_bytes;
    int64_t dst_row_bytes;
    uint32_t chan_stride;
    int32_t n_channels;
    uint32_t n_items;
    uint32_t n_samples;
    uint32_t chunks_per_sample;
    uint32_t pad_bits;
    uint32_t col_off;               // byte offset of the delivered column slice inside a row
    float scale[BDDL_MAX_CHANNELS];
    float bias[BDDL_MAX_CHANNELS];
};

template <int MODE>
struct Cvt;
template <>
struct Cvt<BDDL_CVT_COPY> {
    static constexpr int IN_ES = 1, OUT_ES = 1, PER = 16;
};
template <>
struct Cvt<BDDL_CVT_U8_BF16> {
    static constexpr int IN_ES = 1, OUT_ES = 2, PER = 8;
};
template <>
struct Cvt<BDDL_CVT_F32_BF16> {
    static constexpr int IN_ES = 4, OUT_ES = 2, PER = 8;
};
template <>
struct Cvt<BDDL_CVT_U8_F32> {
    static constexpr int IN_ES = 1, OUT_ES = 4, PER = 4;
};
template <>
struct Cvt<BDDL_CVT_I32_I64> {
    static constexpr int IN_ES = 4, OUT_ES = 8, PER = 2;
};

struct Item {
    const uint8_t* src;
    uint8_t* dst;
    uint32_t in_bytes;      // payload bytes of this chunk (0 = nothing to load)
    uint32_t region_bytes;  // input-byte extent this item is responsible for (payload + padding)
    uint32_t elem0;         // index of the chunk's first element inside the sample
};

// Work item (chunk c of batch row s) for the sample named by `slot`; `dst` is the batch tensor the
// row belongs to (the consumer's own ring slot in the pull form, a peer's in the push form).
template <int MODE, int CHUNK>
__device__ __forceinline__ Item resolve_slot(const KParams& p, uint64_t slot, uint32_t s, uint32_t c, uint8_t* dst) {
    using C = Cvt<MODE>;
    const uint32_t source = (uint32_t)(slot >> BDDL_SLOT_ROW_BITS);
    const uint64_t row = slot & kRowMask;
    const uint8_t* base = (const uint8_t*)__ldg((const unsigned long long*)(p.src_base + source));
    int64_t off, len;
    if (p.src_offsets) {
        const int64_t* offs = (const int64_t*)__ldg((const unsigned long long*)(p.src_offsets + source));
        const int32_t* lens = (const int32_t*)__ldg((const unsigned long long*)(p.src_lengths + source));
        off = __ldg(offs + row);
        len = __ldg(lens + row);
    } else {
        off = (int64_t)row * p.row_bytes;
        len = p.sample_bytes;
    }
    const int64_t start = (int64_t)c * CHUNK;
    Item it;
    it.in_bytes = len > start ? (uint32_t)min((int64_t)CHUNK, len - start) : 0u;
    it.region_bytes = (uint32_t)min((int64_t)CHUNK, p.sample_bytes - start);
    it.src = base + off + p.col_off + start;
    it.elem0 = (uint32_t)(start / C::IN_ES);
    it.dst = dst + (int64_t)s * p.dst_row_bytes + (int64_t)it.elem0 * C::OUT_ES;
    return it;
}

template <int MODE, int CHUNK>
__device__ __forceinline__ Item resolve_item(const KParams& p, uint32_t item) {
    // chunk-major order: neighbouring warps work on the same chunk index of *different* samples,
    // so a batch of few, huge samples still spreads its loads over all owners' NVLink ports
    // instead of draining one owner at a time
    const uint32_t c = item / p.n_samples;
    const uint32_t s = item - c * p.n_samples;
    return resolve_slot<MODE, CHUNK>(p, __ldg(p.slots + s), s, c, p.dst);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// x*scale + bias with two roundings, so results match a plain fp32 PyTorch expression exactly.
__device__ __forceinline__ float affine(float x, float sc, float bi) { return __fadd_rn(__fmul_rn(x, sc), bi); }

// uint8 -> float without the conversion pipe: PRMT builds the float 2^23 + x, and
// fma(2^23 + x, s, -(2^23 * s)) is x*s rounded once, i.e. bit-identical to fmul(float(x), s)
// (2^23 * s is exact).  Three full-rate instructions per element: PRMT, FFMA, FADD.
__device__ __forceinline__ float u8_magic(uint32_t word, int k) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u + (uint32_t)k));
}
__device__ __forceinline__ float affine_u8(float magic, float sc, float noff, float bi) {
    return __fadd_rn(__fmaf_rn(magic, sc, noff), bi);
}

template <bool CHAN>
__device__ __forceinline__ int channel_of(const KParams& p, uint32_t e) {
    if (!CHAN) return 0;
    return (int)((e / p.chan_stride) % (uint32_t)p.n_channels);
}

// Per-chunk, warp-uniform channel context.  When a channel plane is at least one chunk long a
// chunk touches at most two channels, so a group only compares its offset with `boundary`;
// shorter planes fall back to a division per group.
struct ChanCtx {
    float sc0, bi0, no0;
    float sc1, bi1, no1;
    uint32_t boundary;  // chunk-relative element index where the second channel starts
    bool fast;
};

template <bool CHAN>
__device__ __forceinline__ ChanCtx make_ctx(const KParams& p, uint32_t elem0, uint32_t chunk_elems) {
    ChanCtx c;
    int c0 = 0, c1 = 0;
    c.boundary = 0xffffffffu;
    c.fast = true;
    if (CHAN) {
        const uint32_t q = elem0 / p.chan_stride;
        c0 = (int)(q % (uint32_t)p.n_channels);
        c1 = c0 + 1 == p.n_channels ? 0 : c0 + 1;
        c.boundary = p.chan_stride - (elem0 - q * p.chan_stride);
        c.fast = p.chan_stride >= chunk_elems;
    }
    c.sc0 = p.scale[c0], c.bi0 = p.bias[c0], c.no0 = __fmul_rn(-8388608.0f, c.sc0);
    c.sc1 = p.scale[c1], c.bi1 = p.bias[c1], c.no1 = __fmul_rn(-8388608.0f, c.sc1);
    return c;
}

// Source loads: G = global memory through the read-only path, otherwise shared memory.
template <bool G>
__device__ __forceinline__ uint4 ld16(const uint8_t* q) {
    if constexpr (G) return __ldg(reinterpret_cast<const uint4*>(q));
    return *reinterpret_cast<const uint4*>(q);
}
template <bool G>
__device__ __forceinline__ uint2 ld8(const uint8_t* q) {
    if constexpr (G) return __ldg(reinterpret_cast<const uint2*>(q));
    return *reinterpret_cast<const uint2*>(q);
}
template <bool G>
__device__ __forceinline__ uint32_t ld4(const uint8_t* q) {
    if constexpr (G) return __ldg(reinterpret_cast<const uint32_t*>(q));
    return *reinterpret_cast<const uint32_t*>(q);
}
template <bool G>
__device__ __forceinline__ uint8_t ld1(const uint8_t* q) {
    if constexpr (G) return __ldg(q);
    return *q;
}

// One output vector (16 bytes) from PER input elements starting at src.  e_rel / e_abs: index of
// the group's first element relative to the chunk / to the sample.
template <int MODE, bool CHAN, bool G>
__device__ __forceinline__ uint4 convert_group(const KParams& p, const ChanCtx& cx, const uint8_t* src,
                                               uint32_t e_rel, uint32_t e_abs) {
    using C = Cvt<MODE>;
    uint4 out;
    if constexpr (MODE == BDDL_CVT_COPY) {
        out = ld16<G>(src);
    } else if constexpr (MODE == BDDL_CVT_I32_I64) {
        const uint2 v = ld8<G>(src);
        const long long a = (int)v.x, b = (int)v.y;
        out.x = (uint32_t)a;
        out.y = (uint32_t)((unsigned long long)a >> 32);
        out.z = (uint32_t)b;
        out.w = (uint32_t)((unsigned long long)b >> 32);
    } else {
        float f[C::PER];
        if constexpr (C::IN_ES == 1) {
            if constexpr (C::PER == 8) {
                const uint2 v = ld8<G>(src);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    f[k] = u8_magic(v.x, k);
                    f[4 + k] = u8_magic(v.y, k);
                }
            } else {
                const uint32_t v = ld4<G>(src);
#pragma unroll
                for (int k = 0; k < 4; ++k) f[k] = u8_magic(v, k);
            }
        } else {
            const uint4 a = ld16<G>(src);
            const uint4 b = ld16<G>(src + 16);
            f[0] = __uint_as_float(a.x), f[1] = __uint_as_float(a.y);
            f[2] = __uint_as_float(a.z), f[3] = __uint_as_float(a.w);
            f[4] = __uint_as_float(b.x), f[5] = __uint_as_float(b.y);
            f[6] = __uint_as_float(b.z), f[7] = __uint_as_float(b.w);
        }
        auto apply = [](float x, float sc, float no, float bi) {
            if constexpr (C::IN_ES == 1) return affine_u8(x, sc, no, bi);
            else return affine(x, sc, bi);
        };
        if (!CHAN || (cx.fast && (e_rel + C::PER <= cx.boundary || e_rel >= cx.boundary))) {
            const bool second = CHAN && e_rel >= cx.boundary;
            const float sc = second ? cx.sc1 : cx.sc0, bi = second ? cx.bi1 : cx.bi0, no = second ? cx.no1 : cx.no0;
#pragma unroll
            for (int k = 0; k < C::PER; ++k) f[k] = apply(f[k], sc, no, bi);
        } else if (cx.fast) {
#pragma unroll
            for (int k = 0; k < C::PER; ++k) {
                const bool second = e_rel + k >= cx.boundary;
                f[k] = apply(f[k], second ? cx.sc1 : cx.sc0, second ? cx.no1 : cx.no0, second ? cx.bi1 : cx.bi0);
            }
        } else {
#pragma unroll
            for (int k = 0; k < C::PER; ++k) {
                const int c = channel_of<CHAN>(p, e_abs + k);
                const float sc = p.scale[c];
                f[k] = apply(f[k], sc, __fmul_rn(-8388608.0f, sc), p.bias[c]);
            }
        }
        if constexpr (C::OUT_ES == 2) {
            out.x = pack_bf16(f[0], f[1]);
            out.y = pack_bf16(f[2], f[3]);
            out.z = pack_bf16(f[4], f[5]);
            out.w = pack_bf16(f[6], f[7]);
        } else {
            out.x = __float_as_uint(f[0]);
            out.y = __float_as_uint(f[1]);
            out.z = __float_as_uint(f[2]);
            out.w = __float_as_uint(f[3]);
        }
    }
    return out;
}

// One element, for the ragged tail of a chunk.
template <int MODE, bool CHAN, bool G>
__device__ __forceinline__ void convert_one(const KParams& p, const uint8_t* src, uint8_t* dst, uint32_t e) {
    if constexpr (MODE == BDDL_CVT_COPY) {
        *dst = ld1<G>(src);
    } else if constexpr (MODE == BDDL_CVT_I32_I64) {
        *reinterpret_cast<long long*>(dst) = (long long)(int)ld4<G>(src);
    } else {
        float x;
        if constexpr (Cvt<MODE>::IN_ES == 1)
            x = (float)ld1<G>(src);
        else
            x = __uint_as_float(ld4<G>(src));
        const int c = channel_of<CHAN>(p, e);
        x = affine(x, p.scale[c], p.bias[c]);
        if constexpr (Cvt<MODE>::OUT_ES == 2)
            *reinterpret_cast<__nv_bfloat16*>(dst) = __float2bfloat16_rn(x);
        else
            *reinterpret_cast<float*>(dst) = x;
    }
}

// One output vector from PER input elements that all share one channel: no boundary logic at all.
template <int MODE, bool G>
__device__ __forceinline__ uint4 convert_group_uniform(const uint8_t* src, float sc, float no, float bi) {
    using C = Cvt<MODE>;
    uint4 out;
    if constexpr (MODE == BDDL_CVT_COPY) {
        out = ld16<G>(src);
    } else if constexpr (MODE == BDDL_CVT_I32_I64) {
        const uint2 v = ld8<G>(src);
        const long long a = (int)v.x, b = (int)v.y;
        out.x = (uint32_t)a;
        out.y = (uint32_t)((unsigned long long)a >> 32);
        out.z = (uint32_t)b;
        out.w = (uint32_t)((unsigned long long)b >> 32);
    } else {
        float f[C::PER];
        if constexpr (C::IN_ES == 1) {
            if constexpr (C::PER == 8) {
                const uint2 v = ld8<G>(src);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    f[k] = affine_u8(u8_magic(v.x, k), sc, no, bi);
                    f[4 + k] = affine_u8(u8_magic(v.y, k), sc, no, bi);
                }
            } else {
                const uint32_t v = ld4<G>(src);
#pragma unroll
                for (int k = 0; k < 4; ++k) f[k] = affine_u8(u8_magic(v, k), sc, no, bi);
            }
        } else {
            const uint4 a = ld16<G>(src);
            const uint4 b = ld16<G>(src + 16);
            f[0] = affine(__uint_as_float(a.x), sc, bi), f[1] = affine(__uint_as_float(a.y), sc, bi);
            f[2] = affine(__uint_as_float(a.z), sc, bi), f[3] = affine(__uint_as_float(a.w), sc, bi);
            f[4] = affine(__uint_as_float(b.x), sc, bi), f[5] = affine(__uint_as_float(b.y), sc, bi);
            f[6] = affine(__uint_as_float(b.z), sc, bi), f[7] = affine(__uint_as_float(b.w), sc, bi);
        }
        if constexpr (C::OUT_ES == 2) {
            out.x = pack_bf16(f[0], f[1]);
            out.y = pack_bf16(f[2], f[3]);
            out.z = pack_bf16(f[4], f[5]);
            out.w = pack_bf16(f[6], f[7]);
        } else {
            out.x = __float_as_uint(f[0]);
            out.y = __float_as_uint(f[1]);
            out.z = __float_as_uint(f[2]);
            out.w = __float_as_uint(f[3]);
        }
    }
    return out;
}

// Groups [g0, g1) of a chunk, one channel: the hot loop (per 8 pixels of the uint8 -> bf16 form:
// one LDS.64, 8 PRMT, 8 FFMA, 8 FADD, 4 packs, one STG.128; 32-bit index arithmetic).
template <bool MC>
__device__ __forceinline__ void store16(uint8_t* dst, uint4 v) {
    if constexpr (MC)
        st_multimem_v4(dst, v);  // dst is a multicast mapping: one store, every GPU's copy
    else
        *reinterpret_cast<uint4*>(dst) = v;
}

template <int MODE, bool G, int UNROLL, bool MC>
__device__ __forceinline__ void convert_span(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t g0,
                                             uint32_t g1, int lane, float sc, float no, float bi) {
    using C = Cvt<MODE>;
#pragma unroll UNROLL
    for (uint32_t g = g0 + lane; g < g1; g += 32) {
        const uint4 v = convert_group_uniform<MODE, G>(src + g * (uint32_t)(C::PER * C::IN_ES), sc, no, bi);
        store16<MC>(dst + g * 16u, v);
    }
}

// Convert the payload of one chunk (src: shared or global) and pad the rest of its region.
template <int MODE, bool CHAN, bool G, int UNROLL, uint32_t CHUNK_ELEMS, bool MC = false>
__device__ __forceinline__ void convert_chunk(const KParams& p, const uint8_t* __restrict__ src,
                                              const Item& it, int lane) {
    using C = Cvt<MODE>;
    uint8_t* __restrict__ dst = it.dst;
    const uint32_t n_elems = it.in_bytes / C::IN_ES;
    const uint32_t n_groups = n_elems / C::PER;
    const ChanCtx cx = make_ctx<CHAN>(p, it.elem0, CHUNK_ELEMS);
    if (!CHAN || cx.fast) {
        // a chunk touches at most two channels: [0, gb) is the first, the group that straddles the
        // boundary (if any) goes through the per-element selects, the rest is the second channel
        const uint32_t gb = CHAN ? min(n_groups, cx.boundary / (uint32_t)C::PER) : n_groups;
        convert_span<MODE, G, UNROLL, MC>(src, dst, 0, gb, lane, cx.sc0, cx.no0, cx.bi0);
        if (CHAN && gb < n_groups) {
            uint32_t g2 = gb;
            if (cx.boundary % (uint32_t)C::PER) {
                if (lane == 0) {
                    const uint4 v = convert_group<MODE, CHAN, G>(p, cx, src + gb * (uint32_t)(C::PER * C::IN_ES),
                                                                 gb * C::PER, it.elem0 + gb * C::PER);
                    store16<MC>(dst + gb * 16u, v);
                }
                g2 = gb + 1;
            }
            convert_span<MODE, G, UNROLL, MC>(src, dst, g2, n_groups, lane, cx.sc1, cx.no1, cx.bi1);
        }
    } else {
#pragma unroll UNROLL
        for (uint32_t g = lane; g < n_groups; g += 32) {
            const uint4 v = convert_group<MODE, CHAN, G>(p, cx, src + (size_t)g * (C::PER * C::IN_ES), g * C::PER,
                                                         it.elem0 + g * C::PER);
            store16<MC>(dst + (size_t)g * 16, v);
        }
    }
    if constexpr (MC) return;  // multicast rows are whole 16-byte groups (checked by the launcher)
    const uint32_t done = n_groups * C::PER;
    for (uint32_t e = done + lane; e < n_elems; e += 32)
        convert_one<MODE, CHAN, G>(p, src + (size_t)e * C::IN_ES, dst + (size_t)e * C::OUT_ES, it.elem0 + e);
    // Variable-size samples: fill [payload, region) of the batch row with the pad element.
    const uint32_t region_elems = it.region_bytes / C::IN_ES;
    for (uint32_t e = n_elems + lane; e < region_elems; e += 32) {
        uint8_t* d = dst + (size_t)e * C::OUT_ES;
        if constexpr (C::OUT_ES == 1)
            *d = (uint8_t)p.pad_bits;
        else if constexpr (C::OUT_ES == 2)
            *reinterpret_cast<uint16_t*>(d) = (uint16_t)p.pad_bits;
        else if constexpr (C::OUT_ES == 4)
            *reinterpret_cast<uint32_t*>(d) = p.pad_bits;
        else
            *reinterpret_cast<long long*>(d) = (long long)(int32_t)p.pad_bits;
    }
}

// ---- LDG form --------------------------------------------------------------------------------
template <int MODE, bool CHAN, int CHUNK>
__global__ void __launch_bounds__(kWarps * 32) gather_ldg_kernel(const __grid_constant__ KParams p) {
    const int lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * kWarps + (threadIdx.x >> 5);
    const uint32_t nw = gridDim.x * kWarps;
    for (uint32_t item = gw; item < p.n_items; item += nw) {
        const Item it = resolve_item<MODE, CHUNK>(p, item);
        convert_chunk<MODE, CHAN, true, 4, CHUNK / Cvt<MODE>::IN_ES>(p, it.src, it, lane);
    }
}

// ---- TMA / convert form ------------------------------------------------------------------------
// One warp's pipeline over its private ring of STAGES chunks: lane 0 keeps STAGES bulk loads in
// flight, the warp converts a landed stage straight into the batch tensor and re-arms the stage.
// item_at(k) names the warp's k-th work item.
template <int MODE, bool CHAN, int CHUNK, int STAGES, bool MC = false, typename ItemAt>
__device__ __forceinline__ void warp_pipeline(const KParams& p, uint8_t* ring, uint64_t* full, uint32_t n_my,
                                              int lane, ItemAt item_at) {
    auto issue = [&](uint32_t k) {
        const Item it = item_at(k);
        if (it.in_bytes == 0) return;
        const int s = k % STAGES;
        const uint32_t bytes = (it.in_bytes + 15u) & ~15u;
        mbar_arrive_expect_tx(full + s, bytes);
        bulk_g2s(ring + (size_t)s * CHUNK, it.src, bytes, full + s);
    };

    if (lane == 0)
        for (uint32_t k = 0; k < min(n_my, (uint32_t)STAGES); ++k) issue(k);

    uint32_t phase = 0;  // one parity bit per stage
    for (uint32_t k = 0; k < n_my; ++k) {
        const int s = k % STAGES;
        const Item it = item_at(k);
        if (it.in_bytes != 0) {
            mbar_wait(full + s, (phase >> s) & 1u);
            phase ^= 1u << s;
        }
        convert_chunk<MODE, CHAN, false, 4, CHUNK / Cvt<MODE>::IN_ES, MC>(p, ring + (size_t)s * CHUNK, it, lane);
        __syncwarp();
        if (lane == 0 && k + STAGES < n_my) {
            fence_proxy_async_smem();  // the warp's reads of this stage precede the next bulk write
            issue(k + STAGES);
        }
    }
}

template <int MODE, bool CHAN, int CHUNK, int STAGES, int NW>
__global__ void __launch_bounds__(NW * 32) gather_tma_cvt_kernel(const __grid_constant__ KParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    uint8_t* ring = smem + (size_t)warp * STAGES * CHUNK;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NW * STAGES * CHUNK) + warp * STAGES;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
        fence_mbar_init();
    }
    __syncwarp();

    const uint32_t gw = blockIdx.x * NW + warp;
    const uint32_t nw = gridDim.x * NW;
    const uint32_t n_my = gw < p.n_items ? (p.n_items - gw + nw - 1) / nw : 0;
    warp_pipeline<MODE, CHAN, CHUNK, STAGES>(p, ring, full, n_my, lane,
                                             [&](uint32_t k) { return resolve_item<MODE, CHUNK>(p, gw + k * nw); });
}

// ---- push form ---------------------------------------------------------------------------------
// The "DataPusher" direction (reference ddl/datapusher.py:147-170 offers a window, the consumer
// reads it): here the OWNER of a sample delivers it.  One launch per loader step on every GPU:
// the grid is cut into one group of CTAs per consumer; a group
//   1. waits (bounded) until that consumer has released the ring slot (`free` word, written by the
//      consumer's training stream) and published the slot table of the window (`table` word),
//   2. copies the consumer's slice of its slot table into shared memory (one NVLink read) and
//      keeps the samples this GPU owns,
//   3. runs the same per-warp TMA pipeline as the pull form -- bulk loads from the LOCAL shard,
//      conversion, 16-byte stores into the consumer's ring slot through its peer mapping (posted
//      NVLink writes), and
//   4. the last CTA of the group to finish adds 1 to the consumer's `ready` word with
//      red.release.sys after a system-scope fence: the hand-over token of the reference
//      (ddl/connection.py:153-182) is written from inside the kernel that moved the data.
// CTA 0 finally waits (bounded) until all owners have bumped this GPU's own `ready` word, so the
// kernel's completion *is* "my batch is assembled" and the loader needs no second launch.
struct PushParams {
    const uint64_t* const* tables;  // [n_ranks] every consumer's slot-table buffer (peer mappings)
    uint8_t* const* rings;          // [n_ranks] every consumer's batch ring (peer mappings)
    uint32_t* const* flags;         // [n_ranks] every rank's flag page (peer mappings)
    uint8_t* mc_ring;               // multicast mapping of the rings (replicated windows), or null
    int64_t table_off;              // byte offset of the window slice inside a table buffer
    int64_t ring_off;               // byte offset of the ring slot inside a ring
    unsigned long long timeout_ns;
    uint32_t* status;               // outcome | which wait << 4 | peer << 8
    int32_t n_ranks, me, ctas_per_consumer;
    int32_t table_rank;             // replicated window: every consumer gets this rank's window; -1 = its own
    uint32_t owner_lo, owner_hi;    // source ids this GPU owns
    uint32_t ready_word, free_word0, table_word0, done_word0, stop_word;
    uint32_t free_need, table_need, ready_need, done_target;
};

// MC = true: a replicated window delivered through the NVLS multicast mapping of the rings -- the
// whole grid is one group, every sample this GPU owns is converted once and stored once
// (multimem.st), NVSwitch writes it into every consumer's ring slot, and the last CTA bumps every
// consumer's `ready` word.
template <int MODE, bool CHAN, int CHUNK, int STAGES, int NW, bool MC>
__global__ void __launch_bounds__(NW * 32) push_tma_cvt_kernel(const __grid_constant__ KParams p,
                                                               const __grid_constant__ PushParams q) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    uint8_t* ring = smem + (size_t)warp * STAGES * CHUNK;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NW * STAGES * CHUNK) + warp * STAGES;
    uint32_t* ctrl = reinterpret_cast<uint32_t*>(smem + (size_t)NW * STAGES * CHUNK + (size_t)NW * STAGES * 8);
    uint64_t* slot_of = reinterpret_cast<uint64_t*>(ctrl + 4);         // [n_samples] the consumer's slice
    uint16_t* mine = reinterpret_cast<uint16_t*>(slot_of + p.n_samples);  // batch rows this GPU owns

    const int G = MC ? (int)gridDim.x : q.ctas_per_consumer;
    const int group = MC ? 0 : blockIdx.x / G;
    const int lc = blockIdx.x - group * G;
    const int c = (q.me + group) % q.n_ranks;  // staggered: owners do not all start on one consumer
    const int tr = q.table_rank >= 0 ? q.table_rank : c;  // whose window this consumer receives
    uint32_t* my_page = q.flags[q.me];
    const uint32_t* stop = my_page + q.stop_word;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
        fence_mbar_init();
    }
    if (threadIdx.x == 0) ctrl[0] = 0;
    __syncthreads();
    // status = outcome (1 timed out, 2 stopped) | which wait (1 free, 2 table, 3 ready) << 4 | peer << 8
    if (threadIdx.x < (MC ? q.n_ranks : 1)) {
        const int peer = MC ? (int)threadIdx.x : c;  // multicast stores land in every consumer's slot
        uint32_t st = 0;
        if (q.free_need) st = poll_geq(my_page + q.free_word0 + peer, q.free_need, stop, q.timeout_ns);
        if (st) st |= (1u << 4) | ((uint32_t)peer << 8);
        if (!st && threadIdx.x == 0) {
            st = poll_geq(my_page + q.table_word0 + tr, q.table_need, stop, q.timeout_ns);
            if (st) st |= (2u << 4) | ((uint32_t)tr << 8);
        }
        if (st) atomicMax(ctrl, st);
    }
    __syncthreads();
    uint32_t st = ctrl[0];
    if (st == 0) {
        const uint64_t* theirs = reinterpret_cast<const uint64_t*>(reinterpret_cast<const uint8_t*>(q.tables[tr]) +
                                                                   q.table_off);
        for (uint32_t i = threadIdx.x; i < p.n_samples; i += NW * 32) slot_of[i] = __ldcg(theirs + i);
        __syncthreads();
        if (warp == 0) {  // ordered compaction: every CTA of the group builds the identical list
            uint32_t n = 0;
            for (uint32_t b = 0; b < p.n_samples; b += 32) {
                const uint32_t i = b + lane;
                bool own = false;
                if (i < p.n_samples) {
                    const uint32_t source = (uint32_t)(slot_of[i] >> BDDL_SLOT_ROW_BITS);
                    own = source >= q.owner_lo && source < q.owner_hi;
                }
                const unsigned m = __ballot_sync(0xffffffffu, own);
                if (own) mine[n + __popc(m & ((1u << lane) - 1u))] = (uint16_t)i;
                n += __popc(m);
            }
            if (lane == 0) ctrl[1] = n;
        }
        __syncthreads();
        const uint32_t n_mine = ctrl[1];
        const uint32_t n_items = n_mine * p.chunks_per_sample;
        const uint32_t gw = lc * NW + warp;
        const uint32_t nw = G * NW;
        const uint32_t n_my = gw < n_items ? (n_items - gw + nw - 1) / nw : 0;
        uint8_t* dst = (MC ? q.mc_ring : q.rings[c]) + q.ring_off;
        warp_pipeline<MODE, CHAN, CHUNK, STAGES, MC>(p, ring, full, n_my, lane, [&](uint32_t k) {
            const uint32_t item = gw + k * nw;
            const uint32_t ch = item / n_mine;
            const uint32_t s = mine[item - ch * n_mine];
            return resolve_slot<MODE, CHUNK>(p, slot_of[s], s, ch, dst);
        });
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();  // this CTA's stores into the consumers' rings are visible system-wide
            const uint32_t t = atomicAdd(my_page + q.done_word0 + (MC ? 0 : c), 1u) + 1u;
            if (t == q.done_target) {  // last CTA of the group: every CTA's stores precede the token
                __threadfence_system();
                if (MC) {
                    for (int r = 0; r < q.n_ranks; ++r) red_release_sys_add_u32(q.flags[r] + q.ready_word, 1u);
                } else {
                    red_release_sys_add_u32(q.flags[c] + q.ready_word, 1u);
                }
            }
        }
    }
    // status may live in mapped host memory (the loader reads it without synchronising): a plain
    // release store; any non-zero value says that a bounded wait gave up
    if (blockIdx.x == 0 && threadIdx.x == 0 && st == 0) {
        st = poll_geq(my_page + q.ready_word, q.ready_need, stop, q.timeout_ns);
        if (st) st |= (3u << 4) | ((uint32_t)q.me << 8);
    }
    if (st && threadIdx.x == 0) {
        // keep the FIRST failure: later launches only report "stopped"
        if (ld_acquire_sys_u32(q.status) == 0) st_release_sys_u32(q.status, st | (blockIdx.x << 16));
        // a wait gave up: set this rank's stop word so that the launches already queued behind this
        // one return at once instead of timing out one after the other
        st_release_sys_u32(my_page + q.stop_word, 1u);
    }
}

// ---- TMA / copy form ---------------------------------------------------------------------------
// Fixed-size rows, bytes unchanged: one thread per warp moves everything with bulk copies.
template <int CHUNK, int STAGES>
__global__ void __launch_bounds__(kWarps * 32) gather_tma_copy_kernel(const __grid_constant__ KParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    if (lane != 0) return;
    uint8_t* ring = smem + (size_t)warp * STAGES * CHUNK;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)kWarps * STAGES * CHUNK) + warp * STAGES;
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
    fence_mbar_init();

    const uint32_t gw = blockIdx.x * kWarps + warp;
    const uint32_t nw = gridDim.x * kWarps;
    const uint32_t n_my = gw < p.n_items ? (p.n_items - gw + nw - 1) / nw : 0;
    constexpr int LAG = 2;  // bulk stores allowed to still be reading their stage
    static_assert(STAGES > LAG, "ring too shallow");

    auto issue = [&](uint32_t k) {
        const Item it = resolve_item<BDDL_CVT_COPY, CHUNK>(p, gw + k * nw);
        const int s = k % STAGES;
        const uint32_t bytes = (it.in_bytes + 15u) & ~15u;
        mbar_arrive_expect_tx(full + s, bytes);
        bulk_g2s(ring + (size_t)s * CHUNK, it.src, bytes, full + s);
    };

    for (uint32_t k = 0; k < min(n_my, (uint32_t)STAGES); ++k) issue(k);
    for (uint32_t k = 0; k < n_my; ++k) {
        const int s = k % STAGES;
        const Item it = resolve_item<BDDL_CVT_COPY, CHUNK>(p, gw + k * nw);
        mbar_wait(full + s, (k / STAGES) & 1u);
        bulk_s2g(it.dst, ring + (size_t)s * CHUNK, (it.in_bytes + 15u) & ~15u);
        bulk_commit();
        bulk_wait_read<LAG>();  // store k-LAG has drained its stage
        if (k >= (uint32_t)LAG && k - LAG + STAGES < n_my) issue(k - LAG + STAGES);
    }
    bulk_wait<0>();
}

int g_sm_count[64];

int sm_count(int* out) {
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return bddl_fail(BDDL_ERR_INVALID, "device index %d out of range", dev);
    if (g_sm_count[dev] == 0) BDDL_CUDA(cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    *out = g_sm_count[dev];
    return 0;
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    BDDL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return 0;
}

template <int CHUNK>
uint32_t chunks_for(int64_t sample_bytes) {
    return (uint32_t)((sample_bytes + CHUNK - 1) / CHUNK);
}

int grid_for(const bddl_gather_desc* d, uint32_t n_items, int* grid, int warps = kWarps) {
    int sms = 0;
    int rc = sm_count(&sms);
    if (rc) return rc;
    int cap = d->max_ctas > 0 ? d->max_ctas : sms;
    int need = (int)((n_items + warps - 1) / warps);
    *grid = need < cap ? need : cap;
    if (*grid < 1) *grid = 1;
    return 0;
}

template <int MODE, bool CHAN>
int launch_ldg(const bddl_gather_desc* d, KParams& p, cudaStream_t st) {
    constexpr int CHUNK = 4096;
    p.chunks_per_sample = chunks_for<CHUNK>(d->sample_bytes);
    p.n_items = (uint32_t)d->n_samples * p.chunks_per_sample;
    int grid = 0;
    // The LDG form hides latency with occupancy: allow several CTAs per SM.
    bddl_gather_desc dd = *d;
    if (dd.max_ctas == 0) {
        int sms = 0;
        int rc = sm_count(&sms);
        if (rc) return rc;
        dd.max_ctas = sms * 4;
    }
    int rc = grid_for(&dd, p.n_items, &grid);
    if (rc) return rc;
    gather_ldg_kernel<MODE, CHAN, CHUNK><<<grid, kWarps * 32, 0, st>>>(p);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

template <int MODE, bool CHAN, int CHUNK, int STAGES, int NW>
int launch_tma_cvt(const bddl_gather_desc* d, KParams& p, cudaStream_t st) {
    constexpr size_t smem = (size_t)NW * STAGES * CHUNK + (size_t)NW * STAGES * sizeof(uint64_t);
    static_assert(smem <= 227 * 1024, "stage rings exceed shared memory");
    static bool configured[64];
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    auto kernel = gather_tma_cvt_kernel<MODE, CHAN, CHUNK, STAGES, NW>;
    if (!configured[dev & 63]) {
        int rc = set_smem(kernel, smem);
        if (rc) return rc;
        configured[dev & 63] = true;
    }
    p.chunks_per_sample = chunks_for<CHUNK>(d->sample_bytes);
    p.n_items = (uint32_t)d->n_samples * p.chunks_per_sample;
    int grid = 0;
    int rc = grid_for(d, p.n_items, &grid, NW);
    if (rc) return rc;
    kernel<<<grid, NW * 32, smem, st>>>(p);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

template <int CHUNK, int STAGES>
int launch_tma_copy(const bddl_gather_desc* d, KParams& p, cudaStream_t st) {
    constexpr size_t smem = (size_t)kWarps * STAGES * CHUNK + (size_t)kWarps * STAGES * sizeof(uint64_t);
    static bool configured[64];
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    auto kernel = gather_tma_copy_kernel<CHUNK, STAGES>;
    if (!configured[dev & 63]) {
        int rc = set_smem(kernel, smem);
        if (rc) return rc;
        configured[dev & 63] = true;
    }
    p.chunks_per_sample = chunks_for<CHUNK>(d->sample_bytes);
    p.n_items = (uint32_t)d->n_samples * p.chunks_per_sample;
    int grid = 0;
    int rc = grid_for(d, p.n_items, &grid);
    if (rc) return rc;
    kernel<<<grid, kWarps * 32, smem, st>>>(p);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

template <int MODE, bool CHAN>
int launch_mode(const bddl_gather_desc* d, KParams& p, bool tma, cudaStream_t st) {
    if (!tma) return launch_ldg<MODE, CHAN>(d, p, st);
    switch (d->tune) {
        case 1:
            return launch_tma_cvt<MODE, CHAN, 2048, 6, 16>(d, p, st);
        case 2:
            return launch_tma_cvt<MODE, CHAN, 8192, 3, 8>(d, p, st);
        case 3:
            return launch_tma_cvt<MODE, CHAN, 2048, 3, 32>(d, p, st);
        default:
            return launch_tma_cvt<MODE, CHAN, 4096, 3, 16>(d, p, st);
    }
}

template <int MODE, bool CHAN, bool MC>
int launch_push_as(const bddl_gather_desc* d, KParams& p, PushParams& q, cudaStream_t st) {
    constexpr int CHUNK = 4096, STAGES = 3, NW = 16;
    const size_t smem = (size_t)NW * STAGES * CHUNK + (size_t)NW * STAGES * sizeof(uint64_t) + 16 +
                        (size_t)d->n_samples * (sizeof(uint64_t) + sizeof(uint16_t));
    if (smem > 227 * 1024)
        return bddl_fail(BDDL_ERR_UNSUPPORTED, "push: %d samples per batch exceed the kernel's slot list (%d at most)",
                         d->n_samples, (int)((227 * 1024 - (size_t)NW * STAGES * (CHUNK + 8) - 16) / 10));
    static bool configured[64];
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    auto kernel = push_tma_cvt_kernel<MODE, CHAN, CHUNK, STAGES, NW, MC>;
    if (!configured[dev & 63]) {
        int rc = set_smem(kernel, 227 * 1024);
        if (rc) return rc;
        configured[dev & 63] = true;
    }
    p.chunks_per_sample = chunks_for<CHUNK>(d->sample_bytes);
    p.n_items = (uint32_t)d->n_samples * p.chunks_per_sample;
    kernel<<<q.n_ranks * q.ctas_per_consumer, NW * 32, smem, st>>>(p, q);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

template <int MODE, bool CHAN>
int launch_push(const bddl_gather_desc* d, KParams& p, PushParams& q, cudaStream_t st) {
    return q.mc_ring ? launch_push_as<MODE, CHAN, true>(d, p, q, st) : launch_push_as<MODE, CHAN, false>(d, p, q, st);
}

}  // namespace

static int64_t g_launches = 0;
void bddl_count_launch(int64_t n) { __atomic_fetch_add(&g_launches, n, __ATOMIC_RELAXED); }
extern "C" int64_t bddl_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// Validate a descriptor and translate it into kernel parameters (slots / dst are checked by the callers).
static int build_params(const bddl_gather_desc* d, KParams& p, bool& chan, bool& var) {
    if (!d->d_src_base || d->n_samples < 0 || d->n_sources < 1)
        return bddl_fail(BDDL_ERR_INVALID, "gather: null pointer or bad counts");
    if (d->mode < BDDL_CVT_COPY || d->mode > BDDL_CVT_I32_I64)
        return bddl_fail(BDDL_ERR_INVALID, "gather: unknown conversion mode %d", d->mode);
    if (d->sample_bytes <= 0 || d->sample_bytes >= (1LL << 31))
        return bddl_fail(BDDL_ERR_INVALID, "gather: sample_bytes %lld out of range", (long long)d->sample_bytes);
    if (d->dst_row_bytes & 15)
        return bddl_fail(BDDL_ERR_INVALID, "gather: batch rows must be 16-byte aligned (stride %lld)",
                         (long long)d->dst_row_bytes);
    var = d->d_src_offsets != nullptr;
    if (var && !d->d_src_lengths) return bddl_fail(BDDL_ERR_INVALID, "gather: offsets without lengths");
    if (d->col_offset_bytes < 0 || (d->col_offset_bytes & 15) || (var && d->col_offset_bytes))
        return bddl_fail(BDDL_ERR_INVALID, "gather: col_offset_bytes must be a non-negative multiple of 16 (fixed-size rows)");
    if (!var && ((d->row_bytes & 15) || d->row_bytes < d->sample_bytes + d->col_offset_bytes))
        return bddl_fail(BDDL_ERR_INVALID, "gather: shard row stride %lld must be a multiple of 16 and >= %lld",
                         (long long)d->row_bytes, (long long)d->sample_bytes);
    static const int in_es[] = {1, 1, 4, 1, 4}, out_es[] = {1, 2, 2, 4, 8};
    if (d->sample_bytes % in_es[d->mode])
        return bddl_fail(BDDL_ERR_INVALID, "gather: sample_bytes not a multiple of the element size");
    const int64_t out_bytes = d->sample_bytes / in_es[d->mode] * out_es[d->mode];
    if (d->dst_row_bytes < ((out_bytes + 15) & ~15LL))
        return bddl_fail(BDDL_ERR_INVALID, "gather: dst_row_bytes %lld too small for %lld output bytes",
                         (long long)d->dst_row_bytes, (long long)out_bytes);
    chan = d->n_channels > 1 && d->chan_stride > 0;
    if (d->n_channels > BDDL_MAX_CHANNELS) return bddl_fail(BDDL_ERR_INVALID, "gather: too many channels");

    p.src_base = d->d_src_base;
    p.src_offsets = d->d_src_offsets;
    p.src_lengths = d->d_src_lengths;
    p.slots = d->d_slots;
    p.dst = (uint8_t*)d->d_dst;
    p.row_bytes = d->row_bytes;
    p.sample_bytes = d->sample_bytes;
    p.dst_row_bytes = d->dst_row_bytes;
    p.chan_stride = chan ? (uint32_t)d->chan_stride : 1u;
    p.n_channels = chan ? d->n_channels : 1;
    p.pad_bits = d->pad_bits;
    p.n_samples = (uint32_t)d->n_samples;
    p.col_off = (uint32_t)d->col_offset_bytes;
    for (int i = 0; i < BDDL_MAX_CHANNELS; ++i) {
        p.scale[i] = d->scale[i];
        p.bias[i] = d->bias[i];
    }
    return 0;
}

extern "C" int bddl_exchange_gather_collate(const bddl_gather_desc* d, void* stream) {
    if (!d) return bddl_fail(BDDL_ERR_INVALID, "gather: null descriptor");
    if (d->n_samples == 0) return 0;
    if (!d->d_slots || !d->d_dst || ((uintptr_t)d->d_dst & 15))
        return bddl_fail(BDDL_ERR_INVALID, "gather: null slots/dst or unaligned dst");
    KParams p;
    bool chan, var;
    int rc = build_params(d, p, chan, var);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    bool tma;
    if (d->variant == BDDL_VARIANT_LDG)
        tma = false;
    else if (d->variant == BDDL_VARIANT_TMA)
        tma = true;
    else
        tma = d->sample_bytes >= 2048;

    switch (d->mode) {
        case BDDL_CVT_COPY:
            if (tma && !var) {
                switch (d->tune) {
                    case 1:
                        return launch_tma_copy<4096, 6>(d, p, st);
                    case 2:
                        return launch_tma_copy<2048, 12>(d, p, st);
                    default:
                        return launch_tma_copy<8192, 3>(d, p, st);
                }
            }
            return launch_mode<BDDL_CVT_COPY, false>(d, p, tma, st);
        case BDDL_CVT_U8_BF16:
            return chan ? launch_mode<BDDL_CVT_U8_BF16, true>(d, p, tma, st)
                        : launch_mode<BDDL_CVT_U8_BF16, false>(d, p, tma, st);
        case BDDL_CVT_F32_BF16:
            return chan ? launch_mode<BDDL_CVT_F32_BF16, true>(d, p, tma, st)
                        : launch_mode<BDDL_CVT_F32_BF16, false>(d, p, tma, st);
        case BDDL_CVT_U8_F32:
            return chan ? launch_mode<BDDL_CVT_U8_F32, true>(d, p, tma, st)
                        : launch_mode<BDDL_CVT_U8_F32, false>(d, p, tma, st);
        default:
            return launch_mode<BDDL_CVT_I32_I64, false>(d, p, tma, st);
    }
}

// One loader step in one call: [wait until the ring slot is free] -> assemble the batch ->
// [record that it is done].  Either event may be null.
extern "C" int bddl_gather_step(const bddl_gather_desc* d, void* stream, void* wait_event, void* record_event) {
    cudaStream_t st = (cudaStream_t)stream;
    if (wait_event) BDDL_CUDA(cudaStreamWaitEvent(st, (cudaEvent_t)wait_event, 0));
    int rc = bddl_exchange_gather_collate(d, stream);
    if (rc) return rc;
    if (record_event) BDDL_CUDA(cudaEventRecord((cudaEvent_t)record_event, st));
    return 0;
}

// CTAs per consumer group of the push kernel for `n_ranks` ranks and an SM budget (0 = all SMs).
extern "C" int bddl_push_ctas_per_consumer(int32_t n_ranks, int32_t max_ctas, int32_t* out) {
    if (n_ranks < 1 || !out) return bddl_fail(BDDL_ERR_INVALID, "push: bad arguments");
    int sms = 0;
    int rc = sm_count(&sms);
    if (rc) return rc;
    // Push CTAs take a whole SM each (200 KB of shared memory) and may wait inside the kernel for a
    // counter.  The small kernels that move those counters (free tokens, flag barrier, table
    // production) must always find an SM, also when every push CTA of this GPU is waiting -- with a
    // multicast launch all of them wait for the slowest consumer at once -- so eight SMs are never
    // given to push CTAs.
    const int cap = max_ctas > 0 ? max_ctas : (sms > 16 ? sms - 8 : sms);
    *out = cap / n_ranks > 0 ? cap / n_ranks : 1;
    return 0;
}

extern "C" int bddl_exchange_push_collate(const bddl_gather_desc* d, const bddl_push_desc* u, void* stream) {
    if (!d || !u) return bddl_fail(BDDL_ERR_INVALID, "push: null descriptor");
    if (d->n_samples <= 0 || d->n_samples > 65535) return bddl_fail(BDDL_ERR_INVALID, "push: 1..65535 samples per batch");
    if (!u->d_tables || !u->d_rings || !u->d_flags || !u->d_status || u->n_ranks < 1 ||
        u->n_ranks > BDDL_FLAG_MAX_RANKS || u->me < 0 || u->me >= u->n_ranks || u->ring_slot < 0 ||
        u->ring_slot >= BDDL_FLAG_MAX_SLOTS || u->window < 0 || u->window >= BDDL_FLAG_MAX_WINDOWS ||
        u->ctas_per_consumer < 1 || (u->ring_offset_bytes & 15) || (u->table_offset_bytes & 7))
        return bddl_fail(BDDL_ERR_INVALID, "push: bad descriptor");
    KParams p;
    bool chan, var;
    int rc = build_params(d, p, chan, var);
    if (rc) return rc;
    p.slots = nullptr;
    p.dst = nullptr;
    PushParams q;
    q.tables = (const uint64_t* const*)u->d_tables;
    q.rings = (uint8_t* const*)u->d_rings;
    q.flags = (uint32_t* const*)u->d_flags;
    q.mc_ring = (uint8_t*)u->d_mc_ring;
    q.table_rank = u->table_rank;
    if (u->table_rank >= u->n_ranks) return bddl_fail(BDDL_ERR_INVALID, "push: table_rank out of range");
    if (u->d_mc_ring) {
        static const int group_in[] = {16, 8, 32, 4, 8};  // input bytes of one 16-byte output group
        if (var || d->sample_bytes % group_in[d->mode] || ((uintptr_t)u->d_mc_ring & 15))
            return bddl_fail(BDDL_ERR_UNSUPPORTED, "push: multicast delivery needs fixed-size rows of whole 16-byte "
                                                   "output groups");
        if (u->table_rank < 0) return bddl_fail(BDDL_ERR_INVALID, "push: multicast delivery is for replicated windows");
    }
    q.table_off = u->table_offset_bytes;
    q.ring_off = u->ring_offset_bytes;
    q.timeout_ns = u->timeout_ns;
    q.status = u->d_status;
    q.n_ranks = u->n_ranks;
    q.me = u->me;
    q.ctas_per_consumer = u->ctas_per_consumer;
    q.owner_lo = (uint32_t)u->owner_lo;
    q.owner_hi = (uint32_t)u->owner_hi;
    q.ready_word = BDDL_FLAG_READY0 + u->ring_slot;
    q.free_word0 = BDDL_FLAG_FREE0 + u->ring_slot * BDDL_FLAG_MAX_RANKS;
    q.done_word0 = u->d_mc_ring ? BDDL_FLAG_DONE_MC0 + u->ring_slot : BDDL_FLAG_DONE0 + u->ring_slot * BDDL_FLAG_MAX_RANKS;
    q.table_word0 = BDDL_FLAG_TABLE0 + u->window * BDDL_FLAG_MAX_RANKS;
    q.stop_word = BDDL_FLAG_STOP;
    q.free_need = u->use;  // the slot's previous contents (its use-th fill) have been consumed
    q.table_need = u->table_seq;
    q.ready_need = (u->use + 1) * (uint32_t)u->n_ranks;  // every owner bumps `ready` once per fill
    // tickets are monotonic per (slot, group kind): `fills` earlier launches of this kind used the slot
    q.done_target = (u->fills + 1) * (uint32_t)(u->d_mc_ring ? u->n_ranks * u->ctas_per_consumer : u->ctas_per_consumer);
    cudaStream_t st = (cudaStream_t)stream;
    switch (d->mode) {
        case BDDL_CVT_COPY:
            return launch_push<BDDL_CVT_COPY, false>(d, p, q, st);
        case BDDL_CVT_U8_BF16:
            return chan ? launch_push<BDDL_CVT_U8_BF16, true>(d, p, q, st) : launch_push<BDDL_CVT_U8_BF16, false>(d, p, q, st);
        case BDDL_CVT_F32_BF16:
            return chan ? launch_push<BDDL_CVT_F32_BF16, true>(d, p, q, st) : launch_push<BDDL_CVT_F32_BF16, false>(d, p, q, st);
        case BDDL_CVT_U8_F32:
            return chan ? launch_push<BDDL_CVT_U8_F32, true>(d, p, q, st) : launch_push<BDDL_CVT_U8_F32, false>(d, p, q, st);
        default:
            return launch_push<BDDL_CVT_I32_I64, false>(d, p, q, st);
    }
}
