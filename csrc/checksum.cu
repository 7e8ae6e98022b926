// Batch checksum: a 64-bit fingerprint of a delivered batch, computed on the device.
//
// The reference never checks what a consumer receives (its only test asserts exit code 0,
// reference tests/test_ddl.py:8-28).  Here a batch can be fingerprinted where it lies, so that
// tests and the integrity mode of the loader can compare "what arrived" with "what the owner
// holds" without copying the batch to the host, and so that a training loop has a cheap per-step
// result to read back.  The fingerprint is sum over 8-byte words of mix(word ^ position) modulo
// 2^64: position-dependent, yet independent of the order of summation, hence identical whatever
// the launch geometry.
//
// Two-stage reduction: every CTA folds one 64 KiB tile into scratch[1 + tile]; a ticket in
// scratch[0] elects the last CTA, which folds the partials into *d_out and re-arms the ticket.
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"

namespace bddl_checksum {

constexpr int kThreads = 256;
constexpr int64_t kTileBytes = 64 * 1024;

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33;
    return x;
}

// Rows of `payload_bytes` (multiple of 8) at stride `row_bytes`; word w of row r has position
// r * words_per_row + w.
__global__ void __launch_bounds__(kThreads) checksum_kernel(const uint8_t* __restrict__ batch, int64_t n_rows,
                                                            int64_t row_bytes, int64_t payload_bytes,
                                                            unsigned long long* scratch, unsigned long long* out) {
    __shared__ unsigned long long warp_sums[kThreads / 32];
    __shared__ bool is_last;
    const int64_t wpr = payload_bytes / 8;
    const int64_t total = n_rows * wpr;
    const int64_t tile_words = kTileBytes / 8;
    const int64_t w0 = (int64_t)blockIdx.x * tile_words;
    const int64_t w1 = w0 + tile_words < total ? w0 + tile_words : total;
    unsigned long long acc = 0;
    for (int64_t w = w0 + threadIdx.x; w < w1; w += kThreads) {
        const int64_t r = w / wpr, c = w - r * wpr;
        const unsigned long long v = *reinterpret_cast<const unsigned long long*>(batch + r * row_bytes + c * 8);
        acc += mix64(v ^ (0x9e3779b97f4a7c15ULL * (unsigned long long)(w + 1)));
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long s = 0;
        for (int i = 0; i < kThreads / 32; ++i) s += warp_sums[i];
        scratch[1 + blockIdx.x] = s;
        __threadfence();
        const unsigned long long ticket = atomicAdd(&scratch[0], 1ULL);
        is_last = ticket == (unsigned long long)gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        unsigned long long s = 0;
        for (int64_t i = threadIdx.x; i < (int64_t)gridDim.x; i += kThreads) s += __ldcg(&scratch[1 + i]);
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long t = 0;
            for (int i = 0; i < kThreads / 32; ++i) t += warp_sums[i];
            *out = t;
            scratch[0] = 0;  // re-arm the ticket for the next launch
        }
    }
}

int64_t tiles_for(int64_t n_rows, int64_t payload_bytes) {
    const int64_t total = n_rows * payload_bytes;
    return total <= 0 ? 0 : (total + kTileBytes - 1) / kTileBytes;
}

int check_args(const void* d_batch, int64_t n_rows, int64_t row_bytes, int64_t payload_bytes, const void* d_out) {
    if (!d_batch || !d_out || n_rows < 0 || payload_bytes <= 0 || (payload_bytes & 7) || (row_bytes & 7) ||
        row_bytes < payload_bytes || ((uintptr_t)d_batch & 7))
        return bddl_fail(BDDL_ERR_INVALID,
                         "batch_checksum: need 8-byte aligned rows (rows=%lld stride=%lld payload=%lld)",
                         (long long)n_rows, (long long)row_bytes, (long long)payload_bytes);
    return 0;
}

int launch(const void* d_batch, int64_t n_rows, int64_t row_bytes, int64_t payload_bytes, uint64_t* d_scratch,
           uint64_t* d_out, cudaStream_t st) {
    const int64_t tiles = tiles_for(n_rows, payload_bytes);
    if (tiles == 0) {
        BDDL_CUDA(cudaMemsetAsync(d_out, 0, sizeof(uint64_t), st));
        return 0;
    }
    // the ticket must start at zero even if an earlier launch was aborted half-way
    BDDL_CUDA(cudaMemsetAsync(d_scratch, 0, sizeof(uint64_t), st));
    checksum_kernel<<<(unsigned)tiles, kThreads, 0, st>>>((const uint8_t*)d_batch, n_rows, row_bytes, payload_bytes,
                                                          (unsigned long long*)d_scratch, (unsigned long long*)d_out);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace bddl_checksum

extern "C" int64_t bddl_batch_checksum_scratch_words(int64_t n_rows, int64_t payload_bytes) {
    return 1 + bddl_checksum::tiles_for(n_rows, payload_bytes);
}

extern "C" int bddl_batch_checksum(const void* d_batch, int64_t n_rows, int64_t row_bytes, int64_t payload_bytes,
                                   uint64_t* d_scratch, int64_t scratch_words, uint64_t* d_out, void* stream) {
    int rc = bddl_checksum::check_args(d_batch, n_rows, row_bytes, payload_bytes, d_out);
    if (rc) return rc;
    const int64_t need = 1 + bddl_checksum::tiles_for(n_rows, payload_bytes);
    if (!d_scratch || scratch_words < need)
        return bddl_fail(BDDL_ERR_INVALID, "batch_checksum: scratch has %lld words, %lld needed",
                         (long long)scratch_words, (long long)need);
    return bddl_checksum::launch(d_batch, n_rows, row_bytes, payload_bytes, d_scratch, d_out, (cudaStream_t)stream);
}

// ---- Entry point for callers that must not allocate ------------------------------------------------
//
// Some users of this library cannot call cudaMalloc when they want a fingerprint: a training
// process that has handed every free byte of HBM to its framework's caching allocator (a fresh
// cudaMalloc then fails, or forces that allocator to trim); a latency-critical step, because
// cudaMalloc and cudaFree may synchronise the whole device; a host application that forbids its
// plug-ins to allocate after start-up.  For them the scratch of the reduction lives in the
// library image: one file-scope device array of 1 + 8192 words (64 KiB).  Every GPU that loads
// the library holds that array until the process exits.
//
// Consequences of sharing one array:
//   * size limit: 8192 tiles of 64 KiB, i.e. batches up to 512 MiB; larger ones are refused;
//   * the array keeps what the previous call left in it -- the ticket word is cleared on the
//     stream before every launch, the partials are fully rewritten before they are read;
//   * two calls whose GPU work overlaps on one device (different streams, no ordering between
//     them) race on the array: order them;
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress); make that first call before any stream capture starts.
#define BDDL_CHECKSUM_STATIC_MAX_TILES 8192

static __device__ __align__(256) unsigned long long g_checksum_scratch[1 + BDDL_CHECKSUM_STATIC_MAX_TILES];

extern "C" int64_t bddl_batch_checksum_static_max_bytes(void) {
    return (int64_t)BDDL_CHECKSUM_STATIC_MAX_TILES * bddl_checksum::kTileBytes;
}

extern "C" int bddl_batch_checksum_static(const void* d_batch, int64_t n_rows, int64_t row_bytes,
                                          int64_t payload_bytes, uint64_t* d_out, void* stream) {
    int rc = bddl_checksum::check_args(d_batch, n_rows, row_bytes, payload_bytes, d_out);
    if (rc) return rc;
    const int64_t tiles = bddl_checksum::tiles_for(n_rows, payload_bytes);
    if (tiles > BDDL_CHECKSUM_STATIC_MAX_TILES)
        return bddl_fail(BDDL_ERR_INVALID,
                         "batch_checksum_static: batch of %lld bytes exceeds the built-in scratch (limit %lld bytes); "
                         "use bddl_batch_checksum with caller-owned scratch",
                         (long long)(n_rows * payload_bytes), (long long)bddl_batch_checksum_static_max_bytes());
    // In host code `g_checksum_scratch` is only a handle for the symbol, not a device address:
    // resolve it once per device, with that device current, and hand the result to the kernel.
    static unsigned long long* resolved[64];
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return bddl_fail(BDDL_ERR_INVALID, "batch_checksum_static: device index %d out of range", dev);
    unsigned long long* scratch = __atomic_load_n(&resolved[dev], __ATOMIC_ACQUIRE);
    if (!scratch) {
        void* p = nullptr;
        BDDL_CUDA(cudaGetSymbolAddress(&p, g_checksum_scratch));
        scratch = (unsigned long long*)p;
        __atomic_store_n(&resolved[dev], scratch, __ATOMIC_RELEASE);
    }
    return bddl_checksum::launch(d_batch, n_rows, row_bytes, payload_bytes, (uint64_t*)scratch, d_out,
                                 (cudaStream_t)stream);
}
