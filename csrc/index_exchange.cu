// N4 index_exchange: the global shuffle on slot tables.
//
// The reference swaps row blocks between k-th producers with two blocking MPI calls
// (SendRecvReplaceGlobalShuffler.global_shuffle, reference ddl/shuffle.py:92-108):
//   rows [0, ne/2)  <- the same rows of the rank that sends to me   (recv_from)
//   rows [ne/2, ne) <- the same rows of the rank I send to          (send_to)
// Here rows never move; only their 8-byte slot ids do.  The kernel builds the next iteration's
// table from three previous-iteration tables, two of which may live on other GPUs and are read
// with plain loads through NVLink peer mappings (a pull, so no remote completion to wait for).
// Sample bytes cross the fabric later, once, inside exchange_gather_collate.
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"

namespace {

__global__ void __launch_bounds__(256) index_exchange_kernel(uint64_t* __restrict__ dst,
                                                             const uint64_t* __restrict__ own,
                                                             const uint64_t* __restrict__ from_recv,
                                                             const uint64_t* __restrict__ from_send, int64_t n,
                                                             int64_t ne) {
    const int64_t half = ne / 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t* src = i < half ? from_recv : (i < ne ? from_send : own);
        dst[i] = __ldcg(src + i);
    }
}

__global__ void __launch_bounds__(256) slot_iota_kernel(uint64_t* __restrict__ t, int64_t n, uint64_t hi,
                                                        int64_t first_row) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        t[i] = hi | (uint64_t)(first_row + i);
}

int blocks_for(int64_t n) {
    int64_t b = (n + 255) / 256;
    return (int)(b < 1 ? 1 : (b > 1184 ? 1184 : b));
}

}  // namespace

extern "C" int bddl_index_exchange(uint64_t* d_dst, const uint64_t* d_own, const uint64_t* d_from_recv,
                                   const uint64_t* d_from_send, int64_t n, int64_t num_exchange, void* stream) {
    if (!d_dst || !d_own || n < 0 || num_exchange < 0 || num_exchange > n)
        return bddl_fail(BDDL_ERR_INVALID, "index_exchange: bad arguments (n=%lld ne=%lld)", (long long)n,
                         (long long)num_exchange);
    if (num_exchange > 0 && (!d_from_recv || !d_from_send))
        return bddl_fail(BDDL_ERR_INVALID, "index_exchange: partner tables missing");
    if (d_dst == d_own) return bddl_fail(BDDL_ERR_INVALID, "index_exchange: dst must not alias own (double-buffer)");
    if (n == 0) return 0;
    index_exchange_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(d_dst, d_own, d_from_recv, d_from_send,
                                                                            n, num_exchange);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int bddl_slot_iota(uint64_t* d_table, int64_t n, int32_t source, int64_t first_row, void* stream) {
    if (!d_table || n < 0 || source < 0 || source >= (1 << 16) || first_row < 0)
        return bddl_fail(BDDL_ERR_INVALID, "slot_iota: bad arguments");
    if (n == 0) return 0;
    slot_iota_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(d_table, n,
                                                                       (uint64_t)source << BDDL_SLOT_ROW_BITS,
                                                                       first_row);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}
