// N8 flags: ownership hand-over and shutdown without messages.
//
// The reference passes a window between producer and consumer with zero-byte MPI messages on
// tag 7 (Recv/Ssend, Irecv/Issend + Waitany, reference ddl/connection.py:153-182), fences it
// with Win.Sync (ddl/connection.py:61-63) and cancels every wait at shutdown with an Ibarrier
// on a duplicated communicator (ddl/connection.py:32-37,184-187).  On B200 the same protocol is
// a page of monotonically increasing 32-bit counters in the symmetric heap:
//   * signal  : a tiny kernel stores (st.release.sys) or adds (red.release.sys) to counters that
//               may live on other GPUs -- an NVLink posted write, ordered after all prior writes
//               of the stream by a system-scope fence;
//   * wait    : cuStreamWaitValue32 on the waiter's own memory -- no SM is spent spinning; a
//               bounded in-kernel acquire-poll exists for platforms without stream memory ops;
//   * shutdown: a stop word polled by every bounded spin.
#include <cuda.h>
#include <cuda_runtime.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"
#include "ptx.cuh"

namespace {

using namespace bddl;

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*WriteValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

int driver_fn(const char* name, void** fn) {
    cudaDriverEntryPointQueryResult q;
    BDDL_CUDA(cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess || !*fn)
        return bddl_fail(BDDL_ERR_UNSUPPORTED, "driver entry point %s unavailable", name);
    return 0;
}

__global__ void flag_signal_kernel(uint32_t* const* targets, int n, uint32_t value, int add) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* t = targets[i];
    if (!t) return;
    fence_sys();  // everything this stream wrote before is visible before the flag moves
    if (add)
        red_release_sys_add_u32(t, value);
    else
        st_release_sys_u32(t, value);
}

// status: 0 = satisfied, 1 = timed out, 2 = stopped.
__global__ void flag_wait_kernel(const uint32_t* flag, uint32_t value, unsigned long long timeout_ns,
                                 const uint32_t* stop, uint32_t* status) {
    const uint32_t st = poll_geq(flag, value, stop, timeout_ns);
    if (status) st_release_sys_u32(status, st);
}

// All-ranks barrier on the flag page: rank `me` writes `epoch` into slot `me` of every rank's
// page (peer stores), then waits until every slot of its own page has reached `epoch`.
__global__ void flag_barrier_kernel(uint32_t* const* pages, int n_ranks, int me, uint32_t epoch,
                                    unsigned long long timeout_ns, const uint32_t* stop, uint32_t* status) {
    const int t = threadIdx.x;
    if (t < n_ranks) {
        fence_sys();
        st_release_sys_u32(pages[t] + me, epoch);
    }
    __syncthreads();
    uint32_t st = 0;
    if (t < n_ranks) st = poll_geq(pages[me] + t, epoch, stop, timeout_ns);
    if (st && status) st_release_sys_u32(status, st | ((uint32_t)t << 8));  // may be mapped host memory: no atomics
}

}  // namespace

extern "C" {

int bddl_stream_wait_value32(void* stream, uint32_t* d_addr, uint32_t value) {
    static WaitValue32Fn fn = nullptr;
    if (!fn) {
        int rc = driver_fn("cuStreamWaitValue32", (void**)&fn);
        if (rc) return rc;
    }
    CUresult r = fn((CUstream)stream, (CUdeviceptr)d_addr, value, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) return bddl_fail((int)r, "cuStreamWaitValue32 failed with CUresult %d", (int)r);
    return 0;
}

int bddl_stream_write_value32(void* stream, uint32_t* d_addr, uint32_t value) {
    static WriteValue32Fn fn = nullptr;
    if (!fn) {
        int rc = driver_fn("cuStreamWriteValue32", (void**)&fn);
        if (rc) return rc;
    }
    CUresult r = fn((CUstream)stream, (CUdeviceptr)d_addr, value, CU_STREAM_WRITE_VALUE_DEFAULT);
    if (r != CUDA_SUCCESS) return bddl_fail((int)r, "cuStreamWriteValue32 failed with CUresult %d", (int)r);
    return 0;
}

int bddl_flag_signal(uint32_t* const* d_targets, int32_t n, uint32_t value, int32_t add, void* stream) {
    if (!d_targets || n < 1) return bddl_fail(BDDL_ERR_INVALID, "flag_signal: bad arguments");
    flag_signal_kernel<<<(n + 63) / 64, 64, 0, (cudaStream_t)stream>>>(d_targets, n, value, add);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

int bddl_flag_wait(const uint32_t* d_flag, uint32_t value, uint64_t timeout_ns, const uint32_t* d_stop,
                   uint32_t* d_status, void* stream) {
    if (!d_flag) return bddl_fail(BDDL_ERR_INVALID, "flag_wait: null flag");
    flag_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(d_flag, value, timeout_ns, d_stop, d_status);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

int bddl_flag_barrier(uint32_t* const* d_pages, int32_t n_ranks, int32_t me, uint32_t epoch, uint64_t timeout_ns,
                      const uint32_t* d_stop, uint32_t* d_status, void* stream) {
    if (!d_pages || n_ranks < 1 || n_ranks > 1024 || me < 0 || me >= n_ranks)
        return bddl_fail(BDDL_ERR_INVALID, "flag_barrier: bad arguments");
    int threads = ((n_ranks + 31) / 32) * 32;
    flag_barrier_kernel<<<1, threads, 0, (cudaStream_t)stream>>>(d_pages, n_ranks, me, epoch, timeout_ns,
                                                                 d_stop, d_status);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
