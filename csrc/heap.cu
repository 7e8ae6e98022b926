// N7 heap + N9 stage_host: device memory that peers can map, pinned host staging, and the
// small set of CUDA runtime calls the Python runtime needs without going through torch.
//
// The reference keeps every producer's data in an MPI-3 shared-memory window that all ranks of
// an instance map (Win.Allocate_shared / Shared_query, reference ddl/connection.py:88-139).
// The B200 equivalent is a symmetric heap in HBM: each process cudaMallocs its segment, exports
// a CUDA IPC handle, and every peer maps it, after which kernels address remote shards with
// plain pointers over NVLink.  Host-resident datasets live in pinned, device-mapped memory and
// are either pulled by the kernel directly or staged with one cudaMemcpyAsync per contiguous
// run on a side stream (the reference's window fill, tests/run_ddl.py:152-161).
#include <cuda_runtime.h>
#include <string.h>

#include <thread>
#include <vector>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"

extern "C" {

int bddl_device_count(int32_t* n) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) {
        cudaGetLastError();
        c = 0;
    }
    *n = c;
    return 0;
}

int bddl_set_device(int32_t dev) {
    BDDL_CUDA(cudaSetDevice(dev));
    return 0;
}

int bddl_get_device(int32_t* dev) {
    int d = 0;
    BDDL_CUDA(cudaGetDevice(&d));
    *dev = d;
    return 0;
}

int bddl_malloc(void** p, size_t bytes) {
    if (!p) return bddl_fail(BDDL_ERR_INVALID, "malloc: null out pointer");
    BDDL_CUDA(cudaMalloc(p, bytes ? bytes : 16));
    return 0;
}

int bddl_free(void* p) {
    if (p) BDDL_CUDA(cudaFree(p));
    return 0;
}

int bddl_memset_async(void* p, int value, size_t bytes, void* stream) {
    BDDL_CUDA(cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream));
    return 0;
}

int bddl_memcpy_async(void* dst, const void* src, size_t bytes, int32_t kind, void* stream) {
    // kind: 1 = host->device, 2 = device->host, 3 = device->device, 4 = default (UVA)
    cudaMemcpyKind k = kind == 1   ? cudaMemcpyHostToDevice
                       : kind == 2 ? cudaMemcpyDeviceToHost
                       : kind == 3 ? cudaMemcpyDeviceToDevice
                                   : cudaMemcpyDefault;
    BDDL_CUDA(cudaMemcpyAsync(dst, src, bytes, k, (cudaStream_t)stream));
    return 0;
}

// Strided rows in one call (K1 shard_load: a pinned window whose rows are padded to the shard's
// 16-byte row stride on the way into HBM).
int bddl_memcpy2d_async(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes,
                        size_t rows, int32_t kind, void* stream) {
    cudaMemcpyKind k = kind == 1   ? cudaMemcpyHostToDevice
                       : kind == 2 ? cudaMemcpyDeviceToHost
                       : kind == 3 ? cudaMemcpyDeviceToDevice
                                   : cudaMemcpyDefault;
    BDDL_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, rows, k, (cudaStream_t)stream));
    return 0;
}

int bddl_stream_synchronize(void* stream) {
    BDDL_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

int bddl_ipc_get_handle(void* p, uint8_t handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    BDDL_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle, &h, 64);
    return 0;
}

int bddl_ipc_open_handle(const uint8_t handle[64], void** p) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    BDDL_CUDA(cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

int bddl_ipc_close_handle(void* p) {
    if (p) BDDL_CUDA(cudaIpcCloseMemHandle(p));
    return 0;
}

int bddl_can_access_peer(int32_t dev, int32_t peer, int32_t* ok) {
    int v = 0;
    BDDL_CUDA(cudaDeviceCanAccessPeer(&v, dev, peer));
    *ok = v;
    return 0;
}

int bddl_enable_peer_access(int32_t peer) {
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    BDDL_CUDA(e);
    return 0;
}

// Pinned host memory, mapped into the device address space; *dev_ptr is the address kernels use.
int bddl_host_alloc(void** host_ptr, void** dev_ptr, size_t bytes) {
    if (!host_ptr) return bddl_fail(BDDL_ERR_INVALID, "host_alloc: null out pointer");
    BDDL_CUDA(cudaHostAlloc(host_ptr, bytes ? bytes : 16, cudaHostAllocMapped | cudaHostAllocPortable));
    if (dev_ptr) BDDL_CUDA(cudaHostGetDevicePointer(dev_ptr, *host_ptr, 0));
    return 0;
}

int bddl_host_free(void* host_ptr) {
    if (host_ptr) BDDL_CUDA(cudaFreeHost(host_ptr));
    return 0;
}

// Pin and map memory that already exists (a POSIX shared-memory window, a numpy array).
int bddl_host_register(void* host_ptr, size_t bytes, void** dev_ptr) {
    BDDL_CUDA(cudaHostRegister(host_ptr, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable));
    if (dev_ptr) BDDL_CUDA(cudaHostGetDevicePointer(dev_ptr, host_ptr, 0));
    return 0;
}

int bddl_host_unregister(void* host_ptr) {
    if (host_ptr) BDDL_CUDA(cudaHostUnregister(host_ptr));
    return 0;
}

// Raw events for the per-step hot path: the loader issues "wait for slot free, assemble, record
// done" as one native call instead of several framework calls (host time per step matters when a
// batch takes under 20 microseconds on the device).
int bddl_event_create(void** ev) {
    cudaEvent_t e;
    BDDL_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *ev = (void*)e;
    return 0;
}

int bddl_event_destroy(void* ev) {
    if (ev) BDDL_CUDA(cudaEventDestroy((cudaEvent_t)ev));
    return 0;
}

int bddl_event_record(void* ev, void* stream) {
    BDDL_CUDA(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)stream));
    return 0;
}

int bddl_stream_wait_event(void* stream, void* ev) {
    BDDL_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)ev, 0));
    return 0;
}

int bddl_event_synchronize(void* ev) {
    BDDL_CUDA(cudaEventSynchronize((cudaEvent_t)ev));
    return 0;
}

// N9 stage_host: copy n_rows rows (indices rows[], stride row_bytes in the pinned source) into
// consecutive rows of a device staging buffer, merging runs of adjacent source rows into one
// cudaMemcpyAsync each.  Returns the number of copies issued through *n_copies.
int bddl_stage_rows(void* d_dst, int64_t dst_row_bytes, const void* h_src, int64_t row_bytes,
                    const int64_t* rows, int64_t n_rows, int64_t copy_bytes, void* stream, int64_t* n_copies) {
    if (!d_dst || !h_src || (!rows && n_rows > 0) || copy_bytes > row_bytes || copy_bytes > dst_row_bytes)
        return bddl_fail(BDDL_ERR_INVALID, "stage_rows: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    int64_t copies = 0;
    const bool dense = copy_bytes == row_bytes && row_bytes == dst_row_bytes;
    int64_t i = 0;
    while (i < n_rows) {
        int64_t j = i + 1;
        if (dense)
            while (j < n_rows && rows[j] == rows[j - 1] + 1) ++j;
        if (dense) {
            BDDL_CUDA(cudaMemcpyAsync((char*)d_dst + i * dst_row_bytes, (const char*)h_src + rows[i] * row_bytes,
                                      (size_t)((j - i) * row_bytes), cudaMemcpyHostToDevice, st));
        } else {
            BDDL_CUDA(cudaMemcpyAsync((char*)d_dst + i * dst_row_bytes, (const char*)h_src + rows[i] * row_bytes,
                                      (size_t)copy_bytes, cudaMemcpyHostToDevice, st));
        }
        ++copies;
        i = j;
    }
    if (n_copies) *n_copies = copies;
    return 0;
}

// Same for a batch described by slot ids (source << 48 | row) over several pinned shards:
// h_bases[source] is the host address of that shard in this process.
int bddl_stage_slots(void* d_dst, int64_t dst_row_bytes, const void* const* h_bases, int32_t n_sources,
                     int64_t row_bytes, const uint64_t* slots, int64_t n, int64_t copy_bytes, void* stream,
                     int64_t* n_copies) {
    if (!d_dst || !h_bases || (!slots && n > 0) || copy_bytes > row_bytes || copy_bytes > dst_row_bytes)
        return bddl_fail(BDDL_ERR_INVALID, "stage_slots: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const uint64_t row_mask = (1ULL << BDDL_SLOT_ROW_BITS) - 1;
    const bool dense = copy_bytes == row_bytes && row_bytes == dst_row_bytes;
    int64_t copies = 0, i = 0;
    while (i < n) {
        const uint64_t src = slots[i] >> BDDL_SLOT_ROW_BITS;
        if ((int64_t)src >= n_sources) return bddl_fail(BDDL_ERR_INVALID, "stage_slots: source %llu out of range",
                                                        (unsigned long long)src);
        int64_t j = i + 1;
        if (dense)
            while (j < n && slots[j] == slots[j - 1] + 1) ++j;
        const char* from = (const char*)h_bases[src] + (int64_t)(slots[i] & row_mask) * row_bytes;
        BDDL_CUDA(cudaMemcpyAsync((char*)d_dst + i * dst_row_bytes, from,
                                  (size_t)(dense ? (j - i) * row_bytes : copy_bytes), cudaMemcpyHostToDevice, st));
        ++copies;
        i = j;
    }
    if (n_copies) *n_copies = copies;
    return 0;
}

// Multi-threaded form: the per-copy driver cost (several microseconds) is what limits a single
// issuing thread, so n_threads host threads issue the row copies on private streams that are
// forked from and joined back into `stream` with events.  PCIe then runs near its DMA rate.
int bddl_stage_slots_mt(void* d_dst, int64_t dst_row_bytes, const void* const* h_bases, int32_t n_sources,
                        int64_t row_bytes, const uint64_t* slots, int64_t n, int64_t copy_bytes, void* stream,
                        int32_t n_threads) {
    if (!d_dst || !h_bases || (!slots && n > 0) || copy_bytes > row_bytes || copy_bytes > dst_row_bytes)
        return bddl_fail(BDDL_ERR_INVALID, "stage_slots_mt: bad arguments");
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 16) n_threads = 16;
    struct Lanes {
        int dev = -1;
        std::vector<cudaStream_t> streams;
        std::vector<cudaEvent_t> done;
        cudaEvent_t fork = nullptr;
    };
    static thread_local Lanes lanes;
    int dev = 0;
    BDDL_CUDA(cudaGetDevice(&dev));
    if (lanes.dev != dev || (int)lanes.streams.size() < n_threads) {
        lanes.dev = dev;
        if (!lanes.fork) BDDL_CUDA(cudaEventCreateWithFlags(&lanes.fork, cudaEventDisableTiming));
        while ((int)lanes.streams.size() < n_threads) {
            cudaStream_t st;
            cudaEvent_t ev;
            BDDL_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            BDDL_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            lanes.streams.push_back(st);
            lanes.done.push_back(ev);
        }
    }
    const uint64_t row_mask = (1ULL << BDDL_SLOT_ROW_BITS) - 1;
    for (int64_t i = 0; i < n; ++i)
        if ((int64_t)(slots[i] >> BDDL_SLOT_ROW_BITS) >= n_sources)
            return bddl_fail(BDDL_ERR_INVALID, "stage_slots_mt: source out of range");
    cudaStream_t main = (cudaStream_t)stream;
    BDDL_CUDA(cudaEventRecord(lanes.fork, main));
    std::vector<int> rc((size_t)n_threads, 0);
    std::vector<std::thread> th;
    Lanes* const L = &lanes;  // thread_local: workers must use the caller's instance
    for (int t = 0; t < n_threads; ++t) {
        th.emplace_back([&, t] {
            cudaSetDevice(dev);
            cudaStream_t ws = L->streams[t];
            cudaError_t e = cudaStreamWaitEvent(ws, L->fork, 0);
            const int64_t i0 = n * t / n_threads, i1 = n * (t + 1) / n_threads;
            for (int64_t i = i0; i < i1 && e == cudaSuccess; ++i) {
                const char* from = (const char*)h_bases[slots[i] >> BDDL_SLOT_ROW_BITS] +
                                   (int64_t)(slots[i] & row_mask) * row_bytes;
                e = cudaMemcpyAsync((char*)d_dst + i * dst_row_bytes, from, (size_t)copy_bytes,
                                    cudaMemcpyHostToDevice, ws);
            }
            if (e == cudaSuccess) e = cudaEventRecord(L->done[t], ws);
            rc[t] = (int)e;
        });
    }
    for (auto& x : th) x.join();
    for (int t = 0; t < n_threads; ++t) {
        if (rc[t]) return bddl_fail(rc[t], "stage_slots_mt: worker %d failed: %s", t, cudaGetErrorString((cudaError_t)rc[t]));
        BDDL_CUDA(cudaStreamWaitEvent(main, lanes.done[t], 0));
    }
    return 0;
}

}  // extern "C"
