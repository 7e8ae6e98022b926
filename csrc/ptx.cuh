// Thin inline-PTX wrappers for the sm_100a features the loader uses: mbarrier, bulk async
// copies (TMA, SASS UBLKCP), proxy fences, system-scope release/acquire (NVLink-visible flags)
// and multimem stores.  No counterpart in the reference (pure Python over MPI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bddl {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

// Bounded: a bulk load whose bytes never arrive (a peer that went away) ends the kernel with a
// trap -- a loud launch failure on the host -- instead of spinning forever.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 26)) __trap();
    }
}

// global -> shared bulk copy, completion counted in bytes on an mbarrier (SASS UBLKCP.S.G).
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(__cvta_generic_to_global(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// shared -> global bulk copy tracked by the thread's bulk async-group (SASS UBLKCP.G.S).
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(__cvta_generic_to_global(gdst)),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

template <int N>
__device__ __forceinline__ void bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// Order generic-proxy shared-memory accesses against later async-proxy (TMA) accesses.
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(__cvta_generic_to_global(p)) : "memory");
    return v;
}

__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "r"(v) : "memory");
}

// 16-byte store through an NVLS multicast mapping: NVSwitch replicates it into every bound GPU's memory.
__device__ __forceinline__ void st_multimem_v4(void* mc_ptr, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(mc_ptr)),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Bounded acquire-poll of a counter that a peer (or the host) advances: 0 = reached `need`,
// 1 = `timeout_ns` passed (0 = no limit), 2 = the stop word was set.
__device__ __forceinline__ uint32_t poll_geq(const uint32_t* flag, uint32_t need, const uint32_t* stop,
                                             unsigned long long timeout_ns) {
    if ((int32_t)(ld_acquire_sys_u32(flag) - need) >= 0) return 0;
    const unsigned long long t0 = global_timer_ns();
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys_u32(flag) - need) < 0) {
        if ((++spins & 63u) == 0) {
            if (stop && ld_acquire_sys_u32(stop) != 0) return 2;
            if (timeout_ns && global_timer_ns() - t0 > timeout_ns) return 1;
        }
        __nanosleep(100);
    }
    return 0;
}

__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(__cvta_generic_to_global(p)));
    return v;
}

}  // namespace bddl
