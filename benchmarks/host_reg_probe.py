"""Why is zero-copy pull slower from cudaHostRegister'ed POSIX shm than from cudaHostAlloc memory?
Compares page backings: shm (4K), shm + MADV_HUGEPAGE, anonymous + MADV_HUGEPAGE, cudaHostAlloc."""
import ctypes as C, json, mmap, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200 import _native as N
from ddl_b200.ops import Convert, GatherPlan, SourceTable

for f in ("/sys/kernel/mm/transparent_hugepage/enabled", "/sys/kernel/mm/transparent_hugepage/shmem_enabled",
          "/proc/sys/vm/nr_hugepages"):
    try:
        print(f, open(f).read().strip())
    except Exception as e:
        print(f, "unreadable", e)
libc = C.CDLL("libc.so.6", use_errno=True)
dev = torch.device("cuda:0")
SB, rows, B = 150528, 4096, 256
nbytes = rows * SB
L = N.lib()


def pull_rate(dev_ptr):
    plan = GatherPlan(SourceTable([dev_ptr], dev), SB, SB, Convert.copy(), variant=2, tune=0)
    slots = torch.randperm(rows, device=dev)[: 8 * B].to(torch.int64).view(8, B).contiguous()
    out = plan.alloc_batch(B)
    plan.run(slots[0], out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(1, 8):
        plan.run(slots[i], out)
    b.record()
    torch.cuda.synchronize()
    return round(B * SB / (a.elapsed_time(b) / 7) / 1e6, 1)


def dma_rate(host_ptr):
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(L.bddl_memcpy_async(C.c_void_p(dst.data_ptr()), C.c_void_p(host_ptr), nbytes, 1, s))
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    N.check(L.bddl_memcpy_async(C.c_void_p(dst.data_ptr()), C.c_void_p(host_ptr), nbytes, 1, s))
    b.record()
    torch.cuda.synchronize()
    return round(nbytes / a.elapsed_time(b) / 1e6, 1)


def registered(ptr):
    d = C.c_void_p()
    N.check(L.bddl_host_register(C.c_void_p(ptr), nbytes, C.byref(d)))
    C.memset(ptr, 1, nbytes)
    return d.value

# (a) POSIX shm, default pages
p = C.c_void_p()
N.check(L.bddl_shm_create(b"/bddl_probe_a", nbytes, C.byref(p)))
C.memset(p.value, 1, nbytes)
d = registered(p.value)
print(json.dumps({"backing": "posix shm 4K + cudaHostRegister", "pull_gbs": pull_rate(d), "dma_gbs": dma_rate(p.value)}))
L.bddl_host_unregister(p); L.bddl_shm_detach(p, nbytes); L.bddl_shm_unlink(b"/bddl_probe_a")
# (b) POSIX shm + MADV_HUGEPAGE before first touch
p = C.c_void_p()
N.check(L.bddl_shm_create(b"/bddl_probe_b", nbytes, C.byref(p)))
rc = libc.madvise(p, C.c_size_t(nbytes), 14)
d = registered(p.value)
print(json.dumps({"backing": f"posix shm + MADV_HUGEPAGE(rc={rc}) + register", "pull_gbs": pull_rate(d), "dma_gbs": dma_rate(p.value)}))
L.bddl_host_unregister(p); L.bddl_shm_detach(p, nbytes); L.bddl_shm_unlink(b"/bddl_probe_b")
# (c) anonymous private + MADV_HUGEPAGE
m = mmap.mmap(-1, nbytes + (2 << 20))
addr = C.addressof(C.c_char.from_buffer(m))
addr = (addr + (2 << 20) - 1) & ~((2 << 20) - 1)
rc = libc.madvise(C.c_void_p(addr), C.c_size_t(nbytes), 14)
d = registered(addr)
print(json.dumps({"backing": f"anonymous + MADV_HUGEPAGE(rc={rc}) + register", "pull_gbs": pull_rate(d), "dma_gbs": dma_rate(addr)}))
L.bddl_host_unregister(C.c_void_p(addr))
# (d) cudaHostAlloc
h, dd = C.c_void_p(), C.c_void_p()
N.check(L.bddl_host_alloc(C.byref(h), C.byref(dd), nbytes))
C.memset(h.value, 1, nbytes)
print(json.dumps({"backing": "cudaHostAlloc", "pull_gbs": pull_rate(dd.value), "dma_gbs": dma_rate(h.value)}))
