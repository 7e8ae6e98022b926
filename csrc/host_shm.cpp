// Host-side windows and flags for producer *processes* (the reference's own process model).
//
// Replaces MPI.Win.Allocate_shared / Shared_query (reference ddl/connection.py:109-131) with
// POSIX shared memory, and the zero-byte token messages + Waitany/Cancel
// (ddl/connection.py:153-182) with acquire/release counters in a shared flag page that are
// waited on with a bounded back-off and a stop word.
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "../include/ddl_b200.h"
#include "common.h"

static int map_shm(const char* name, size_t bytes, int create, void** out) {
    if (!name || !out || bytes == 0) return bddl_fail(BDDL_ERR_INVALID, "shm: bad arguments");
    int fd = shm_open(name, create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return bddl_fail(BDDL_ERR_INVALID, "shm_open(%s) failed: %s", name, strerror(errno));
    if (create && ftruncate(fd, (off_t)bytes) != 0) {
        int e = errno;
        close(fd);
        shm_unlink(name);
        return bddl_fail(BDDL_ERR_INVALID, "ftruncate(%s, %zu) failed: %s", name, bytes, strerror(e));
    }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    int e = errno;
    close(fd);
    if (p == MAP_FAILED) {
        if (create) shm_unlink(name);
        return bddl_fail(BDDL_ERR_INVALID, "mmap(%s, %zu) failed: %s", name, bytes, strerror(e));
    }
    *out = p;
    return 0;
}

static double now_us() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}

extern "C" {

int bddl_shm_create(const char* name, size_t bytes, void** ptr) { return map_shm(name, bytes, 1, ptr); }
int bddl_shm_attach(const char* name, size_t bytes, void** ptr) { return map_shm(name, bytes, 0, ptr); }

int bddl_shm_detach(void* ptr, size_t bytes) {
    if (ptr && munmap(ptr, bytes) != 0) return bddl_fail(BDDL_ERR_INVALID, "munmap failed: %s", strerror(errno));
    return 0;
}

int bddl_shm_unlink(const char* name) {
    if (shm_unlink(name) != 0 && errno != ENOENT)
        return bddl_fail(BDDL_ERR_INVALID, "shm_unlink(%s) failed: %s", name, strerror(errno));
    return 0;
}

void bddl_hflag_store(uint32_t* p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
uint32_t bddl_hflag_load(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
uint32_t bddl_hflag_add(uint32_t* p, uint32_t v) { return __atomic_add_fetch(p, v, __ATOMIC_ACQ_REL); }

// Wait until *p >= value (wrap-safe).  Returns 0, BDDL_ERR_STOPPED when *stop != 0, or
// BDDL_ERR_TIMEOUT after timeout_us (0 = wait forever).  Spins briefly, then yields, then sleeps.
int bddl_hflag_wait_geq(const uint32_t* p, uint32_t value, int64_t timeout_us, const uint32_t* stop) {
    if (!p) return bddl_fail(BDDL_ERR_INVALID, "hflag_wait: null flag");
    const double t0 = now_us();
    for (uint64_t it = 0;; ++it) {
        if ((int32_t)(__atomic_load_n(p, __ATOMIC_ACQUIRE) - value) >= 0) return 0;
        if (stop && __atomic_load_n(stop, __ATOMIC_ACQUIRE) != 0) return BDDL_ERR_STOPPED;
        if (it < 2000) continue;
        if (it < 4000) {
            sched_yield();
        } else {
            timespec ts = {0, 50000};
            nanosleep(&ts, nullptr);
        }
        if (timeout_us > 0 && (it & 63) == 0 && now_us() - t0 > (double)timeout_us)
            return bddl_fail(BDDL_ERR_TIMEOUT, "flag wait timed out after %lld us (have %u, want %u)",
                             (long long)timeout_us, __atomic_load_n(p, __ATOMIC_ACQUIRE), value);
    }
}

}  // extern "C"
