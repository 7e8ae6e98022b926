// N9 stage_host, CPU half: gather one batch's rows from pinned host shards into one contiguous
// pinned staging buffer with a small persistent thread pool, so that the host->device transfer
// is a single cudaMemcpyAsync at full PCIe rate instead of one small copy per sample.
// (The reference leaves host-to-device delivery to the user: tests/run_ddl.py:233-235.)
#include <string.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/ddl_b200.h"
#include "common.h"

namespace {

struct Job {
    char* dst;
    int64_t dst_row_bytes;
    const void* const* bases;
    int64_t row_bytes;
    const uint64_t* slots;
    int64_t n;
    int64_t copy_bytes;
};

class Pool {
public:
    explicit Pool(int n) : stop_(false), generation_(0), pending_(0) {
        for (int i = 0; i < n; ++i) workers_.emplace_back([this, i] { loop(i); });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> g(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    int size() const { return (int)workers_.size(); }
    void run(const Job& job) {
        std::unique_lock<std::mutex> g(m_);
        job_ = job;
        next_.store(0);
        pending_ = (int)workers_.size();
        ++generation_;
        cv_.notify_all();
        done_.wait(g, [this] { return pending_ == 0; });
    }

private:
    void loop(int) {
        uint64_t seen = 0;
        for (;;) {
            Job job;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [&] { return stop_ || generation_ != seen; });
                if (stop_) return;
                seen = generation_;
                job = job_;
            }
            const uint64_t mask = (1ULL << BDDL_SLOT_ROW_BITS) - 1;
            for (;;) {
                const int64_t i0 = next_.fetch_add(8);
                if (i0 >= job.n) break;
                const int64_t i1 = i0 + 8 < job.n ? i0 + 8 : job.n;
                for (int64_t i = i0; i < i1; ++i) {
                    const uint64_t s = job.slots[i];
                    const char* from = (const char*)job.bases[s >> BDDL_SLOT_ROW_BITS] + (int64_t)(s & mask) * job.row_bytes;
                    memcpy(job.dst + i * job.dst_row_bytes, from, (size_t)job.copy_bytes);
                }
            }
            {
                std::lock_guard<std::mutex> g(m_);
                if (--pending_ == 0) done_.notify_all();
            }
        }
    }
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    bool stop_;
    uint64_t generation_;
    int pending_;
    Job job_;
    std::atomic<int64_t> next_;
};

Pool* g_pool = nullptr;
std::mutex g_pool_mutex;

}  // namespace

extern "C" int bddl_host_gather_slots(void* h_dst, int64_t dst_row_bytes, const void* const* h_bases,
                                      int32_t n_sources, int64_t row_bytes, const uint64_t* slots, int64_t n,
                                      int64_t copy_bytes, int32_t n_threads) {
    if (!h_dst || !h_bases || (!slots && n > 0) || copy_bytes > row_bytes || copy_bytes > dst_row_bytes || n_sources < 1)
        return bddl_fail(BDDL_ERR_INVALID, "host_gather_slots: bad arguments");
    for (int64_t i = 0; i < n; ++i)
        if ((int64_t)(slots[i] >> BDDL_SLOT_ROW_BITS) >= n_sources)
            return bddl_fail(BDDL_ERR_INVALID, "host_gather_slots: slot %lld names source %llu of %d", (long long)i,
                             (unsigned long long)(slots[i] >> BDDL_SLOT_ROW_BITS), n_sources);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    std::lock_guard<std::mutex> g(g_pool_mutex);  // one batch at a time per process
    if (!g_pool || g_pool->size() != n_threads) {
        delete g_pool;
        g_pool = new Pool(n_threads);
    }
    Job job{(char*)h_dst, dst_row_bytes, h_bases, row_bytes, slots, n, copy_bytes};
    g_pool->run(job);
    return 0;
}

// Out-of-place row permutation for declared shuffles of host windows: dst[i] = X[perm[i]] where X is
// `src` with its first `ne` rows replaced by `recv_block` (the rows received in the global exchange;
// null = no exchange).  Threaded: one pass over the window instead of numpy's three row copies per
// Fisher-Yates step (reference tests/run_ddl.py:163-167).
extern "C" int bddl_host_permute_rows(void* dst, const void* src, const void* recv_block, int64_t ne,
                                      const int64_t* perm, int64_t n, int64_t row_bytes, int32_t n_threads) {
    if (!dst || !src || (!perm && n > 0) || n < 0 || row_bytes <= 0 || ne < 0 || ne > n || dst == src)
        return bddl_fail(BDDL_ERR_INVALID, "host_permute_rows: bad arguments");
    for (int64_t i = 0; i < n; ++i)
        if (perm[i] < 0 || perm[i] >= n) return bddl_fail(BDDL_ERR_INVALID, "host_permute_rows: perm[%lld] out of range", (long long)i);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; ++t) {
        th.emplace_back([=] {
            const int64_t i0 = n * t / n_threads, i1 = n * (t + 1) / n_threads;
            for (int64_t i = i0; i < i1; ++i) {
                const int64_t j = perm[i];
                const char* from = (recv_block && j < ne) ? (const char*)recv_block + j * row_bytes
                                                          : (const char*)src + j * row_bytes;
                memcpy((char*)dst + i * row_bytes, from, (size_t)row_bytes);
            }
        });
    }
    for (auto& x : th) x.join();
    return 0;
}
