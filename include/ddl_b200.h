/* ddl_b200 native library: C ABI.
 *
 * One shared object (ddl_b200/lib/libddl_b200.so), built for sm_100a only, loaded from Python
 * with ctypes.  Every function returns 0 on success; on failure it returns a non-zero code
 * (a cudaError_t value, or a negative BDDL_ERR_* code) and bddl_last_error() describes it.
 * Pointers named d_* are device addresses; `stream` is a cudaStream_t passed as void*.
 *
 * The reference (maximilian-tech/ddl) has no native code; each entry point cites the Python
 * it replaces (paths relative to the reference tree).
 */
#ifndef DDL_B200_H
#define DDL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BDDL_ABI_VERSION 2

#define BDDL_ERR_INVALID (-1)
#define BDDL_ERR_UNSUPPORTED (-2)
#define BDDL_ERR_TIMEOUT (-3)
#define BDDL_ERR_PARTNER (-4)
#define BDDL_ERR_STOPPED (-5)

const char* bddl_last_error(void);
int bddl_abi_version(void);

/* ---- N1: numpy-compatible RNG (replaces np.random.default_rng, ddl/shuffle.py:28-30 and
 *      tests/run_ddl.py:122). ------------------------------------------------------------- */
typedef struct bddl_rng_state {
    uint64_t state_lo, state_hi;
    uint64_t inc_lo, inc_hi;
    uint32_t has_uint32;
    uint32_t uinteger;
    uint64_t draws32;
} bddl_rng_state;

void bddl_rng_seed(bddl_rng_state* s, uint64_t seed);
uint32_t bddl_rng_next32(bddl_rng_state* s);
uint64_t bddl_rng_next64(bddl_rng_state* s);
uint64_t bddl_rng_interval(bddl_rng_state* s, uint64_t max_inclusive);
/* Skip n 64-bit outputs in O(log n). */
void bddl_rng_advance(bddl_rng_state* s, uint64_t n_outputs64);

/* Host Fisher-Yates, same draws and swaps as Generator.shuffle (tests/run_ddl.py:163-167).
 * _u64 permutes a slot table; _rows permutes n rows of row_bytes bytes in place. */
int bddl_host_shuffle_u64(bddl_rng_state* s, uint64_t* table, int64_t n);
int bddl_host_shuffle_i32(bddl_rng_state* s, int32_t* table, int64_t n);
int bddl_host_shuffle_rows(bddl_rng_state* s, void* rows, int64_t n, int64_t row_bytes);

/* ---- N3: partner selection (replaces GlobalShuffler._calculate_comm_partner,
 *      ddl/shuffle.py:32-79).  perm_out (optional, n_instance entries) receives the accepted
 *      send_to permutation; tries_out (optional) the number of rejected shuffles. ------------ */
int bddl_partner_select(bddl_rng_state* s, int32_t n_instance, int32_t my_rank, int32_t* send_to,
                        int32_t* recv_from, int32_t* perm_out, int32_t* tries_out);

/* ---- N2: perm_gen (replaces the in-place row shuffle rng.shuffle(my_ary), invoked at
 *      ddl/datapusher.py:154).  Shuffles n_tables slot tables of n entries each (table t starts
 *      at d_tables + t*table_stride) with the RNG states d_states[t], which are advanced in
 *      place on the device.  Bit-exact with bddl_host_shuffle_u64.  One thread-block cluster
 *      (8 CTAs x 512 threads) per table: block-parallel draw generation and accept/reject, then
 *      the swaps applied by the whole cluster in O(log n) rounds of deterministic reservations.  d_scratch: caller-owned device memory,
 *      n_tables * round_up(bddl_perm_gen_scratch_bytes(n), 64) bytes, 16-byte aligned. -------- */
int64_t bddl_perm_gen_scratch_bytes(int64_t n);
int bddl_perm_gen(uint64_t* d_tables, int64_t table_stride, int64_t n, bddl_rng_state* d_states,
                  int32_t n_tables, void* d_scratch, int64_t scratch_bytes, void* stream);

/* ---- N4: index_exchange (replaces the two Sendrecv_replace calls of
 *      SendRecvReplaceGlobalShuffler.global_shuffle, ddl/shuffle.py:92-108) on slot tables:
 *      dst[0,h) = from_recv[0,h); dst[h,ne) = from_send[h,ne); dst[ne,n) = own[ne,n), h = ne/2.
 *      from_recv / from_send may be peer-mapped (NVLink loads). ------------------------------ */
int bddl_index_exchange(uint64_t* d_dst, const uint64_t* d_own, const uint64_t* d_from_recv,
                        const uint64_t* d_from_send, int64_t n, int64_t num_exchange, void* stream);

/* Fill d_table[i] = (source << 48) | (first_row + i). */
int bddl_slot_iota(uint64_t* d_table, int64_t n, int32_t source, int64_t first_row, void* stream);

/* ---- N5: exchange_gather_collate (replaces batch extraction ddl/mpi_dataloader.py:179-198,
 *      the data half of ddl/shuffle.py:92-108 and the absent H2D copy tests/run_ddl.py:233-235).
 *      Pull form: for each of n_samples slot ids (source << 48 | row) read the sample from
 *      src_base[source] (local HBM, NVLink peer mapping or mapped pinned host memory), apply the
 *      conversion, and write row i of the batch tensor at dst + i*dst_row_bytes. ------------- */
enum {
    BDDL_CVT_COPY = 0,     /* bytes unchanged                                   */
    BDDL_CVT_U8_BF16 = 1,  /* out = bf16(float(u8) * scale[c] + bias[c])        */
    BDDL_CVT_F32_BF16 = 2, /* out = bf16(f32 * scale[c] + bias[c])              */
    BDDL_CVT_U8_F32 = 3,   /* out = float(u8) * scale[c] + bias[c]              */
    BDDL_CVT_I32_I64 = 4   /* sign-extend token ids                             */
};
enum { BDDL_VARIANT_AUTO = 0, BDDL_VARIANT_LDG = 1, BDDL_VARIANT_TMA = 2 };
#define BDDL_MAX_CHANNELS 8
#define BDDL_SLOT_ROW_BITS 48

typedef struct bddl_gather_desc {
    const void* const* d_src_base;   /* device array [n_sources] of shard base addresses          */
    const int64_t* const* d_src_offsets; /* device array [n_sources] of CSR byte offsets, or NULL */
    const int32_t* const* d_src_lengths; /* device array [n_sources] of payload bytes, or NULL    */
    const uint64_t* d_slots;         /* [n_samples] slot ids                                      */
    void* d_dst;                     /* batch tensor                                              */
    int64_t row_bytes;               /* fixed-size rows: shard row stride, multiple of 16         */
    int64_t sample_bytes;            /* payload bytes per sample (variable: the maximum)          */
    int64_t dst_row_bytes;           /* output row stride, multiple of 16                         */
    int64_t chan_stride;             /* elements per channel plane; 0 = one scale/bias            */
    int32_t n_samples;
    int32_t n_sources;
    int32_t mode;                    /* BDDL_CVT_*                                                */
    int32_t n_channels;
    float scale[BDDL_MAX_CHANNELS];
    float bias[BDDL_MAX_CHANNELS];
    uint32_t pad_bits;               /* variable-size: output element used for padding            */
    int32_t variant;                 /* BDDL_VARIANT_*                                            */
    int32_t max_ctas;                /* SM budget; 0 = one CTA per SM                             */
    int32_t reserved0;               /* (was the source filter of the two-launch push form)          */
    int32_t reserved1;
    int32_t col_offset_bytes;        /* deliver sample bytes [col_offset, col_offset + sample_bytes) (multiple of 16) */
    int32_t tune;                    /* stage ring per warp; convert: 0 = 16 warps x 3 x 4 KiB, 1 = 16 x 6 x 2 KiB, 2 = 8 x 3 x 8 KiB, 3 = 32 x 3 x 2 KiB */
} bddl_gather_desc;

int bddl_exchange_gather_collate(const bddl_gather_desc* desc, void* stream);
/* One loader step in one call: [cudaStreamWaitEvent(wait_event)] -> assemble -> [cudaEventRecord(record_event)]. */
int bddl_gather_step(const bddl_gather_desc* desc, void* stream, void* wait_event, void* record_event);

/* Push form of N5 (the "DataPusher" direction, ddl/datapusher.py:147-170, with the lend/return
 * tokens of ddl/connection.py:153-182 written from inside the kernel).  One launch per loader
 * step on every GPU: each GPU scans every consumer's slice of its slot table, delivers the samples
 * it owns into that consumer's ring slot with stores over NVLink, and the last CTA working for a
 * consumer adds 1 to that consumer's `ready` word (red.release.sys).  The kernel first waits,
 * bounded, for the consumer's `free` word of the ring slot and for its `table` word; CTA 0 finally
 * waits, bounded, until all n_ranks owners have bumped this GPU's own `ready` word, so the kernel's
 * completion means "my batch is assembled".  *d_status (device or mapped host memory) receives
 * 1 (timed out) or 2 (stopped) when a wait gives up.  `desc` is a gather descriptor whose d_slots and
 * d_dst are ignored.  Flag page layout (32-bit words, one page of BDDL_FLAG_PAGE_BYTES per rank): */
#define BDDL_FLAG_PAGE_BYTES 16384
#define BDDL_FLAG_MAX_RANKS 64
#define BDDL_FLAG_MAX_SLOTS 16
#define BDDL_FLAG_MAX_WINDOWS 24
#define BDDL_FLAG_BARRIER0 0      /* [rank]              flag barrier slots                          */
#define BDDL_FLAG_READY0 256      /* [slot]              += 1 by every owner per fill of my slot     */
#define BDDL_FLAG_STOP 272        /*                     non-zero cancels every wait                 */
#define BDDL_FLAG_DONE_MC0 288    /* [slot]              local CTA tickets of multicast launches     */
#define BDDL_FLAG_FREE0 512       /* [slot][consumer]    = fills consumed, stored by that consumer   */
#define BDDL_FLAG_DONE0 1536      /* [slot][consumer]    local CTA tickets of the push kernel        */
#define BDDL_FLAG_TABLE0 2560     /* [window][consumer]  = tables produced, stored by that consumer  */
typedef struct bddl_push_desc {
    const void* const* d_tables;  /* device array [n_ranks]: every consumer's slot-table buffer      */
    void* const* d_rings;         /* device array [n_ranks]: every consumer's batch ring             */
    uint32_t* const* d_flags;     /* device array [n_ranks]: every rank's flag page                  */
    int64_t table_offset_bytes;   /* where this step's slice starts inside a table buffer            */
    int64_t ring_offset_bytes;    /* where the ring slot starts inside a ring                        */
    uint64_t timeout_ns;          /* bound of every in-kernel wait, in nanoseconds (0 = unbounded)   */
    uint32_t* d_status;
    int32_t n_ranks, me;
    int32_t ring_slot, window;
    int32_t owner_lo, owner_hi;   /* source ids [lo, hi) whose shards live on this GPU               */
    int32_t ctas_per_consumer;    /* from bddl_push_ctas_per_consumer; the same on every launch      */
    uint32_t use;                 /* how often this ring slot has been filled before                 */
    uint32_t table_seq;           /* the consumers' `table` words must have reached this value       */
    /* replicated windows (K9/N6 multicast_put fused with the N5 epilogues): every consumer receives
     * rank `table_rank`'s window; with d_mc_ring (the NVLS multicast mapping of the rings) each sample
     * is stored once with multimem.st and replicated by NVSwitch, otherwise once per consumer. */
    void* d_mc_ring;              /* or NULL                                                         */
    int32_t table_rank;           /* -1: every consumer receives its own window                      */
    uint32_t fills;               /* earlier launches of this kind (unicast / multicast) into the slot */
} bddl_push_desc;
int bddl_push_ctas_per_consumer(int32_t n_ranks, int32_t max_ctas, int32_t* out);
int bddl_exchange_push_collate(const bddl_gather_desc* desc, const bddl_push_desc* push, void* stream);

/* Number of kernels this library has launched in this process (for gpu_launches accounting). */
int64_t bddl_launch_count(void);

/* ---- Batch checksum (no counterpart: the reference never verifies what a consumer receives,
 *      tests/test_ddl.py:8-28).  *d_out = sum over the 8-byte words of the batch of
 *      mix64(word ^ position), rows of payload_bytes at stride row_bytes.  d_scratch: caller-owned
 *      device memory of at least bddl_batch_checksum_scratch_words() 64-bit words. ------------- */
int64_t bddl_batch_checksum_scratch_words(int64_t n_rows, int64_t payload_bytes);
int bddl_batch_checksum(const void* d_batch, int64_t n_rows, int64_t row_bytes, int64_t payload_bytes,
                        uint64_t* d_scratch, int64_t scratch_words, uint64_t* d_out, void* stream);

/* Same fingerprint for callers that cannot allocate device memory (all HBM already owned by a
 * framework allocator, no tolerance for the device synchronisation cudaMalloc/cudaFree can cause,
 * or a host that forbids allocation after start-up).  This entry point has no counterpart in the
 * reference.  Scratch is a 64 KiB array inside the library image, held by every GPU that loads
 * the library; batches up to bddl_batch_checksum_static_max_bytes() (512 MiB), larger ones are
 * refused.  Calls that overlap on one GPU race on that array: order them.  The first call on a
 * device resolves the array's address; make it before any stream capture. */
int64_t bddl_batch_checksum_static_max_bytes(void);
int bddl_batch_checksum_static(const void* d_batch, int64_t n_rows, int64_t row_bytes, int64_t payload_bytes,
                               uint64_t* d_out, void* stream);

/* ---- N7: symmetric heap (replaces MPI.Win.Allocate_shared / Shared_query,
 *      ddl/connection.py:88-139): cudaMalloc segments exported as CUDA IPC handles and mapped
 *      by every peer; pinned, device-mapped host memory for host-resident datasets. ---------- */
int bddl_device_count(int32_t* n);
int bddl_set_device(int32_t dev);
int bddl_get_device(int32_t* dev);
int bddl_malloc(void** p, size_t bytes);
int bddl_free(void* p);
int bddl_memset_async(void* p, int value, size_t bytes, void* stream);
int bddl_memcpy_async(void* dst, const void* src, size_t bytes, int32_t kind, void* stream);
int bddl_memcpy2d_async(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes,
                        size_t rows, int32_t kind, void* stream);
int bddl_stream_synchronize(void* stream);
int bddl_event_create(void** ev);               /* cudaEventDisableTiming */
int bddl_event_destroy(void* ev);
int bddl_event_record(void* ev, void* stream);
int bddl_stream_wait_event(void* stream, void* ev);
int bddl_event_synchronize(void* ev);
int bddl_ipc_get_handle(void* p, uint8_t handle[64]);
int bddl_ipc_open_handle(const uint8_t handle[64], void** p);
int bddl_ipc_close_handle(void* p);
int bddl_can_access_peer(int32_t dev, int32_t peer, int32_t* ok);
int bddl_enable_peer_access(int32_t peer);
int bddl_host_alloc(void** host_ptr, void** dev_ptr, size_t bytes);
int bddl_host_free(void* host_ptr);
int bddl_host_register(void* host_ptr, size_t bytes, void** dev_ptr);
int bddl_host_unregister(void* host_ptr);

/* VMM form of the heap and N6 multicast_put (no counterpart in the reference): physical
 * allocations shared between processes as POSIX file descriptors, and an NVLS multicast object
 * every GPU binds memory to; a store through the multicast mapping (multimem.st) is replicated
 * by NVSwitch into every GPU's copy.  All feature-detected (bddl_vmm_query). */
int bddl_vmm_query(int32_t dev, int32_t* vmm, int32_t* posix_fd, int32_t* multicast);
int bddl_vmm_alloc(size_t bytes, int32_t dev, size_t granularity, void** ptr, uint64_t* handle, int32_t* fd,
                   size_t* rounded);
int bddl_vmm_import(int32_t fd, size_t size, int32_t dev, void** ptr, uint64_t* handle);
int bddl_vmm_free(void* ptr, size_t size, uint64_t handle);
int bddl_mc_granularity(size_t bytes, int32_t n_devices, size_t* gran);
int bddl_mc_create(size_t size, int32_t n_devices, uint64_t* mc, int32_t* fd);
int bddl_mc_import(int32_t fd, uint64_t* mc);
int bddl_mc_add_device(uint64_t mc, int32_t dev);
int bddl_mc_bind_and_map(uint64_t mc, uint64_t mem_handle, size_t size, int32_t dev, void** mc_ptr);
int bddl_multicast_put(void* mc_dst, const void* d_src, int64_t bytes, void* stream);

/* ---- N9: stage_host (replaces the window fill, tests/run_ddl.py:152-161, and the absent
 *      .to(device), tests/run_ddl.py:233-235, for host-resident datasets): one cudaMemcpyAsync
 *      per run of adjacent rows from pinned memory into a device staging buffer. ------------- */
int bddl_stage_rows(void* d_dst, int64_t dst_row_bytes, const void* h_src, int64_t row_bytes,
                    const int64_t* rows, int64_t n_rows, int64_t copy_bytes, void* stream, int64_t* n_copies);

int bddl_stage_slots(void* d_dst, int64_t dst_row_bytes, const void* const* h_bases, int32_t n_sources,
                     int64_t row_bytes, const uint64_t* slots, int64_t n, int64_t copy_bytes, void* stream,
                     int64_t* n_copies);

int bddl_stage_slots_mt(void* d_dst, int64_t dst_row_bytes, const void* const* h_bases, int32_t n_sources,
                        int64_t row_bytes, const uint64_t* slots, int64_t n, int64_t copy_bytes, void* stream,
                        int32_t n_threads);
/* CPU half of stage_host: gather a batch's rows from pinned shards into one contiguous pinned
 * buffer with a persistent thread pool (then a single cudaMemcpyAsync moves it). */
int bddl_host_gather_slots(void* h_dst, int64_t dst_row_bytes, const void* const* h_bases, int32_t n_sources,
                           int64_t row_bytes, const uint64_t* slots, int64_t n, int64_t copy_bytes,
                           int32_t n_threads);

/* Declared shuffle of a host window, out of place and threaded: dst[i] = X[perm[i]], X = src with
 * its first ne rows replaced by recv_block (rows received in the global exchange, or NULL).
 * Replaces rng.shuffle(my_ary) (tests/run_ddl.py:163-167) + the in-place Sendrecv_replace
 * (ddl/shuffle.py:95-108) for producer processes with two windows. */
int bddl_host_permute_rows(void* dst, const void* src, const void* recv_block, int64_t ne, const int64_t* perm,
                           int64_t n, int64_t row_bytes, int32_t n_threads);

/* ---- N8: flags (replace the zero-byte token messages, Win.Sync and the Ibarrier shutdown,
 *      ddl/connection.py:61-63,153-187).  Device side: counters in the symmetric heap. ------- */
int bddl_stream_wait_value32(void* stream, uint32_t* d_addr, uint32_t value);   /* wait *addr >= value */
int bddl_stream_write_value32(void* stream, uint32_t* d_addr, uint32_t value);
/* Store (add=0) or add (add=1) `value` to n counters (device array of addresses, peers allowed). */
int bddl_flag_signal(uint32_t* const* d_targets, int32_t n, uint32_t value, int32_t add, void* stream);
/* Bounded in-kernel wait (timeout in nanoseconds of the GPU's global timer, 0 = unbounded);
 * *d_status (device or mapped host memory): 0 satisfied, 1 timed out, 2 stopped. */
int bddl_flag_wait(const uint32_t* d_flag, uint32_t value, uint64_t timeout_ns, const uint32_t* d_stop,
                   uint32_t* d_status, void* stream);
/* All-ranks barrier over per-rank flag pages (d_pages[r] = rank r's page, n_ranks words each). */
int bddl_flag_barrier(uint32_t* const* d_pages, int32_t n_ranks, int32_t me, uint32_t epoch, uint64_t timeout_ns,
                      const uint32_t* d_stop, uint32_t* d_status, void* stream);

/* Host side: POSIX shared-memory windows and flag words for producer processes. */
int bddl_shm_create(const char* name, size_t bytes, void** ptr);
int bddl_shm_attach(const char* name, size_t bytes, void** ptr);
int bddl_shm_detach(void* ptr, size_t bytes);
int bddl_shm_unlink(const char* name);
void bddl_hflag_store(uint32_t* p, uint32_t v);
uint32_t bddl_hflag_load(const uint32_t* p);
uint32_t bddl_hflag_add(uint32_t* p, uint32_t v);
int bddl_hflag_wait_geq(const uint32_t* p, uint32_t value, int64_t timeout_us, const uint32_t* stop);

#ifdef __cplusplus
}
#endif
#endif /* DDL_B200_H */
