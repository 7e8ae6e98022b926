// Host side of N1/N3: RNG entry points, host Fisher-Yates and partner selection.
// Replaces numpy Generator.shuffle (reference tests/run_ddl.py:163-167) and
// GlobalShuffler._calculate_comm_partner (reference ddl/shuffle.py:32-79).
#include <string.h>

#include <vector>

#include "../include/ddl_b200.h"
#include "common.h"
#include "rng_pcg64.h"

static_assert(sizeof(bddl_rng_state) == sizeof(BddlRng), "rng state layouts must match");

static thread_local char g_err[512];
char* bddl_err_buf() { return g_err; }

int bddl_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

template <typename T>
static void shuffle_table(BddlRng* r, T* t, int64_t n) {
    for (int64_t i = n - 1; i >= 1; --i) {
        int64_t j = (int64_t)bddl_interval(r, (uint64_t)i);
        T a = t[i];
        t[i] = t[j];
        t[j] = a;
    }
}

extern "C" {

const char* bddl_last_error(void) { return g_err; }
int bddl_abi_version(void) { return BDDL_ABI_VERSION; }

void bddl_rng_seed(bddl_rng_state* s, uint64_t seed) { bddl_seed_u64((BddlRng*)s, seed); }
uint32_t bddl_rng_next32(bddl_rng_state* s) { return bddl_next32((BddlRng*)s); }
uint64_t bddl_rng_next64(bddl_rng_state* s) { return bddl_next64((BddlRng*)s); }
uint64_t bddl_rng_interval(bddl_rng_state* s, uint64_t max_inclusive) {
    return bddl_interval((BddlRng*)s, max_inclusive);
}

void bddl_rng_advance(bddl_rng_state* s, uint64_t n) {
    U128 am, ap, st = {s->state_lo, s->state_hi}, inc = {s->inc_lo, s->inc_hi};
    pcg_jump_coeffs(inc, n, &am, &ap);
    st = u128_add(u128_mul(am, st), ap);
    s->state_lo = st.lo;
    s->state_hi = st.hi;
}

int bddl_host_shuffle_u64(bddl_rng_state* s, uint64_t* table, int64_t n) {
    if (!s || (!table && n > 0) || n < 0) return bddl_fail(BDDL_ERR_INVALID, "host_shuffle_u64: bad arguments");
    shuffle_table((BddlRng*)s, table, n);
    return 0;
}

int bddl_host_shuffle_i32(bddl_rng_state* s, int32_t* table, int64_t n) {
    if (!s || (!table && n > 0) || n < 0) return bddl_fail(BDDL_ERR_INVALID, "host_shuffle_i32: bad arguments");
    shuffle_table((BddlRng*)s, table, n);
    return 0;
}

int bddl_host_shuffle_rows(bddl_rng_state* s, void* rows, int64_t n, int64_t row_bytes) {
    if (!s || (!rows && n > 0) || n < 0 || row_bytes <= 0)
        return bddl_fail(BDDL_ERR_INVALID, "host_shuffle_rows: bad arguments");
    BddlRng* r = (BddlRng*)s;
    std::vector<char> tmp((size_t)row_bytes);
    char* base = (char*)rows;
    for (int64_t i = n - 1; i >= 1; --i) {
        int64_t j = (int64_t)bddl_interval(r, (uint64_t)i);
        if (i == j) continue;
        char* a = base + i * row_bytes;
        char* b = base + j * row_bytes;
        memcpy(tmp.data(), b, (size_t)row_bytes);
        memcpy(b, a, (size_t)row_bytes);
        memcpy(a, tmp.data(), (size_t)row_bytes);
    }
    return 0;
}

int bddl_partner_select(bddl_rng_state* s, int32_t n, int32_t me, int32_t* send_to,
                        int32_t* recv_from, int32_t* perm_out, int32_t* tries_out) {
    if (!s || !send_to || !recv_from || n < 1 || me < 0 || me >= n)
        return bddl_fail(BDDL_ERR_INVALID, "partner_select: bad arguments (n=%d me=%d)", n, me);
    if (tries_out) *tries_out = 0;
    if (n == 1) {
        *send_to = 0;
        *recv_from = 0;
        if (perm_out) perm_out[0] = 0;
        return 0;
    }
    if (n == 2) {
        // The reference hands rank 0 (1,1) and rank 1 (0,0); no random draw is consumed.
        *send_to = me == 0 ? 1 : 0;
        *recv_from = *send_to;
        if (perm_out) {
            perm_out[0] = 1;
            perm_out[1] = 0;
        }
        return 0;
    }
    // The permutation persists across rejected attempts: each try reshuffles the array the
    // previous try left behind, starting from the identity.
    std::vector<int32_t> col((size_t)n), inv((size_t)n);
    for (int32_t i = 0; i < n; ++i) col[i] = i;
    for (int cnt = 0;; ++cnt) {
        shuffle_table((BddlRng*)s, col.data(), n);
        for (int32_t i = 0; i < n; ++i) inv[col[i]] = i;
        bool ok = true;
        for (int32_t r = 0; r < n && ok; ++r)
            if (col[r] == r || col[r] == inv[r]) ok = false;
        if (ok) {
            *send_to = col[me];
            *recv_from = inv[me];
            if (perm_out) memcpy(perm_out, col.data(), sizeof(int32_t) * (size_t)n);
            if (tries_out) *tries_out = cnt;
            return 0;
        }
        if (cnt + 1 > 1000)
            return bddl_fail(BDDL_ERR_PARTNER,
                             "partner_select: no valid communication pattern after %d tries", cnt + 1);
    }
}

}  // extern "C"
