// SeedSequence + PCG64 (XSL-RR 128/64) + masked-rejection bounded draws, usable from host
// and device code.  Bit-exact with numpy.random.default_rng(seed): the reference's only
// source of randomness (reference ddl/shuffle.py:28-30 partner RNG, tests/run_ddl.py:122,167
// per-epoch row shuffle).  Algorithm as pinned in SURVEY.md section 3.6.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define BDDL_HD __host__ __device__ __forceinline__
#else
#define BDDL_HD inline
#endif

// C-ABI mirror lives in include/ddl_b200.h (bddl_rng_state); keep the layouts identical.
struct BddlRng {
    uint64_t state_lo, state_hi;  // 128-bit LCG state
    uint64_t inc_lo, inc_hi;      // 128-bit odd increment
    uint32_t has_uint32;          // numpy's buffered upper half of the last 64-bit output
    uint32_t uinteger;
    uint64_t draws32;             // statistics only: number of 32-bit draws handed out
};

struct U128 {
    uint64_t lo, hi;
};

BDDL_HD uint64_t bddl_mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}

BDDL_HD U128 u128_mul(U128 a, U128 b) {
    U128 r;
    r.lo = a.lo * b.lo;
    r.hi = bddl_mulhi64(a.lo, b.lo) + a.hi * b.lo + a.lo * b.hi;
    return r;
}

BDDL_HD U128 u128_add(U128 a, U128 b) {
    U128 r;
    r.lo = a.lo + b.lo;
    r.hi = a.hi + b.hi + (r.lo < a.lo ? 1u : 0u);
    return r;
}

// PCG 128-bit default multiplier 0x2360ED051FC65DA4_4385DF649FCCF645
#define BDDL_PCG_MULT_HI 0x2360ED051FC65DA4ULL
#define BDDL_PCG_MULT_LO 0x4385DF649FCCF645ULL

BDDL_HD U128 pcg_mult() {
    U128 m;
    m.lo = BDDL_PCG_MULT_LO;
    m.hi = BDDL_PCG_MULT_HI;
    return m;
}

BDDL_HD U128 pcg_step(U128 s, U128 inc) { return u128_add(u128_mul(s, pcg_mult()), inc); }

BDDL_HD uint64_t pcg_output(U128 s) {
    uint64_t x = s.hi ^ s.lo;
    unsigned rot = (unsigned)(s.hi >> 58);
    return (x >> rot) | (x << ((64u - rot) & 63u));
}

// (acc_mult, acc_plus) such that state advanced by `delta` steps = acc_mult*state + acc_plus.
BDDL_HD void pcg_jump_coeffs(U128 inc, uint64_t delta, U128* acc_mult, U128* acc_plus) {
    U128 am = {1, 0}, ap = {0, 0};
    U128 cm = pcg_mult(), cp = inc;
    while (delta > 0) {
        if (delta & 1) {
            am = u128_mul(am, cm);
            ap = u128_add(u128_mul(ap, cm), cp);
        }
        U128 cm1 = cm;
        cm1.lo += 1;
        if (cm1.lo == 0) cm1.hi += 1;
        cp = u128_mul(cm1, cp);
        cm = u128_mul(cm, cm);
        delta >>= 1;
    }
    *acc_mult = am;
    *acc_plus = ap;
}

BDDL_HD uint64_t bddl_next64(BddlRng* r) {
    U128 s = {r->state_lo, r->state_hi}, inc = {r->inc_lo, r->inc_hi};
    s = pcg_step(s, inc);
    r->state_lo = s.lo;
    r->state_hi = s.hi;
    return pcg_output(s);
}

// numpy hands out the low half of a fresh 64-bit output first, then the buffered high half.
BDDL_HD uint32_t bddl_next32(BddlRng* r) {
    r->draws32++;
    if (r->has_uint32) {
        r->has_uint32 = 0;
        return r->uinteger;
    }
    uint64_t v = bddl_next64(r);
    r->has_uint32 = 1;
    r->uinteger = (uint32_t)(v >> 32);
    return (uint32_t)v;
}

BDDL_HD uint64_t bddl_mask_for(uint64_t max) {
    uint64_t m = max;
    m |= m >> 1;
    m |= m >> 2;
    m |= m >> 4;
    m |= m >> 8;
    m |= m >> 16;
    m |= m >> 32;
    return m;
}

// numpy random_interval(): uniform integer in [0, max], masked rejection; no draw when max == 0.
BDDL_HD uint64_t bddl_interval(BddlRng* r, uint64_t max) {
    if (max == 0) return 0;
    uint64_t mask = bddl_mask_for(max);
    uint64_t v;
    if (max <= 0xffffffffULL) {
        do {
            v = (uint64_t)bddl_next32(r) & mask;
        } while (v > max);
    } else {
        do {
            v = bddl_next64(r) & mask;
        } while (v > max);
    }
    return v;
}

// ---- SeedSequence (pool of four uint32 words) ---------------------------------------------
#define BDDL_SS_INIT_A 0x43b0d7e5u
#define BDDL_SS_MULT_A 0x931e8875u
#define BDDL_SS_INIT_B 0x8b51f9ddu
#define BDDL_SS_MULT_B 0x58f38dedu
#define BDDL_SS_MIX_L 0xca01f9ddu
#define BDDL_SS_MIX_R 0x4973f715u

BDDL_HD uint32_t bddl_ss_hashmix(uint32_t v, uint32_t* hc) {
    v ^= *hc;
    *hc *= BDDL_SS_MULT_A;
    v *= *hc;
    v ^= v >> 16;
    return v;
}

BDDL_HD uint32_t bddl_ss_mix(uint32_t x, uint32_t y) {
    uint32_t r = BDDL_SS_MIX_L * x - BDDL_SS_MIX_R * y;
    r ^= r >> 16;
    return r;
}

// entropy: little-endian uint32 words of the (non-negative) integer seed; n_words >= 1.
BDDL_HD void bddl_seed_words(BddlRng* r, const uint32_t* entropy, int n_words) {
    uint32_t pool[4];
    uint32_t hc = BDDL_SS_INIT_A;
    for (int i = 0; i < 4; ++i) pool[i] = bddl_ss_hashmix(i < n_words ? entropy[i] : 0u, &hc);
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            if (i != j) pool[j] = bddl_ss_mix(pool[j], bddl_ss_hashmix(pool[i], &hc));
    for (int i = 4; i < n_words; ++i)
        for (int j = 0; j < 4; ++j) pool[j] = bddl_ss_mix(pool[j], bddl_ss_hashmix(entropy[i], &hc));

    uint32_t out[8];
    uint32_t hb = BDDL_SS_INIT_B;
    for (int i = 0; i < 8; ++i) {
        uint32_t d = pool[i & 3] ^ hb;
        hb *= BDDL_SS_MULT_B;
        d *= hb;
        d ^= d >> 16;
        out[i] = d;
    }
    uint64_t v0 = (uint64_t)out[0] | ((uint64_t)out[1] << 32);
    uint64_t v1 = (uint64_t)out[2] | ((uint64_t)out[3] << 32);
    uint64_t v2 = (uint64_t)out[4] | ((uint64_t)out[5] << 32);
    uint64_t v3 = (uint64_t)out[6] | ((uint64_t)out[7] << 32);

    U128 initstate = {v1, v0};  // (v0 << 64) | v1
    U128 initseq = {v3, v2};    // (v2 << 64) | v3
    U128 inc;
    inc.hi = (initseq.hi << 1) | (initseq.lo >> 63);
    inc.lo = (initseq.lo << 1) | 1u;
    U128 s = {0, 0};
    s = pcg_step(s, inc);
    s = u128_add(s, initstate);
    s = pcg_step(s, inc);
    r->state_lo = s.lo;
    r->state_hi = s.hi;
    r->inc_lo = inc.lo;
    r->inc_hi = inc.hi;
    r->has_uint32 = 0;
    r->uinteger = 0;
    r->draws32 = 0;
}

BDDL_HD void bddl_seed_u64(BddlRng* r, uint64_t seed) {
    uint32_t w[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    bddl_seed_words(r, w, w[1] ? 2 : 1);
}
