/* vsp_oracle.c — CPU ORACLE for the VoteSAVER Groth16/SAVER hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's algorithms, used (a) as the checker in tests/ and in
 * __graft_entry__.smoke(), and (b) as the timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`).
 * The product path (vote-saver-protocol_b200/) never links, loads or calls it.
 *
 * The reference itself cannot be compiled here (Boost headers absent; crypto3-algebra / -zk / -multiprecision
 * submodules empty: SURVEY.md 0.1-0.2), so this is a "port" oracle.  What it follows, paths relative to
 * /root/reference:
 *   [SRC]  libs/math/include/nil/crypto3/math/domains/detail/basic_radix2_domain_aux.hpp:46-83   orc_fr_basic_radix2_fft
 *   [SRC]  libs/math/include/nil/crypto3/math/domains/basic_radix2_domain.hpp:69-96,118-124      basic domain fft/ifft/divide
 *   [SRC]  libs/math/include/nil/crypto3/math/domains/step_radix2_domain.hpp:73-166,222-246     step domain fft/ifft/divide
 *   [SRC]  libs/math/include/nil/crypto3/math/algorithms/make_evaluation_domain.hpp:51-106       orc_domain_plan
 *   [SRC]  libs/math/include/nil/crypto3/math/algorithms/unity_root.hpp:52-71, coset.hpp:37-43
 *   [SRC]  libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:341-352, 632-667         SAVER ciphertext, rerandomize
 *   [UPSTREAM-RECALL] (crypto3-zk @ d335bb8b, crypto3-algebra @ e4fa2c5e, .SUBMODULES.json:199,17)
 *          r1cs_to_qap::witness_map, multiexp<BDLO12> / multiexp_with_mixed_addition, r1cs_gg_ppzksnark prover.
 * Pinning: Fr FFT, G1/G2 addition and encodings are checked against the reference's own known answers
 * (tests/golden/, tests/test_oracle.py).  Whole proofs / ciphertexts: PARITY UNPINNED — the reference has no test
 * fixing r, s or proof bytes; the gate there is generate -> prove -> verify self-consistency (oracle/pyref.py).
 *
 * ABI: canonical little-endian u64 limbs.  Fr = 4 limbs, Fq = 6 limbs, G1 affine = x|y (12 limbs),
 * G2 affine = x.c0|x.c1|y.c0|y.c1 (24 limbs); the point at infinity is all-zero.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint64_t u64;
typedef unsigned __int128 u128;

/* ------------------------------------------------------------------------------------------ */
/* generic N-limb Montgomery field                                                             */
/* ------------------------------------------------------------------------------------------ */
#define FQ_N 6
#define FR_N 4
typedef struct { u64 v[FQ_N]; } fq_t;
typedef struct { u64 v[FR_N]; } fr_t;

static const u64 FQ_P[FQ_N] = {0xb9feffffffffaaabULL, 0x1eabfffeb153ffffULL, 0x6730d2a0f6b0f624ULL,
                               0x64774b84f38512bfULL, 0x4b1ba7b6434bacd7ULL, 0x1a0111ea397fe69aULL};
static const u64 FR_P[FR_N] = {0xffffffff00000001ULL, 0x53bda402fffe5bfeULL, 0x3339d80809a1d805ULL,
                               0x73eda753299d7d48ULL};
static u64 FQ_INV, FR_INV;     /* -p^-1 mod 2^64 */
static fq_t FQ_R1, FQ_R2;      /* R mod p, R^2 mod p */
static fr_t FR_R1, FR_R2;
static fr_t FR_ROOT;           /* 7^((r-1)/2^32), Montgomery form */
static fr_t FR_GEN;            /* multiplicative generator 7, Montgomery form */
static int g_init_done = 0;

static inline int n_geq(const u64 *a, const u64 *b, int N) {
    for (int i = N - 1; i >= 0; --i) {
        if (a[i] != b[i]) return a[i] > b[i];
    }
    return 1;
}
static inline u64 n_add(u64 *r, const u64 *a, const u64 *b, int N) {
    u128 c = 0;
    for (int i = 0; i < N; ++i) {
        c += (u128)a[i] + b[i];
        r[i] = (u64)c;
        c >>= 64;
    }
    return (u64)c;
}
static inline u64 n_sub(u64 *r, const u64 *a, const u64 *b, int N) {
    u64 borrow = 0;
    for (int i = 0; i < N; ++i) {
        u128 d = (u128)a[i] - b[i] - borrow;
        r[i] = (u64)d;
        borrow = (u64)(d >> 64) & 1;
    }
    return borrow;
}
static inline void m_add(u64 *r, const u64 *a, const u64 *b, const u64 *p, int N) {
    u64 t[FQ_N];
    u64 c = n_add(t, a, b, N);
    if (c || n_geq(t, p, N)) n_sub(t, t, p, N);
    memcpy(r, t, 8 * N);
}
static inline void m_sub(u64 *r, const u64 *a, const u64 *b, const u64 *p, int N) {
    u64 t[FQ_N];
    if (n_sub(t, a, b, N)) n_add(t, t, p, N);
    memcpy(r, t, 8 * N);
}
static inline int n_iszero(const u64 *a, int N) {
    u64 x = 0;
    for (int i = 0; i < N; ++i) x |= a[i];
    return x == 0;
}
/* CIOS Montgomery multiplication */
static inline __attribute__((always_inline)) void m_mul(u64 *r, const u64 *a, const u64 *b, const u64 *p, u64 inv,
                                                        const int N) {
    u64 t[FQ_N + 2];
    for (int i = 0; i < N + 2; ++i) t[i] = 0;
    for (int i = 0; i < N; ++i) {
        u128 c = 0;
        for (int j = 0; j < N; ++j) {
            c += (u128)a[j] * b[i] + t[j];
            t[j] = (u64)c;
            c >>= 64;
        }
        c += t[N];
        t[N] = (u64)c;
        t[N + 1] = (u64)(c >> 64);
        u64 m = t[0] * inv;
        c = (u128)m * p[0] + t[0];
        c >>= 64;
        for (int j = 1; j < N; ++j) {
            c += (u128)m * p[j] + t[j];
            t[j - 1] = (u64)c;
            c >>= 64;
        }
        c += t[N];
        t[N - 1] = (u64)c;
        t[N] = t[N + 1] + (u64)(c >> 64);
    }
    if (t[N] || n_geq(t, p, N)) n_sub(t, t, p, N);
    memcpy(r, t, 8 * N);
}

/* Fq */
static inline void fq_add(fq_t *r, const fq_t *a, const fq_t *b) { m_add(r->v, a->v, b->v, FQ_P, FQ_N); }
static inline void fq_sub(fq_t *r, const fq_t *a, const fq_t *b) { m_sub(r->v, a->v, b->v, FQ_P, FQ_N); }
static void fq_mul(fq_t *r, const fq_t *a, const fq_t *b) { m_mul(r->v, a->v, b->v, FQ_P, FQ_INV, FQ_N); }
static inline void fq_sqr(fq_t *r, const fq_t *a) { fq_mul(r, a, a); }
static inline int fq_iszero(const fq_t *a) { return n_iszero(a->v, FQ_N); }
static inline int fq_eq(const fq_t *a, const fq_t *b) { return memcmp(a, b, sizeof(fq_t)) == 0; }
static inline void fq_zero(fq_t *a) { memset(a, 0, sizeof(*a)); }
static inline void fq_one(fq_t *a) { *a = FQ_R1; }
static inline void fq_neg(fq_t *r, const fq_t *a) {
    if (fq_iszero(a)) {
        *r = *a;
    } else {
        n_sub(r->v, FQ_P, a->v, FQ_N);
    }
}
static void fq_pow(fq_t *r, const fq_t *a, const u64 *e, int elimbs) {
    fq_t acc = FQ_R1, base = *a;
    for (int i = 0; i < 64 * elimbs; ++i) {
        if ((e[i >> 6] >> (i & 63)) & 1) fq_mul(&acc, &acc, &base);
        fq_mul(&base, &base, &base);
    }
    *r = acc;
}
static void fq_inv(fq_t *r, const fq_t *a) {
    u64 e[FQ_N], two[FQ_N] = {2, 0, 0, 0, 0, 0};
    n_sub(e, FQ_P, two, FQ_N);
    fq_pow(r, a, e, FQ_N);
}
static inline void fq_from_canon(fq_t *r, const u64 *c) {
    fq_t t;
    memcpy(t.v, c, sizeof(t));
    fq_mul(r, &t, &FQ_R2);
}
static inline void fq_to_canon(u64 *c, const fq_t *a) {
    fq_t one;
    fq_zero(&one);
    one.v[0] = 1;
    fq_t t;
    fq_mul(&t, a, &one);
    memcpy(c, t.v, sizeof(t));
}

/* Fr */
static inline void fr_add(fr_t *r, const fr_t *a, const fr_t *b) { m_add(r->v, a->v, b->v, FR_P, FR_N); }
static inline void fr_sub(fr_t *r, const fr_t *a, const fr_t *b) { m_sub(r->v, a->v, b->v, FR_P, FR_N); }
static void fr_mul(fr_t *r, const fr_t *a, const fr_t *b) { m_mul(r->v, a->v, b->v, FR_P, FR_INV, FR_N); }
static inline int fr_iszero(const fr_t *a) { return n_iszero(a->v, FR_N); }
static inline void fr_zero(fr_t *a) { memset(a, 0, sizeof(*a)); }
static void fr_pow_u64(fr_t *r, const fr_t *a, u64 e) {
    fr_t acc = FR_R1, base = *a;
    while (e) {
        if (e & 1) fr_mul(&acc, &acc, &base);
        fr_mul(&base, &base, &base);
        e >>= 1;
    }
    *r = acc;
}
static void fr_inv(fr_t *r, const fr_t *a) {
    u64 e[FR_N], two[FR_N] = {2, 0, 0, 0};
    n_sub(e, FR_P, two, FR_N);
    fr_t acc = FR_R1, base = *a;
    for (int i = 0; i < 256; ++i) {
        if ((e[i >> 6] >> (i & 63)) & 1) fr_mul(&acc, &acc, &base);
        fr_mul(&base, &base, &base);
    }
    *r = acc;
}
static inline void fr_from_canon(fr_t *r, const u64 *c) {
    fr_t t;
    memcpy(t.v, c, sizeof(t));
    fr_mul(r, &t, &FR_R2);
}
static inline void fr_to_canon(u64 *c, const fr_t *a) {
    fr_t one, t;
    fr_zero(&one);
    one.v[0] = 1;
    fr_mul(&t, a, &one);
    memcpy(c, t.v, sizeof(t));
}
static inline void fr_from_u64(fr_t *r, u64 x) {
    u64 c[FR_N] = {x, 0, 0, 0};
    fr_from_canon(r, c);
}

/* Fq2 = Fq[u]/(u^2 + 1) */
typedef struct { fq_t c0, c1; } fq2_t;
static inline void fq2_add(fq2_t *r, const fq2_t *a, const fq2_t *b) {
    fq_add(&r->c0, &a->c0, &b->c0);
    fq_add(&r->c1, &a->c1, &b->c1);
}
static inline void fq2_sub(fq2_t *r, const fq2_t *a, const fq2_t *b) {
    fq_sub(&r->c0, &a->c0, &b->c0);
    fq_sub(&r->c1, &a->c1, &b->c1);
}
static void fq2_mul(fq2_t *r, const fq2_t *a, const fq2_t *b) {
    fq_t t0, t1, t2, t3;
    fq_mul(&t0, &a->c0, &b->c0);
    fq_mul(&t1, &a->c1, &b->c1);
    fq_add(&t2, &a->c0, &a->c1);
    fq_add(&t3, &b->c0, &b->c1);
    fq_mul(&t2, &t2, &t3);
    fq_sub(&t2, &t2, &t0);
    fq_sub(&r->c1, &t2, &t1);
    fq_sub(&r->c0, &t0, &t1);
}
static void fq2_sqr(fq2_t *r, const fq2_t *a) {
    fq_t s, d, m;
    fq_add(&s, &a->c0, &a->c1);
    fq_sub(&d, &a->c0, &a->c1);
    fq_mul(&m, &a->c0, &a->c1);
    fq_mul(&r->c0, &s, &d);
    fq_add(&r->c1, &m, &m);
}
static inline void fq2_neg(fq2_t *r, const fq2_t *a) {
    fq_neg(&r->c0, &a->c0);
    fq_neg(&r->c1, &a->c1);
}
static void fq2_inv(fq2_t *r, const fq2_t *a) {
    fq_t n, t;
    fq_sqr(&n, &a->c0);
    fq_sqr(&t, &a->c1);
    fq_add(&n, &n, &t);
    fq_inv(&n, &n);
    fq_mul(&r->c0, &a->c0, &n);
    fq_mul(&t, &a->c1, &n);
    fq_neg(&r->c1, &t);
}
static inline int fq2_iszero(const fq2_t *a) { return fq_iszero(&a->c0) && fq_iszero(&a->c1); }
static inline int fq2_eq(const fq2_t *a, const fq2_t *b) { return fq_eq(&a->c0, &b->c0) && fq_eq(&a->c1, &b->c1); }
static inline void fq2_zero(fq2_t *a) { memset(a, 0, sizeof(*a)); }
static inline void fq2_one(fq2_t *a) {
    a->c0 = FQ_R1;
    fq_zero(&a->c1);
}

/* ------------------------------------------------------------------------------------------ */
/* curves                                                                                      */
/* ------------------------------------------------------------------------------------------ */
#define CV(n) g1_##n
#define F_T fq_t
#define F_ADD fq_add
#define F_SUB fq_sub
#define F_MUL fq_mul
#define F_SQR fq_sqr
#define F_NEG fq_neg
#define F_INV fq_inv
#define F_ISZERO fq_iszero
#define F_EQ fq_eq
#define F_ONE fq_one
#define F_ZERO fq_zero
#include "curve_tmpl.h"
#undef CV
#undef F_T
#undef F_ADD
#undef F_SUB
#undef F_MUL
#undef F_SQR
#undef F_NEG
#undef F_INV
#undef F_ISZERO
#undef F_EQ
#undef F_ONE
#undef F_ZERO

#define CV(n) g2_##n
#define F_T fq2_t
#define F_ADD fq2_add
#define F_SUB fq2_sub
#define F_MUL fq2_mul
#define F_SQR fq2_sqr
#define F_NEG fq2_neg
#define F_INV fq2_inv
#define F_ISZERO fq2_iszero
#define F_EQ fq2_eq
#define F_ONE fq2_one
#define F_ZERO fq2_zero
#include "curve_tmpl.h"
#undef CV

static void g1_aff_from_canon(g1_aff *r, const u64 *c) {
    fq_from_canon(&r->x, c);
    fq_from_canon(&r->y, c + 6);
}
static void g1_aff_to_canon(u64 *c, const g1_aff *a) {
    fq_to_canon(c, &a->x);
    fq_to_canon(c + 6, &a->y);
}
static void g2_aff_from_canon(g2_aff *r, const u64 *c) {
    fq_from_canon(&r->x.c0, c);
    fq_from_canon(&r->x.c1, c + 6);
    fq_from_canon(&r->y.c0, c + 12);
    fq_from_canon(&r->y.c1, c + 18);
}
static void g2_aff_to_canon(u64 *c, const g2_aff *a) {
    fq_to_canon(c, &a->x.c0);
    fq_to_canon(c + 6, &a->x.c1);
    fq_to_canon(c + 12, &a->y.c0);
    fq_to_canon(c + 18, &a->y.c1);
}
static void g1_jac_to_canon(u64 *c, const g1_jac *p) {
    g1_aff a;
    g1_to_aff(&a, p);
    g1_aff_to_canon(c, &a);
}
static void g2_jac_to_canon(u64 *c, const g2_jac *p) {
    g2_aff a;
    g2_to_aff(&a, p);
    g2_aff_to_canon(c, &a);
}
static void g1_jac_from_canon(g1_jac *p, const u64 *c) {
    g1_aff a;
    g1_aff_from_canon(&a, c);
    g1_from_aff(p, &a);
}
static void g2_jac_from_canon(g2_jac *p, const u64 *c) {
    g2_aff a;
    g2_aff_from_canon(&a, c);
    g2_from_aff(p, &a);
}

/* ------------------------------------------------------------------------------------------ */
/* init                                                                                        */
/* ------------------------------------------------------------------------------------------ */
static u64 neg_inv64(u64 p0) {
    u64 x = 1;
    for (int i = 0; i < 7; ++i) x *= 2 - p0 * x;
    return (u64)0 - x;
}
static void pow2_mod(u64 *r, int bits, const u64 *p, int N) {
    u64 t[FQ_N] = {1, 0, 0, 0, 0, 0};
    for (int i = 0; i < bits; ++i) m_add(t, t, t, p, N);
    memcpy(r, t, 8 * N);
}
void orc_init(void) {
    if (g_init_done) return;
    FQ_INV = neg_inv64(FQ_P[0]);
    FR_INV = neg_inv64(FR_P[0]);
    pow2_mod(FQ_R1.v, 64 * FQ_N, FQ_P, FQ_N);
    pow2_mod(FQ_R2.v, 128 * FQ_N, FQ_P, FQ_N);
    pow2_mod(FR_R1.v, 64 * FR_N, FR_P, FR_N);
    pow2_mod(FR_R2.v, 128 * FR_N, FR_P, FR_N);
    fr_from_u64(&FR_GEN, 7);
    /* root_of_unity = 7^((r-1) / 2^32) */
    u64 e[FR_N], one[FR_N] = {1, 0, 0, 0};
    n_sub(e, FR_P, one, FR_N);
    for (int i = 0; i < FR_N; ++i) e[i] = (e[i] >> 32) | (i + 1 < FR_N ? e[i + 1] << 32 : 0);
    fr_t acc = FR_R1, base = FR_GEN;
    for (int i = 0; i < 256; ++i) {
        if ((e[i >> 6] >> (i & 63)) & 1) fr_mul(&acc, &acc, &base);
        fr_mul(&base, &base, &base);
    }
    FR_ROOT = acc;
    g_init_done = 1;
}

/* ------------------------------------------------------------------------------------------ */
/* FFT and evaluation domains  [SRC]                                                           */
/* ------------------------------------------------------------------------------------------ */
static size_t ceil_log2(size_t n) {
    size_t l = 0;
    while (((size_t)1 << l) < n) ++l;
    return l;
}
static size_t bitreverse(size_t n, size_t l) {
    size_t r = 0;
    for (size_t k = 0; k < l; ++k) {
        r = (r << 1) | (n & 1);
        n >>= 1;
    }
    return r;
}
/* unity_root.hpp:52-71 */
static int unity_root(fr_t *w, size_t n) {
    size_t logn = ceil_log2(n);
    if (n != ((size_t)1 << logn) || logn > 32) return -1;
    *w = FR_ROOT;
    for (size_t i = 32; i > logn; --i) fr_mul(w, w, w);
    return 0;
}
/* basic_radix2_domain_aux.hpp:46-83.  Montgomery-form in, Montgomery-form out. */
static int basic_radix2_fft(fr_t *a, size_t n, const fr_t *omega) {
    size_t logn = ceil_log2(n);
    if (n != ((size_t)1 << logn)) return -1;
    for (size_t k = 0; k < n; ++k) {
        size_t rk = bitreverse(k, logn);
        if (k < rk) {
            fr_t t = a[k];
            a[k] = a[rk];
            a[rk] = t;
        }
    }
    size_t m = 1;
    for (size_t s = 1; s <= logn; ++s) {
        fr_t w_m;
        fr_pow_u64(&w_m, omega, n / (2 * m));
        for (size_t k = 0; k < n; k += 2 * m) {
            fr_t w = FR_R1;
            for (size_t j = 0; j < m; ++j) {
                fr_t t;
                fr_mul(&t, &w, &a[k + j + m]);
                fr_sub(&a[k + j + m], &a[k + j], &t);
                fr_add(&a[k + j], &a[k + j], &t);
                fr_mul(&w, &w, &w_m);
            }
        }
        m *= 2;
    }
    return 0;
}

enum { ORC_DOMAIN_BASIC = 1, ORC_DOMAIN_EXTENDED = 2, ORC_DOMAIN_STEP = 3 };

static int is_basic(size_t m) { return m > 1 && ceil_log2(m) <= 32 && m == ((size_t)1 << ceil_log2(m)); }
static int is_extended(size_t m) {
    if (m <= 1) return 0;
    size_t small_m = m / 2;
    return ceil_log2(m) == 33 && small_m == ((size_t)1 << ceil_log2(small_m)) && ceil_log2(small_m) <= 32;
}
static int is_step(size_t m) {
    if (m <= 1) return 0;
    size_t log_m = ceil_log2(m);
    size_t small_m = m - ((size_t)1 << (log_m - 1));
    if (small_m == 0) return 0;
    return small_m == ((size_t)1 << ceil_log2(small_m)) && log_m <= 32;
}
/* make_evaluation_domain.hpp:51-106: which domain class and which size a request for m gets */
int orc_domain_plan(u64 m, int *kind, u64 *size) {
    if (m <= 1) return -1;
    size_t big = (size_t)1 << (ceil_log2(m) - 1);
    size_t rounded_small = (size_t)1 << ceil_log2(m - big);
    size_t cand[2] = {m, big + rounded_small};
    for (int i = 0; i < 2; ++i) {
        if (is_basic(cand[i])) { *kind = ORC_DOMAIN_BASIC; *size = cand[i]; return 0; }
        if (is_extended(cand[i])) { *kind = ORC_DOMAIN_EXTENDED; *size = cand[i]; return 0; }
        if (is_step(cand[i])) { *kind = ORC_DOMAIN_STEP; *size = cand[i]; return 0; }
    }
    return -1;
}

typedef struct {
    int kind;
    size_t m, big_m, small_m;
    fr_t omega, big_omega, small_omega;
} domain_t;

static int domain_init(domain_t *d, size_t m) {
    memset(d, 0, sizeof(*d));
    d->m = m;
    if (is_basic(m)) {
        d->kind = ORC_DOMAIN_BASIC;
        return unity_root(&d->omega, m);
    }
    if (is_step(m)) {
        d->kind = ORC_DOMAIN_STEP;
        d->big_m = (size_t)1 << (ceil_log2(m) - 1);
        d->small_m = m - d->big_m;
        if (unity_root(&d->omega, (size_t)1 << ceil_log2(m))) return -1;
        fr_mul(&d->big_omega, &d->omega, &d->omega);
        if (d->small_m > 1) {
            if (unity_root(&d->small_omega, d->small_m)) return -1;
        } else {
            d->small_omega = FR_R1;
        }
        return 0;
    }
    return -1;
}

static void scale_all(fr_t *a, size_t n, const fr_t *k) {
    for (size_t i = 0; i < n; ++i) fr_mul(&a[i], &a[i], k);
}

/* a has exactly d->m elements (callers zero-pad as basic_radix2_domain.hpp:70-77 does) */
static int domain_fft(const domain_t *d, fr_t *a) {
    if (d->kind == ORC_DOMAIN_BASIC) return basic_radix2_fft(a, d->m, &d->omega);
    /* step_radix2_domain.hpp:73-111 */
    size_t big_m = d->big_m, small_m = d->small_m;
    fr_t *c = (fr_t *)calloc(big_m, sizeof(fr_t));
    fr_t *dd = (fr_t *)calloc(big_m, sizeof(fr_t));
    fr_t *e = (fr_t *)calloc(small_m, sizeof(fr_t));
    fr_t omega_i = FR_R1;
    for (size_t i = 0; i < big_m; ++i) {
        if (i < small_m) {
            fr_t t;
            fr_add(&c[i], &a[i], &a[i + big_m]);
            fr_sub(&t, &a[i], &a[i + big_m]);
            fr_mul(&dd[i], &omega_i, &t);
        } else {
            c[i] = a[i];
            fr_mul(&dd[i], &omega_i, &a[i]);
        }
        fr_mul(&omega_i, &omega_i, &d->omega);
    }
    size_t compr = (size_t)1 << (ceil_log2(big_m) - ceil_log2(small_m));
    for (size_t i = 0; i < small_m; ++i)
        for (size_t j = 0; j < compr; ++j) fr_add(&e[i], &e[i], &dd[i + j * small_m]);
    basic_radix2_fft(c, big_m, &d->big_omega);
    basic_radix2_fft(e, small_m, &d->small_omega);
    memcpy(a, c, big_m * sizeof(fr_t));
    memcpy(a + big_m, e, small_m * sizeof(fr_t));
    free(c);
    free(dd);
    free(e);
    return 0;
}

static int domain_ifft(const domain_t *d, fr_t *a) {
    if (d->kind == ORC_DOMAIN_BASIC) {
        /* basic_radix2_domain.hpp:81-96 */
        fr_t winv, ninv;
        fr_inv(&winv, &d->omega);
        if (basic_radix2_fft(a, d->m, &winv)) return -1;
        fr_from_u64(&ninv, (u64)d->m);
        fr_inv(&ninv, &ninv);
        scale_all(a, d->m, &ninv);
        return 0;
    }
    /* step_radix2_domain.hpp:112-166 */
    size_t big_m = d->big_m, small_m = d->small_m;
    fr_t *U0 = (fr_t *)malloc(big_m * sizeof(fr_t));
    fr_t *U1 = (fr_t *)malloc(small_m * sizeof(fr_t));
    fr_t *tmp = (fr_t *)malloc(big_m * sizeof(fr_t));
    memcpy(U0, a, big_m * sizeof(fr_t));
    memcpy(U1, a + big_m, small_m * sizeof(fr_t));
    fr_t w;
    fr_inv(&w, &d->big_omega);
    basic_radix2_fft(U0, big_m, &w);
    fr_inv(&w, &d->small_omega);
    basic_radix2_fft(U1, small_m, &w);
    fr_t k;
    fr_from_u64(&k, (u64)big_m);
    fr_inv(&k, &k);
    scale_all(U0, big_m, &k);
    fr_from_u64(&k, (u64)small_m);
    fr_inv(&k, &k);
    scale_all(U1, small_m, &k);
    fr_t omega_i = FR_R1;
    for (size_t i = 0; i < big_m; ++i) {
        fr_mul(&tmp[i], &U0[i], &omega_i);
        fr_mul(&omega_i, &omega_i, &d->omega);
    }
    for (size_t i = small_m; i < big_m; ++i) a[i] = U0[i];
    size_t compr = (size_t)1 << (ceil_log2(big_m) - ceil_log2(small_m));
    for (size_t i = 0; i < small_m; ++i)
        for (size_t j = 1; j < compr; ++j) fr_sub(&U1[i], &U1[i], &tmp[i + j * small_m]);
    fr_t omega_inv, omega_inv_i = FR_R1;
    fr_inv(&omega_inv, &d->omega);
    for (size_t i = 0; i < small_m; ++i) {
        fr_mul(&U1[i], &U1[i], &omega_inv_i);
        fr_mul(&omega_inv_i, &omega_inv_i, &omega_inv);
    }
    fr_t over_two;
    fr_from_u64(&over_two, 2);
    fr_inv(&over_two, &over_two);
    for (size_t i = 0; i < small_m; ++i) {
        fr_t s, df;
        fr_add(&s, &U0[i], &U1[i]);
        fr_sub(&df, &U0[i], &U1[i]);
        fr_mul(&a[i], &s, &over_two);
        fr_mul(&a[big_m + i], &df, &over_two);
    }
    free(U0);
    free(U1);
    free(tmp);
    return 0;
}

/* coset.hpp:37-43 */
static void multiply_by_coset(fr_t *a, size_t n, const fr_t *g) {
    fr_t u = *g;
    for (size_t i = 1; i < n; ++i) {
        fr_mul(&a[i], &a[i], &u);
        fr_mul(&u, &u, g);
    }
}

static void fr_pow_sz(fr_t *r, const fr_t *a, size_t e) { fr_pow_u64(r, a, (u64)e); }

/* basic_radix2_domain.hpp:118-124, step_radix2_domain.hpp:222-246 */
static void divide_by_z_on_coset(const domain_t *d, fr_t *P) {
    if (d->kind == ORC_DOMAIN_BASIC) {
        fr_t z;
        fr_pow_sz(&z, &FR_GEN, d->m);
        fr_sub(&z, &z, &FR_R1);
        fr_inv(&z, &z);
        scale_all(P, d->m, &z);
        return;
    }
    size_t big_m = d->big_m, small_m = d->small_m;
    fr_t Z0, c_small_Z0, w_small, w_small_Z0, w_2small, elt = FR_R1, t;
    fr_pow_sz(&Z0, &FR_GEN, big_m);
    fr_sub(&Z0, &Z0, &FR_R1);
    fr_pow_sz(&c_small_Z0, &FR_GEN, small_m);
    fr_mul(&c_small_Z0, &c_small_Z0, &Z0);
    fr_pow_sz(&w_small, &d->omega, small_m);
    fr_mul(&w_small_Z0, &w_small, &Z0);
    fr_pow_sz(&w_2small, &d->omega, 2 * small_m);
    for (size_t i = 0; i < big_m; ++i) {
        fr_mul(&t, &c_small_Z0, &elt);
        fr_sub(&t, &t, &w_small_Z0);
        fr_inv(&t, &t);
        fr_mul(&P[i], &P[i], &t);
        fr_mul(&elt, &elt, &w_2small);
    }
    fr_t cw, a1, a2, Z1;
    fr_mul(&cw, &FR_GEN, &d->omega);
    fr_pow_sz(&a1, &cw, big_m);
    fr_sub(&a1, &a1, &FR_R1);
    fr_pow_sz(&a2, &cw, small_m);
    fr_sub(&a2, &a2, &w_small);
    fr_mul(&Z1, &a1, &a2);
    fr_inv(&Z1, &Z1);
    for (size_t i = 0; i < small_m; ++i) fr_mul(&P[big_m + i], &P[big_m + i], &Z1);
}

/* exported: canonical in / canonical out, a has exactly m entries, m a valid domain size */
static fr_t *load_fr(const u64 *c, size_t n, size_t cap) {
    fr_t *a = (fr_t *)calloc(cap ? cap : 1, sizeof(fr_t));
    for (size_t i = 0; i < n; ++i) fr_from_canon(&a[i], c + 4 * i);
    return a;
}
static void store_fr(u64 *c, const fr_t *a, size_t n) {
    for (size_t i = 0; i < n; ++i) fr_to_canon(c + 4 * i, &a[i]);
}

/* evaluation_domain<Fr>::fft / inverse_fft on make_evaluation_domain-compatible size m.
 * n_in <= m entries are read (zero padded for the basic domain as the reference does; the step domain's
 * inverse_fft throws on short input: step_radix2_domain.hpp:112-114 -> returns -2). */
int orc_domain_fft(u64 m, int inverse, const u64 *in, u64 n_in, u64 *out) {
    orc_init();
    domain_t d;
    if (domain_init(&d, m)) return -1;
    if (n_in > m) return -2;
    if (inverse && d.kind == ORC_DOMAIN_STEP && n_in != m) return -2;
    fr_t *a = load_fr(in, n_in, m);
    int rc = inverse ? domain_ifft(&d, a) : domain_fft(&d, a);
    if (!rc) store_fr(out, a, m);
    free(a);
    return rc;
}
/* detail::basic_radix2_fft(a, omega) with an explicit root (canonical) */
int orc_fr_basic_radix2_fft(u64 n, const u64 *omega, u64 *a_inout) {
    orc_init();
    fr_t w;
    fr_from_canon(&w, omega);
    fr_t *a = load_fr(a_inout, n, n);
    int rc = basic_radix2_fft(a, n, &w);
    if (!rc) store_fr(a_inout, a, n);
    free(a);
    return rc;
}
int orc_multiply_by_coset(u64 n, const u64 *g, u64 *a_inout) {
    orc_init();
    fr_t gg;
    fr_from_canon(&gg, g);
    fr_t *a = load_fr(a_inout, n, n);
    multiply_by_coset(a, n, &gg);
    store_fr(a_inout, a, n);
    free(a);
    return 0;
}
int orc_divide_by_z_on_coset(u64 m, u64 *a_inout) {
    orc_init();
    domain_t d;
    if (domain_init(&d, m)) return -1;
    fr_t *a = load_fr(a_inout, m, m);
    divide_by_z_on_coset(&d, a);
    store_fr(a_inout, a, m);
    free(a);
    return 0;
}
int orc_unity_root(u64 n, u64 *out) {
    orc_init();
    fr_t w;
    if (unity_root(&w, n)) return -1;
    fr_to_canon(out, &w);
    return 0;
}
/* Fr helpers for tests */
int orc_fr_mul(const u64 *a, const u64 *b, u64 *out) {
    orc_init();
    fr_t x, y;
    fr_from_canon(&x, a);
    fr_from_canon(&y, b);
    fr_mul(&x, &x, &y);
    fr_to_canon(out, &x);
    return 0;
}
int orc_fq_mul(const u64 *a, const u64 *b, u64 *out) {
    orc_init();
    fq_t x, y;
    fq_from_canon(&x, a);
    fq_from_canon(&y, b);
    fq_mul(&x, &x, &y);
    fq_to_canon(out, &x);
    return 0;
}
int orc_fr_inv(const u64 *a, u64 *out) {
    orc_init();
    fr_t x;
    fr_from_canon(&x, a);
    fr_inv(&x, &x);
    fr_to_canon(out, &x);
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* multiexp exports                                                                            */
/* ------------------------------------------------------------------------------------------ */
static g1_aff *load_g1(const u64 *c, size_t n) {
    g1_aff *b = (g1_aff *)malloc((n ? n : 1) * sizeof(g1_aff));
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; ++i) g1_aff_from_canon(&b[i], c + 12 * i);
    return b;
}
static g2_aff *load_g2(const u64 *c, size_t n) {
    g2_aff *b = (g2_aff *)malloc((n ? n : 1) * sizeof(g2_aff));
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; ++i) g2_aff_from_canon(&b[i], c + 24 * i);
    return b;
}

/* algebra::multiexp<BDLO12>(bases, scalars, chunks); mixed != 0 -> multiexp_with_mixed_addition */
int orc_msm_g1(u64 n, const u64 *bases, const u64 *scalars, u64 chunks, int mixed, u64 *out) {
    orc_init();
    g1_aff *b = load_g1(bases, n);
    g1_jac r;
    if (mixed)
        g1_multiexp_mixed(&r, b, scalars, n, chunks);
    else
        g1_multiexp(&r, b, scalars, n, chunks);
    g1_jac_to_canon(out, &r);
    free(b);
    return 0;
}
int orc_msm_g2(u64 n, const u64 *bases, const u64 *scalars, u64 chunks, int mixed, u64 *out) {
    orc_init();
    g2_aff *b = load_g2(bases, n);
    g2_jac r;
    if (mixed)
        g2_multiexp_mixed(&r, b, scalars, n, chunks);
    else
        g2_multiexp(&r, b, scalars, n, chunks);
    g2_jac_to_canon(out, &r);
    free(b);
    return 0;
}
/* timing-only variant: bases already converted once (the reference holds Montgomery-form points in memory,
 * so conversion is not part of its multiexp).  handle = orc_g1_bases_load(...) */
void *orc_g1_bases_load(u64 n, const u64 *bases) {
    orc_init();
    return load_g1(bases, n);
}
void orc_bases_free(void *h) { free(h); }
int orc_msm_g1_loaded(u64 n, const void *h, const u64 *scalars, u64 chunks, int mixed, u64 *out) {
    g1_jac r;
    if (mixed)
        g1_multiexp_mixed(&r, (const g1_aff *)h, scalars, n, chunks);
    else
        g1_multiexp(&r, (const g1_aff *)h, scalars, n, chunks);
    g1_jac_to_canon(out, &r);
    return 0;
}

/* small group helpers (tests, generators of synthetic bases) */
int orc_g1_add(const u64 *a, const u64 *b, u64 *out) {
    orc_init();
    g1_jac p, q;
    g1_jac_from_canon(&p, a);
    g1_jac_from_canon(&q, b);
    g1_add(&p, &p, &q);
    g1_jac_to_canon(out, &p);
    return 0;
}
int orc_g2_add(const u64 *a, const u64 *b, u64 *out) {
    orc_init();
    g2_jac p, q;
    g2_jac_from_canon(&p, a);
    g2_jac_from_canon(&q, b);
    g2_add(&p, &p, &q);
    g2_jac_to_canon(out, &p);
    return 0;
}
int orc_g1_mul(const u64 *k, const u64 *a, u64 *out) {
    orc_init();
    g1_jac p;
    g1_jac_from_canon(&p, a);
    g1_mul(&p, k, &p);
    g1_jac_to_canon(out, &p);
    return 0;
}
int orc_g2_mul(const u64 *k, const u64 *a, u64 *out) {
    orc_init();
    g2_jac p;
    g2_jac_from_canon(&p, a);
    g2_mul(&p, k, &p);
    g2_jac_to_canon(out, &p);
    return 0;
}
/* Jacobian (X, Y, Z canonical, 18 limbs) -> canonical affine: the to_affine() the encoders call
 * (processing/curve_element.hpp:112-113) */
int orc_g1_jacobian_to_affine(const u64 *xyz, u64 *out) {
    orc_init();
    g1_jac p;
    fq_from_canon(&p.X, xyz);
    fq_from_canon(&p.Y, xyz + 6);
    fq_from_canon(&p.Z, xyz + 12);
    g1_jac_to_canon(out, &p);
    return 0;
}
/* out[i] = k_i * base for n 64-bit k_i: synthetic "random point" generation for benches/tests.
 * Uses a running table so it costs ~64 doublings once + <=64 adds per point. */
int orc_g1_fixed_base_mul64(const u64 *base, u64 n, const u64 *k, u64 *out) {
    orc_init();
    g1_jac tbl[64];
    g1_jac_from_canon(&tbl[0], base);
    for (int i = 1; i < 64; ++i) g1_dbl(&tbl[i], &tbl[i - 1]);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; ++i) {
        g1_jac acc;
        g1_set_inf(&acc);
        for (int b = 0; b < 64; ++b)
            if ((k[i] >> b) & 1) g1_add(&acc, &acc, &tbl[b]);
        g1_jac_to_canon(out + 12 * i, &acc);
    }
    return 0;
}
int orc_g2_fixed_base_mul64(const u64 *base, u64 n, const u64 *k, u64 *out) {
    orc_init();
    g2_jac tbl[64];
    g2_jac_from_canon(&tbl[0], base);
    for (int i = 1; i < 64; ++i) g2_dbl(&tbl[i], &tbl[i - 1]);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; ++i) {
        g2_jac acc;
        g2_set_inf(&acc);
        for (int b = 0; b < 64; ++b)
            if ((k[i] >> b) & 1) g2_add(&acc, &acc, &tbl[b]);
        g2_jac_to_canon(out + 24 * i, &acc);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* R1CS -> QAP witness map, Groth16 prover  [UPSTREAM-RECALL]                                  */
/* ------------------------------------------------------------------------------------------ */
/* Mirrors vsp_r1cs_t / vsp_proving_key_t in include/vsp_b200.h field for field (ctypes shares one Structure). */
typedef struct {
    u64 num_inputs;      /* primary_input_size */
    u64 num_aux;         /* auxiliary_input_size */
    u64 num_constraints;
    /* CSR per matrix; column 0 is the constant-one variable (r1cs_examples.hpp:104) */
    const u64 *a_row, *b_row, *c_row; /* num_constraints + 1 each */
    const uint32_t *a_col, *b_col, *c_col;
    const u64 *a_val, *b_val, *c_val; /* canonical Fr, 4 limbs per entry */
} orc_r1cs_t;

typedef struct {
    orc_r1cs_t cs;
    const u64 *alpha_g1, *beta_g1, *beta_g2, *delta_g1, *delta_g2;
    const u64 *A_query; /* num_variables + 1 G1 */
    u64 B_count;        /* sparse B_query (fast_proving_key.hpp:85-89): indices + (G2, G1) pairs */
    const uint32_t *B_idx;
    const u64 *B_g2, *B_g1;
    u64 H_count; /* m - 1 */
    const u64 *H_query;
    const u64 *L_query; /* num_aux G1 */
} orc_proving_key_t;

static void eval_rows(fr_t *dst, const u64 *row, const uint32_t *col, const u64 *val, size_t nrows, const fr_t *w,
                      int accumulate) {
    for (size_t i = 0; i < nrows; ++i) {
        fr_t acc;
        fr_zero(&acc);
        for (u64 k = row[i]; k < row[i + 1]; ++k) {
            fr_t c, t;
            fr_from_canon(&c, val + 4 * k);
            fr_mul(&t, &c, &w[col[k]]);
            fr_add(&acc, &acc, &t);
        }
        if (accumulate)
            fr_add(&dst[i], &dst[i], &acc);
        else
            dst[i] = acc;
    }
}

/* returns coefficients_for_H[0..m] (canonical, m+1 entries); *m_out = domain size.  h_out sized by caller via
 * orc_domain_plan(num_constraints + num_inputs + 1). */
int orc_witness_map(const orc_r1cs_t *cs, const u64 *primary, const u64 *aux, u64 *h_out, u64 *m_out) {
    orc_init();
    int kind;
    u64 m;
    if (orc_domain_plan(cs->num_constraints + cs->num_inputs + 1, &kind, &m)) return -1;
    domain_t d;
    if (domain_init(&d, m)) return -1;
    size_t nv = cs->num_inputs + cs->num_aux, nc = cs->num_constraints;
    fr_t *w = (fr_t *)malloc((nv + 1) * sizeof(fr_t)); /* const-padded assignment (1 || full) */
    w[0] = FR_R1;
    for (size_t i = 0; i < cs->num_inputs; ++i) fr_from_canon(&w[1 + i], primary + 4 * i);
    for (size_t i = 0; i < cs->num_aux; ++i) fr_from_canon(&w[1 + cs->num_inputs + i], aux + 4 * i);
    fr_t *aA = (fr_t *)calloc(m, sizeof(fr_t)), *aB = (fr_t *)calloc(m, sizeof(fr_t)),
         *aC = (fr_t *)calloc(m, sizeof(fr_t));
    for (size_t i = 0; i <= cs->num_inputs; ++i) aA[i + nc] = w[i];
    eval_rows(aA, cs->a_row, cs->a_col, cs->a_val, nc, w, 1);
    eval_rows(aB, cs->b_row, cs->b_col, cs->b_val, nc, w, 1);
    domain_ifft(&d, aA);
    domain_ifft(&d, aB);
    multiply_by_coset(aA, m, &FR_GEN);
    domain_fft(&d, aA);
    multiply_by_coset(aB, m, &FR_GEN);
    domain_fft(&d, aB);
    for (size_t i = 0; i < m; ++i) fr_mul(&aA[i], &aA[i], &aB[i]); /* H_tmp */
    eval_rows(aC, cs->c_row, cs->c_col, cs->c_val, nc, w, 1);
    domain_ifft(&d, aC);
    multiply_by_coset(aC, m, &FR_GEN);
    domain_fft(&d, aC);
    for (size_t i = 0; i < m; ++i) fr_sub(&aA[i], &aA[i], &aC[i]);
    divide_by_z_on_coset(&d, aA);
    domain_ifft(&d, aA);
    fr_t ginv;
    fr_inv(&ginv, &FR_GEN);
    multiply_by_coset(aA, m, &ginv);
    store_fr(h_out, aA, m);
    memset(h_out + 4 * m, 0, 32); /* coefficients_for_H[m] = 0 when d1 = d2 = 0 */
    *m_out = m;
    free(w);
    free(aA);
    free(aB);
    free(aC);
    return 0;
}

static void g1_mul_canon(g1_jac *r, const u64 *k, const g1_jac *p) { g1_mul(r, k, p); }

/* zk::prove<r1cs_gg_ppzksnark<...>> with explicit r, s.
 * out_proof = g_A (12 limbs) | g_B (24 limbs) | g_C (12 limbs), canonical affine.
 * r_saver / gamma_inverse_sum_s_g1 non-NULL: encrypted_input mode, g_C += r_saver * gamma_inverse_sum_s_g1. */
int orc_groth16_prove(const orc_proving_key_t *pk, const u64 *primary, const u64 *aux, const u64 *r, const u64 *s,
                      const u64 *r_saver, const u64 *gamma_inverse_sum_s_g1, u64 chunks, u64 *out_proof) {
    orc_init();
    const orc_r1cs_t *cs = &pk->cs;
    size_t ni = cs->num_inputs, na = cs->num_aux, nv = ni + na;
    int kind;
    u64 m;
    if (orc_domain_plan(cs->num_constraints + ni + 1, &kind, &m)) return -1;
    if (pk->H_count < m - 1) return -3;
    u64 *H = (u64 *)malloc((m + 1) * 32);
    if (orc_witness_map(cs, primary, aux, H, &m)) {
        free(H);
        return -1;
    }
    u64 *w = (u64 *)calloc((nv + 1) * 4, 8);
    w[0] = 1;
    memcpy(w + 4, primary, ni * 32);
    memcpy(w + 4 + 4 * ni, aux, na * 32);

    g1_jac evA, evBg1, evH, evL;
    g2_jac evBg2;
    {
        g1_aff *b = load_g1(pk->A_query, nv + 1);
        g1_multiexp_mixed(&evA, b, w, nv + 1, chunks);
        free(b);
    }
    {
        u64 *ws = (u64 *)malloc((pk->B_count ? pk->B_count : 1) * 32);
        for (size_t i = 0; i < pk->B_count; ++i) memcpy(ws + 4 * i, w + 4 * pk->B_idx[i], 32);
        g2_aff *b2 = load_g2(pk->B_g2, pk->B_count);
        g1_aff *b1 = load_g1(pk->B_g1, pk->B_count);
        g2_multiexp_mixed(&evBg2, b2, ws, pk->B_count, chunks);
        g1_multiexp_mixed(&evBg1, b1, ws, pk->B_count, chunks);
        free(b2);
        free(b1);
        free(ws);
    }
    {
        g1_aff *b = load_g1(pk->H_query, m - 1);
        g1_multiexp(&evH, b, H, m - 1, chunks);
        free(b);
    }
    {
        g1_aff *b = load_g1(pk->L_query, na);
        g1_multiexp_mixed(&evL, b, w + 4 * (ni + 1), na, chunks);
        free(b);
    }
    g1_jac alpha, beta1, delta1, t, gA, gB1, gC;
    g2_jac beta2, delta2, t2, gB2;
    g1_jac_from_canon(&alpha, pk->alpha_g1);
    g1_jac_from_canon(&beta1, pk->beta_g1);
    g1_jac_from_canon(&delta1, pk->delta_g1);
    g2_jac_from_canon(&beta2, pk->beta_g2);
    g2_jac_from_canon(&delta2, pk->delta_g2);
    /* g_A = alpha + A + r delta */
    g1_add(&gA, &alpha, &evA);
    g1_mul_canon(&t, r, &delta1);
    g1_add(&gA, &gA, &t);
    /* g1_B = beta + B.h + s delta ; g2_B = beta2 + B.g + s delta2 */
    g1_add(&gB1, &beta1, &evBg1);
    g1_mul_canon(&t, s, &delta1);
    g1_add(&gB1, &gB1, &t);
    g2_add(&gB2, &beta2, &evBg2);
    g2_mul(&t2, s, &delta2);
    g2_add(&gB2, &gB2, &t2);
    /* g_C = H + L + s g_A + r g1_B - (r s) delta */
    g1_add(&gC, &evH, &evL);
    g1_mul_canon(&t, s, &gA);
    g1_add(&gC, &gC, &t);
    g1_mul_canon(&t, r, &gB1);
    g1_add(&gC, &gC, &t);
    fr_t rr, ss;
    u64 rs[4];
    fr_from_canon(&rr, r);
    fr_from_canon(&ss, s);
    fr_mul(&rr, &rr, &ss);
    fr_to_canon(rs, &rr);
    g1_mul_canon(&t, rs, &delta1);
    g1_neg(&t, &t);
    g1_add(&gC, &gC, &t);
    if (r_saver && gamma_inverse_sum_s_g1) {
        g1_jac gis;
        g1_jac_from_canon(&gis, gamma_inverse_sum_s_g1);
        g1_mul_canon(&t, r_saver, &gis);
        g1_add(&gC, &gC, &t);
    }
    g1_jac_to_canon(out_proof, &gA);
    g2_jac_to_canon(out_proof + 12, &gB2);
    g1_jac_to_canon(out_proof + 36, &gC);
    free(H);
    free(w);
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* SAVER  [SRC] elgamal_verifiable.hpp                                                         */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    u64 msg_size;
    const u64 *delta_g1;               /* G1 */
    const u64 *delta_s_g1;             /* msg_size G1 */
    const u64 *t_g1;                   /* msg_size G1 */
    const u64 *t_g2;                   /* msg_size + 1 G2 (unused by encrypt) */
    const u64 *delta_sum_s_g1;         /* G1 */
    const u64 *gamma_inverse_sum_s_g1; /* G1 */
} orc_saver_pubkey_t;

/* encrypt_op::process ciphertext, :341-352.  ct_out: (msg_size + 2) G1. */
int orc_saver_ciphertext(const orc_saver_pubkey_t *pub, const u64 *gamma_abc_rest, const u64 *msg, const u64 *r,
                         u64 *ct_out) {
    orc_init();
    size_t n = pub->msg_size;
    g1_jac p, t, sum;
    g1_jac_from_canon(&p, pub->delta_g1);
    g1_mul(&t, r, &p);
    g1_jac_to_canon(ct_out, &t);
    g1_jac_from_canon(&p, pub->delta_sum_s_g1);
    g1_mul(&sum, r, &p);
    for (size_t i = 0; i < n; ++i) {
        g1_jac a, b;
        g1_jac_from_canon(&p, pub->delta_s_g1 + 12 * i);
        g1_mul(&a, r, &p);
        g1_jac_from_canon(&p, gamma_abc_rest + 12 * i);
        g1_mul(&b, msg + 4 * i, &p);
        g1_add(&a, &a, &b);
        g1_jac_to_canon(ct_out + 12 * (1 + i), &a);
        g1_jac_from_canon(&p, pub->t_g1 + 12 * i);
        g1_mul(&b, msg + 4 * i, &p);
        g1_add(&sum, &sum, &b);
    }
    g1_jac_to_canon(ct_out + 12 * (1 + n), &sum);
    return 0;
}

/* rerandomize_op::process, :632-667.  rnd = r | z1 | z2; proof layout as orc_groth16_prove. */
int orc_saver_rerandomize(const orc_saver_pubkey_t *pub, const u64 *delta_g2, const u64 *rnd, const u64 *ct_in,
                          const u64 *proof_in, u64 *ct_out, u64 *proof_out) {
    orc_init();
    size_t n = pub->msg_size;
    const u64 *r = rnd, *z1 = rnd + 4, *z2 = rnd + 8;
    u64 z1_inv[4];
    orc_fr_inv(z1, z1_inv);
    g1_jac p, c, t;
    g1_jac_from_canon(&p, pub->delta_g1);
    g1_mul(&t, r, &p);
    g1_jac_from_canon(&c, ct_in);
    g1_add(&c, &c, &t);
    g1_jac_to_canon(ct_out, &c);
    for (size_t i = 1; i <= n; ++i) {
        g1_jac_from_canon(&p, pub->delta_s_g1 + 12 * (i - 1));
        g1_mul(&t, r, &p);
        g1_jac_from_canon(&c, ct_in + 12 * i);
        g1_add(&c, &c, &t);
        g1_jac_to_canon(ct_out + 12 * i, &c);
    }
    g1_jac_from_canon(&p, pub->delta_sum_s_g1);
    g1_mul(&t, r, &p);
    g1_jac_from_canon(&c, ct_in + 12 * (n + 1));
    g1_add(&c, &c, &t);
    g1_jac_to_canon(ct_out + 12 * (n + 1), &c);

    g1_jac gA, gC, A2;
    g2_jac gB, B2, d2, t2;
    g1_jac_from_canon(&gA, proof_in);
    g2_jac_from_canon(&gB, proof_in + 12);
    g1_jac_from_canon(&gC, proof_in + 36);
    g1_mul(&A2, z1, &gA);
    g2_mul(&B2, z1_inv, &gB);
    g2_jac_from_canon(&d2, delta_g2);
    g2_mul(&t2, z2, &d2);
    g2_add(&B2, &B2, &t2);
    g1_mul(&t, z2, &A2);
    g1_add(&gC, &gC, &t);
    g1_jac_from_canon(&p, pub->gamma_inverse_sum_s_g1);
    g1_mul(&t, r, &p);
    g1_add(&gC, &gC, &t);
    g1_jac_to_canon(proof_out, &A2);
    g2_jac_to_canon(proof_out + 12, &B2);
    g1_jac_to_canon(proof_out + 36, &gC);
    return 0;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
