// fp.cuh — Montgomery prime-field arithmetic on 32-bit limbs for sm_100a (BLS12-381 Fq: 12 limbs, Fr: 8 limbs).
//
// Replaces the arithmetic of crypto3-algebra's `fields::bls12_fq<381>` / `bls12_fr<381>` value types (un-vendored
// submodule crypto3-algebra @ e4fa2c5e, /root/reference/.SUBMODULES.json:17; used through every `*`, `+`,
// `.inversed()`, `.pow()` in /root/reference/libs/math/include/nil/crypto3/math/domains/*.hpp).
//
// Multiplication keeps two partial accumulators (products landing on even / odd 32-bit columns) so every
// mad.lo.cc / madc.hi.cc pair is a 64-bit-aligned multiply-accumulate that ptxas maps to IMAD.WIDE.U32 with the
// carry in a predicate; the Montgomery shift by one limb is a role swap of the two accumulators.
#pragma once
#include <cstdint>

#include "bls12_381_constants.cuh"

namespace vsp {

#ifdef VSP_HOST_EMUL
// g++ build of the same templates (tests/host_emul): CUDA keywords and the few builtins used become plain C++
#define VSP_DEV inline
#define __device__
#define __host__
#define __noinline__
#define __forceinline__ inline
struct uint4 {
    uint32_t x, y, z, w;
};
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
inline int __clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
#else
#define VSP_DEV __device__ __forceinline__
#endif

#ifdef VSP_HOST_EMUL
// Host emulation of the PTX carry-chain instructions (tests/host_emul: the same templates are compiled with g++ so
// the limb schedules can be checked against the oracle without a GPU).  CC is the PTX condition-code carry flag.
namespace ptx {
static thread_local uint32_t CC = 0;
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)((uint64_t(a) * b) >> 32); }
inline uint32_t acc3(uint64_t x, uint64_t c, uint64_t cin, bool set) {
    const uint64_t t = x + c + cin;
    if (set) CC = (uint32_t)(t >> 32);
    return (uint32_t)t;
}
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return acc3(mul_lo(a, b), c, 0, true); }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return acc3(mul_lo(a, b), c, CC, true); }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return acc3(mul_hi(a, b), c, CC, true); }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return acc3(mul_hi(a, b), c, CC, false); }
inline uint32_t add_cc(uint32_t a, uint32_t b) { return acc3(a, b, 0, true); }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { return acc3(a, b, CC, true); }
inline uint32_t addc(uint32_t a, uint32_t b) { return acc3(a, b, CC, false); }
inline uint32_t sub3(uint64_t a, uint64_t b, uint64_t bin, bool set) {
    const uint64_t t = a - b - bin;
    if (set) CC = (uint32_t)(t >> 63);
    return (uint32_t)t;
}
inline uint32_t sub_cc(uint32_t a, uint32_t b) { return sub3(a, b, 0, true); }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { return sub3(a, b, CC, true); }
inline uint32_t subc(uint32_t a, uint32_t b) { return sub3(a, b, CC, false); }
}  // namespace ptx
#else
namespace ptx {
VSP_DEV uint32_t mul_lo(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t mul_hi(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
VSP_DEV uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
VSP_DEV uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
VSP_DEV uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
VSP_DEV uint32_t add_cc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t addc_cc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t addc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t sub_cc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t subc_cc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
VSP_DEV uint32_t subc(uint32_t a, uint32_t b) {
    uint32_t r;
    asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
}  // namespace ptx
#endif

struct FqCfg {
    static constexpr int N = VSP_FQ_LIMBS;
    static constexpr uint32_t INV = VSP_FQ_INV;
    __host__ __device__ static constexpr uint32_t mod(int i) { return VSP_FQ_MOD(i); }
    __host__ __device__ static constexpr uint32_t r1(int i) { return VSP_FQ_R1(i); }
    __host__ __device__ static constexpr uint32_t r2(int i) { return VSP_FQ_R2(i); }
    __host__ __device__ static constexpr uint32_t p2(int i) { return VSP_FQ_P2(i); }
};
struct FrCfg {
    static constexpr int N = VSP_FR_LIMBS;
    static constexpr uint32_t INV = VSP_FR_INV;
    __host__ __device__ static constexpr uint32_t mod(int i) { return VSP_FR_MOD(i); }
    __host__ __device__ static constexpr uint32_t r1(int i) { return VSP_FR_R1(i); }
    __host__ __device__ static constexpr uint32_t r2(int i) { return VSP_FR_R2(i); }
    __host__ __device__ static constexpr uint32_t p2(int i) { return VSP_FR_P2(i); }
};

template <class C>
struct alignas(16) Fp {
    static constexpr int N = C::N;
    uint32_t l[C::N];

    VSP_DEV static Fp zero() {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) r.l[i] = 0;
        return r;
    }
    // Montgomery form of 1
    VSP_DEV static Fp one() {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) r.l[i] = C::r1(i);
        return r;
    }
    VSP_DEV static Fp r2() {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) r.l[i] = C::r2(i);
        return r;
    }
    VSP_DEV bool is_zero() const {
        uint32_t x = 0;
#pragma unroll
        for (int i = 0; i < N; ++i) x |= l[i];
        return x == 0;
    }
    VSP_DEV bool operator==(const Fp& o) const {
        uint32_t x = 0;
#pragma unroll
        for (int i = 0; i < N; ++i) x |= l[i] ^ o.l[i];
        return x == 0;
    }
    VSP_DEV bool operator!=(const Fp& o) const { return !(*this == o); }
};

// r = (a >= p) ? a - p : a      (a < 2p)
template <class C>
VSP_DEV void fp_reduce_once(Fp<C>& a) {
    constexpr int N = C::N;
    uint32_t t[N];
    t[0] = ptx::sub_cc(a.l[0], C::mod(0));
#pragma unroll
    for (int i = 1; i < N; ++i) t[i] = ptx::subc_cc(a.l[i], C::mod(i));
    uint32_t borrow = ptx::subc(0u, 0u);  // 0xffffffff if a < p
#pragma unroll
    for (int i = 0; i < N; ++i) a.l[i] = borrow ? a.l[i] : t[i];
}

template <class C>
VSP_DEV Fp<C> fp_add(const Fp<C>& a, const Fp<C>& b) {
    constexpr int N = C::N;
    Fp<C> r;
    r.l[0] = ptx::add_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < N - 1; ++i) r.l[i] = ptx::addc_cc(a.l[i], b.l[i]);
    r.l[N - 1] = ptx::addc(a.l[N - 1], b.l[N - 1]);  // 2p < 2^(32N): no carry out
    fp_reduce_once(r);
    return r;
}

template <class C>
VSP_DEV Fp<C> fp_sub(const Fp<C>& a, const Fp<C>& b) {
    constexpr int N = C::N;
    Fp<C> r;
    r.l[0] = ptx::sub_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < N; ++i) r.l[i] = ptx::subc_cc(a.l[i], b.l[i]);
    uint32_t mask = ptx::subc(0u, 0u);  // all ones on borrow
    r.l[0] = ptx::add_cc(r.l[0], C::mod(0) & mask);
#pragma unroll
    for (int i = 1; i < N - 1; ++i) r.l[i] = ptx::addc_cc(r.l[i], C::mod(i) & mask);
    r.l[N - 1] = ptx::addc(r.l[N - 1], C::mod(N - 1) & mask);
    return r;
}

// a + b without the conditional subtraction (< 2p < 2^(32N)): an operand for fp_mul_wide
template <class C>
VSP_DEV Fp<C> fp_add_raw(const Fp<C>& a, const Fp<C>& b) {
    constexpr int N = C::N;
    Fp<C> r;
    r.l[0] = ptx::add_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < N - 1; ++i) r.l[i] = ptx::addc_cc(a.l[i], b.l[i]);
    r.l[N - 1] = ptx::addc(a.l[N - 1], b.l[N - 1]);
    return r;
}

template <class C>
VSP_DEV Fp<C> fp_neg(const Fp<C>& a) {
    constexpr int N = C::N;
    Fp<C> r;
    r.l[0] = ptx::sub_cc(C::mod(0), a.l[0]);
#pragma unroll
    for (int i = 1; i < N - 1; ++i) r.l[i] = ptx::subc_cc(C::mod(i), a.l[i]);
    r.l[N - 1] = ptx::subc(C::mod(N - 1), a.l[N - 1]);
    const bool z = a.is_zero();
#pragma unroll
    for (int i = 0; i < N; ++i) r.l[i] = z ? 0u : r.l[i];
    return r;
}

template <class C>
VSP_DEV Fp<C> fp_dbl(const Fp<C>& a) {
    return fp_add(a, a);
}

// One Montgomery reduction column: m = E[0] * INV;  O += m * p_odd;  E += m * p_even  (E[0] becomes 0).
// CARRY_IN: the O chain starts from the carry flag left by the caller's last .cc instruction (a carry into column 1 of
// the frame, which is where O[0] sits), so a frame shift needs no ripple of its own.
template <class C, bool CARRY_IN = false>
VSP_DEV void mont_reduce_column(uint32_t (&E)[C::N], uint32_t (&O)[C::N]) {
    constexpr int N = C::N;
    const uint32_t m = E[0] * C::INV;
    O[0] = CARRY_IN ? ptx::madc_lo_cc(C::mod(1), m, O[0]) : ptx::mad_lo_cc(C::mod(1), m, O[0]);
    O[1] = ptx::madc_hi_cc(C::mod(1), m, O[1]);
#pragma unroll
    for (int k = 2; k < N - 2; k += 2) {
        O[k] = ptx::madc_lo_cc(C::mod(k + 1), m, O[k]);
        O[k + 1] = ptx::madc_hi_cc(C::mod(k + 1), m, O[k + 1]);
    }
    O[N - 2] = ptx::madc_lo_cc(C::mod(N - 1), m, O[N - 2]);
    O[N - 1] = ptx::madc_hi(C::mod(N - 1), m, O[N - 1]);
    E[0] = ptx::mad_lo_cc(C::mod(0), m, E[0]);
    E[1] = ptx::madc_hi_cc(C::mod(0), m, E[1]);
#pragma unroll
    for (int k = 2; k < N; k += 2) {
        E[k] = ptx::madc_lo_cc(C::mod(k), m, E[k]);
        E[k + 1] = ptx::madc_hi_cc(C::mod(k), m, E[k + 1]);
    }
    O[N - 1] = ptx::addc(O[N - 1], 0u);
}

// Montgomery product a * b / 2^(32N) mod p, inputs and output fully reduced.
template <class C>
VSP_DEV Fp<C> fp_mul(const Fp<C>& a, const Fp<C>& b) {
    constexpr int N = C::N;
    static_assert(N % 2 == 0, "even limb count expected");
    uint32_t E[N], O[N];
    {
        const uint32_t bi = b.l[0];
#pragma unroll
        for (int k = 0; k < N; k += 2) {
            E[k] = ptx::mul_lo(a.l[k], bi);
            E[k + 1] = ptx::mul_hi(a.l[k], bi);
            O[k] = ptx::mul_lo(a.l[k + 1], bi);
            O[k + 1] = ptx::mul_hi(a.l[k + 1], bi);
        }
        mont_reduce_column<C>(E, O);
    }
#pragma unroll
    for (int i = 1; i < N; ++i) {
        const uint32_t bi = b.l[i];
        uint32_t nO[N];
        // shift by one limb: new even accumulator = O (+ E[1] at column 0), new odd accumulator = E >> 64
        O[0] = ptx::add_cc(O[0], E[1]);
#pragma unroll
        for (int k = 0; k < N - 2; k += 2) {
            nO[k] = ptx::madc_lo_cc(a.l[k + 1], bi, E[k + 2]);
            nO[k + 1] = ptx::madc_hi_cc(a.l[k + 1], bi, E[k + 3]);
        }
        nO[N - 2] = ptx::madc_lo_cc(a.l[N - 1], bi, 0u);
        nO[N - 1] = ptx::madc_hi(a.l[N - 1], bi, 0u);
        O[0] = ptx::mad_lo_cc(a.l[0], bi, O[0]);
        O[1] = ptx::madc_hi_cc(a.l[0], bi, O[1]);
#pragma unroll
        for (int k = 2; k < N; k += 2) {
            O[k] = ptx::madc_lo_cc(a.l[k], bi, O[k]);
            O[k + 1] = ptx::madc_hi_cc(a.l[k], bi, O[k + 1]);
        }
        nO[N - 1] = ptx::addc(nO[N - 1], 0u);
#pragma unroll
        for (int k = 0; k < N; ++k) {
            E[k] = O[k];
            O[k] = nO[k];
        }
        mont_reduce_column<C>(E, O);
    }
    Fp<C> r;
    r.l[0] = ptx::add_cc(O[0], E[1]);
#pragma unroll
    for (int k = 1; k < N - 1; ++k) r.l[k] = ptx::addc_cc(O[k], E[k + 1]);
    r.l[N - 1] = ptx::addc(O[N - 1], 0u);
    fp_reduce_once(r);
    return r;
}

// Montgomery reduction of a double-width value: T / 2^(32N) mod p for T < p 2^(32N), fully reduced.  Column by column
// on the sliding (E, O) frame of mont_reduce_column: frame value = E + O 2^32 over limbs [i, i + N]; the frame's shift is
// a role swap, the carry of O[0] + E[1] rides into the next column's multiply chain, and limb N + i of T enters the
// frame's top as row i leaves.  N^2 + N multiply-adds.
template <class C>
VSP_DEV Fp<C> fp_redc(const uint32_t (&T)[2 * C::N]) {
    constexpr int N = C::N;
    uint32_t E[N], O[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
        E[k] = T[k];
        O[k] = 0u;
    }
    mont_reduce_column<C>(E, O);  // E[0] = 0 now
#pragma unroll
    for (int i = 0; i < N - 1; ++i) {
        uint32_t nO[N];
#pragma unroll
        for (int k = 0; k < N - 2; ++k) nO[k] = E[k + 2];
        nO[N - 2] = T[N + i];
        nO[N - 1] = 0u;
        const uint32_t e0 = ptx::add_cc(O[0], E[1]);
#pragma unroll
        for (int k = 1; k < N; ++k) E[k] = O[k];
        E[0] = e0;
#pragma unroll
        for (int k = 0; k < N; ++k) O[k] = nO[k];
        mont_reduce_column<C, true>(E, O);
    }
    Fp<C> r;
    r.l[0] = ptx::add_cc(O[0], E[1]);
#pragma unroll
    for (int k = 1; k < N - 1; ++k) r.l[k] = ptx::addc_cc(O[k], E[k + 1]);
    r.l[N - 1] = ptx::addc(O[N - 1], T[2 * N - 1]);
    fp_reduce_once(r);
    return r;
}

// a * b as a 2N-limb integer, no reduction (N^2 multiply-adds).  Operands may be any N-limb values (e.g. an unreduced
// sum a0 + a1 < 2p).  Same two-accumulator discipline as fp_mul: a product landing on an even column goes to the pair
// Ev[c], Ev[c + 1], on an odd column to Od[c - 1], Od[c]; rows in increasing order, so the limb above a row's chain has
// only ever held carries when the chain's last carry is added to it.
template <class C>
VSP_DEV void fp_mul_wide(const Fp<C>& a, const Fp<C>& b, uint32_t (&T)[2 * C::N]) {
    constexpr int N = C::N;
    uint32_t Ev[2 * N], Od[2 * N];
#pragma unroll
    for (int k = 0; k < 2 * N; ++k) Ev[k] = Od[k] = 0u;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const uint32_t bi = b.l[i];
        // columns i + j with j of i's parity are even when i is even: they go to Ev if (i + j) is even
        {
            const int j0 = i & 1;  // (i + j) even
#pragma unroll
            for (int j = j0; j < N; j += 2) {
                const int c = i + j;
                Ev[c] = (j == j0) ? ptx::mad_lo_cc(a.l[j], bi, Ev[c]) : ptx::madc_lo_cc(a.l[j], bi, Ev[c]);
                Ev[c + 1] = ptx::madc_hi_cc(a.l[j], bi, Ev[c + 1]);
            }
            const int top = i + (N - 2 + j0) + 2;  // limb above the chain
            if (top < 2 * N) Ev[top] = ptx::addc(Ev[top], 0u);
        }
        {
            const int j0 = 1 - (i & 1);  // (i + j) odd
#pragma unroll
            for (int j = j0; j < N; j += 2) {
                const int c = i + j - 1;
                Od[c] = (j == j0) ? ptx::mad_lo_cc(a.l[j], bi, Od[c]) : ptx::madc_lo_cc(a.l[j], bi, Od[c]);
                Od[c + 1] = ptx::madc_hi_cc(a.l[j], bi, Od[c + 1]);
            }
            const int top = i + (N - 2 + j0) - 1 + 2;
            if (top < 2 * N) Od[top] = ptx::addc(Od[top], 0u);
        }
    }
    T[0] = Ev[0];
    T[1] = ptx::add_cc(Ev[1], Od[0]);
#pragma unroll
    for (int k = 2; k < 2 * N - 1; ++k) T[k] = ptx::addc_cc(Ev[k], Od[k - 1]);
    T[2 * N - 1] = ptx::addc(Ev[2 * N - 1], Od[2 * N - 2]);
}
// double-width helpers: T += U, T -= U (no overflow / no borrow by the callers' bounds), T += p^2
template <int M>
VSP_DEV void wide_add(uint32_t (&T)[M], const uint32_t (&U)[M]) {
    T[0] = ptx::add_cc(T[0], U[0]);
#pragma unroll
    for (int k = 1; k < M - 1; ++k) T[k] = ptx::addc_cc(T[k], U[k]);
    T[M - 1] = ptx::addc(T[M - 1], U[M - 1]);
}
template <int M>
VSP_DEV void wide_sub(uint32_t (&T)[M], const uint32_t (&U)[M]) {
    T[0] = ptx::sub_cc(T[0], U[0]);
#pragma unroll
    for (int k = 1; k < M - 1; ++k) T[k] = ptx::subc_cc(T[k], U[k]);
    T[M - 1] = ptx::subc(T[M - 1], U[M - 1]);
}
template <class C>
VSP_DEV void wide_add_p2(uint32_t (&T)[2 * C::N]) {
    constexpr int M = 2 * C::N;
    T[0] = ptx::add_cc(T[0], C::p2(0));
#pragma unroll
    for (int k = 1; k < M - 1; ++k) T[k] = ptx::addc_cc(T[k], C::p2(k));
    T[M - 1] = ptx::addc(T[M - 1], C::p2(M - 1));
}
// a b - c d mod p on ONE reduction: (a b - c d + p^2) / 2^(32N), all four operands fully reduced
template <class C>
VSP_DEV Fp<C> fp_mul_sub_mul(const Fp<C>& a, const Fp<C>& b, const Fp<C>& c, const Fp<C>& d) {
    uint32_t T[2 * C::N], U[2 * C::N];
    fp_mul_wide<C>(a, b, T);
    fp_mul_wide<C>(c, d, U);
    wide_add_p2<C>(T);
    wide_sub(T, U);
    return fp_redc<C>(T);
}

// Montgomery square a^2 / 2^(32N) mod p.  The N(N-1)/2 cross products a_i a_j (i < j) are formed once and doubled, so a
// square costs N(N-1)/2 + N + N^2 + N multiply-adds (Fq: 234, Fr: 108) against the product's 2N^2 + N (300 / 136).
// Same two-accumulator discipline as fp_mul so every lo/hi pair is one IMAD.WIDE: a product landing on an even 32-bit
// column goes to Ev (pair Ev[c], Ev[c+1]), on an odd column to Od (Od[k] is column k + 1), rows i in increasing order so
// the limb above a row's chain has only ever held carries.  Then T = 2 (Ev + Od 2^32) + sum a_i^2 2^(64 i) as a flat
// 2N-limb integer, reduced column by column on the sliding (E, O) frame of mont_reduce_column; the frame's shift is a
// role swap plus one carry ripple, and limb N + i of T enters the frame's top as row i leaves.
template <class C>
VSP_DEV Fp<C> fp_sqr(const Fp<C>& a) {
    constexpr int N = C::N;
    static_assert(N % 2 == 0 && N >= 4, "even limb count expected");
    uint32_t Ev[2 * N], Od[2 * N];
#pragma unroll
    for (int k = 0; k < 2 * N; ++k) Ev[k] = Od[k] = 0u;
#pragma unroll
    for (int i = 0; i < N - 1; ++i) {
        // partners j = i + 1, i + 3, ...: odd columns i + j -> Od[i + j - 1], Od[i + j]
        {
            const int jl = ((N - 1 - (i + 1)) & 1) ? N - 2 : N - 1;  // last partner of this parity
#pragma unroll
            for (int j = i + 1; j < N; j += 2) {
                const int c = i + j - 1;
                Od[c] = (j == i + 1) ? ptx::mad_lo_cc(a.l[i], a.l[j], Od[c]) : ptx::madc_lo_cc(a.l[i], a.l[j], Od[c]);
                Od[c + 1] = ptx::madc_hi_cc(a.l[i], a.l[j], Od[c + 1]);
            }
            Od[i + jl + 1] = ptx::addc(Od[i + jl + 1], 0u);
        }
        // partners j = i + 2, i + 4, ...: even columns i + j -> Ev[i + j], Ev[i + j + 1]
        if (i + 2 < N) {
            const int jl = ((N - 1 - (i + 2)) & 1) ? N - 2 : N - 1;
#pragma unroll
            for (int j = i + 2; j < N; j += 2) {
                const int c = i + j;
                Ev[c] = (j == i + 2) ? ptx::mad_lo_cc(a.l[i], a.l[j], Ev[c]) : ptx::madc_lo_cc(a.l[i], a.l[j], Ev[c]);
                Ev[c + 1] = ptx::madc_hi_cc(a.l[i], a.l[j], Ev[c + 1]);
            }
            Ev[i + jl + 2] = ptx::addc(Ev[i + jl + 2], 0u);
        }
    }
    // T = Ev + Od 2^32 (the cross sum is < 2^(64N - 1): no carry out), doubled, plus the diagonal
    uint32_t T[2 * N];
    T[0] = Ev[0];
    T[1] = ptx::add_cc(Ev[1], Od[0]);
#pragma unroll
    for (int k = 2; k < 2 * N - 1; ++k) T[k] = ptx::addc_cc(Ev[k], Od[k - 1]);
    T[2 * N - 1] = ptx::addc(Ev[2 * N - 1], Od[2 * N - 2]);
#pragma unroll
    for (int k = 2 * N - 1; k > 0; --k) T[k] = (T[k] << 1) | (T[k - 1] >> 31);
    T[0] <<= 1;
    T[0] = ptx::mad_lo_cc(a.l[0], a.l[0], T[0]);
    T[1] = ptx::madc_hi_cc(a.l[0], a.l[0], T[1]);
#pragma unroll
    for (int i = 1; i < N - 1; ++i) {
        T[2 * i] = ptx::madc_lo_cc(a.l[i], a.l[i], T[2 * i]);
        T[2 * i + 1] = ptx::madc_hi_cc(a.l[i], a.l[i], T[2 * i + 1]);
    }
    T[2 * N - 2] = ptx::madc_lo_cc(a.l[N - 1], a.l[N - 1], T[2 * N - 2]);
    T[2 * N - 1] = ptx::madc_hi(a.l[N - 1], a.l[N - 1], T[2 * N - 1]);
    return fp_redc<C>(T);
}

template <class C>
VSP_DEV Fp<C> fp_to_mont(const Fp<C>& a) {
    return fp_mul(a, Fp<C>::r2());
}
template <class C>
VSP_DEV Fp<C> fp_from_mont(const Fp<C>& a) {
    Fp<C> one = Fp<C>::zero();
    one.l[0] = 1;
    return fp_mul(a, one);
}

// a^(p-2) by plain square-and-multiply over the constant exponent bits (not on the hot path: batch inversion
// amortises it).
template <class C>
__device__ __noinline__ Fp<C> fp_inv(const Fp<C>& a) {
    constexpr int N = C::N;
    Fp<C> acc = Fp<C>::one();
    Fp<C> base = a;
    uint32_t borrow = 2;  // exponent = p - 2
    for (int i = 0; i < N; ++i) {
        uint32_t limb = C::mod(0);
        // runtime-indexed modulus limb without a table: select
#pragma unroll
        for (int k = 0; k < N; ++k) limb = (i == k) ? C::mod(k) : limb;
        uint32_t e = limb - borrow;
        borrow = (limb < borrow) ? 1u : 0u;
        for (int bit = 0; bit < 32; ++bit) {
            if ((e >> bit) & 1u) acc = fp_mul(acc, base);
            base = fp_sqr(base);
        }
    }
    return acc;
}

// The same inverse by the binary extended Euclid (right shifts and subtractions on the limbs, ~45k plain integer
// instructions against ~190k multiply-adds for the power): the branches follow the data, so it only pays where ONE
// thread of a warp is working — the serial tails of a single multiexp or proof — and the uniform power above stays
// the choice wherever the lanes of a warp invert different values.  In: a R (Montgomery); out: a^-1 R; 0 -> 0.
namespace gcd_detail {
template <int N>
VSP_DEV void shr1(uint32_t (&u)[N], uint32_t top) {
#pragma unroll
    for (int k = 0; k < N - 1; ++k) u[k] = (u[k] >> 1) | (u[k + 1] << 31);
    u[N - 1] = (u[N - 1] >> 1) | (top << 31);
}
template <int N>
VSP_DEV uint32_t sub_to(uint32_t (&u)[N], const uint32_t (&v)[N]) {  // u -= v, returns the borrow
    uint32_t r[N];
    r[0] = ptx::sub_cc(u[0], v[0]);
#pragma unroll
    for (int k = 1; k < N; ++k) r[k] = ptx::subc_cc(u[k], v[k]);
    const uint32_t borrow = ptx::subc(0u, 0u);
#pragma unroll
    for (int k = 0; k < N; ++k) u[k] = r[k];
    return borrow;  // 0 or 0xffffffff
}
template <class C>
VSP_DEV uint32_t add_mod_to(uint32_t (&x)[C::N]) {  // x += p, returns the carry
    x[0] = ptx::add_cc(x[0], C::mod(0));
#pragma unroll
    for (int k = 1; k < C::N; ++k) x[k] = ptx::addc_cc(x[k], C::mod(k));
    return ptx::addc(0u, 0u);
}
template <int N>
VSP_DEV bool geq(const uint32_t (&u)[N], const uint32_t (&v)[N]) {
    uint32_t t = ptx::sub_cc(u[0], v[0]);
#pragma unroll
    for (int k = 1; k < N; ++k) t = ptx::subc_cc(u[k], v[k]);
    (void)t;
    return ptx::subc(0u, 0u) == 0u;
}
template <int N>
VSP_DEV bool is_one(const uint32_t (&u)[N]) {
    uint32_t acc = u[0] ^ 1u;
#pragma unroll
    for (int k = 1; k < N; ++k) acc |= u[k];
    return acc == 0u;
}
}  // namespace gcd_detail

template <class C>
#ifdef __CUDACC__
__device__ __noinline__
#else
inline
#endif
Fp<C> fp_inv_gcd(const Fp<C>& a) {
    using namespace gcd_detail;
    constexpr int N = C::N;
    if (a.is_zero()) return a;
    // invariants: u = a x1, v = a x2 (mod p); u, v > 0; gcd(u, v) = 1
    uint32_t u[N], v[N], x1[N], x2[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
        u[k] = a.l[k];
        v[k] = C::mod(k);
        x1[k] = k == 0 ? 1u : 0u;
        x2[k] = 0u;
    }
    while (!is_one<N>(u) && !is_one<N>(v)) {
        if (!(u[0] & 1u)) {
            shr1<N>(u, 0u);
            const uint32_t c = (x1[0] & 1u) ? add_mod_to<C>(x1) : 0u;
            shr1<N>(x1, c);
        } else if (!(v[0] & 1u)) {
            shr1<N>(v, 0u);
            const uint32_t c = (x2[0] & 1u) ? add_mod_to<C>(x2) : 0u;
            shr1<N>(x2, c);
        } else if (geq<N>(u, v)) {
            sub_to<N>(u, v);
            if (sub_to<N>(x1, x2)) add_mod_to<C>(x1);
        } else {
            sub_to<N>(v, u);
            if (sub_to<N>(x2, x1)) add_mod_to<C>(x2);
        }
    }
    Fp<C> r;
    const bool from_u = is_one<N>(u);
#pragma unroll
    for (int k = 0; k < N; ++k) r.l[k] = from_u ? x1[k] : x2[k];
    // r = (a R)^-1 = a^-1 R^-1; one Montgomery product with R^3 puts it back in Montgomery form
    const Fp<C> r2 = Fp<C>::r2();
    return fp_mul(r, fp_mul(r2, r2));
}

// The inverse by Bernstein-Yang division steps ("safegcd", eprint 2019/266), constant time: the same instruction stream
// for every input, so the 32 lanes of a warp invert 32 different values in lock step.  Layout after libsecp256k1's
// modinv32 (signed 30-bit limbs; 30 division steps at a time on the low words, collected in a 2x2 transition matrix that
// is then applied to the full-width f, g and — modulo p — to d, e).  Number of steps: the paper's bound for the delta = 1
// start, floor((49 d + 57) / 17) for a d-bit modulus (d >= 46): 1,101 for Fq, 746 for Fr; groups of 30.
// Cost for Fq: 37 groups x (130 signed 32x32->64 products + ~700 plain integer instructions) — about 4.8k heavy-pipe
// multiplies against ~146k for the Fermat power, which is what makes a shared inversion per few dozen affine additions
// pay (msm_impl.cuh).  In: a R (Montgomery form); out: a^-1 R; 0 -> 0.
namespace safegcd {
constexpr int32_t M30 = 0x3fffffff;
struct Trans {
    int32_t u, v, q, r;
};
// 30 division steps on the low 30+ bits; eta = -delta
VSP_DEV int32_t divsteps_30(int32_t eta, uint32_t f0, uint32_t g0, Trans& t) {
    uint32_t u = 1, v = 0, q = 0, r = 1, f = f0, g = g0;
#pragma unroll 6
    for (int i = 0; i < 30; ++i) {
        uint32_t c1 = (uint32_t)(eta >> 31);  // delta > 0
        const uint32_t c2 = 0u - (g & 1u);    // g odd
        const uint32_t x = (f ^ c1) - c1, y = (u ^ c1) - c1, z = (v ^ c1) - c1;
        g += x & c2;
        q += y & c2;
        r += z & c2;
        c1 &= c2;
        eta = (int32_t)(((uint32_t)eta ^ c1) - (c1 + 1u));
        f += g & c1;
        u += q & c1;
        v += r & c1;
        g >>= 1;
        u <<= 1;
        v <<= 1;
    }
    t.u = (int32_t)u;
    t.v = (int32_t)v;
    t.q = (int32_t)q;
    t.r = (int32_t)r;
    return eta;
}
// limb j (30 bits) of the modulus
template <class C>
__host__ __device__ constexpr uint32_t mod30(int j) {
    return (uint32_t)(((((30 * j) / 32 + 1 < C::N ? (uint64_t)C::mod((30 * j) / 32 + 1) : 0ull) << 32 |
                        ((30 * j) / 32 < C::N ? (uint64_t)C::mod((30 * j) / 32) : 0ull)) >>
                       ((30 * j) % 32)) &
                      0x3fffffffull);
}
}  // namespace safegcd

template <class C>
#ifdef __CUDACC__
__device__ __noinline__
#else
inline
#endif
Fp<C> fp_inv_safegcd(const Fp<C>& a) {
    using namespace safegcd;
    constexpr int N = C::N;
    constexpr int L = (32 * N + 29) / 30 + ((32 * N) % 30 == 0 ? 1 : 0);  // 30-bit limbs, at least 2 spare bits: Fq 13, Fr 9
    static_assert(30 * L >= 32 * N + 2, "headroom for values in (-2p, p)");
    constexpr int BITS = 32 * N - (C::mod(N - 1) >> 31 ? 0 : (C::mod(N - 1) >> 30 ? 1 : (C::mod(N - 1) >> 29 ? 2 : 3)));
    static_assert((C::mod(N - 1) >> 28) != 0, "modulus top limb expected to have at least 29 bits");
    constexpr int STEPS = (49 * BITS + 57) / 17;
    constexpr int GROUPS = (STEPS + 29) / 30;
    constexpr uint32_t PINV30 = (0u - C::INV) & 0x3fffffffu;  // p^-1 mod 2^30 (C::INV = -p^-1 mod 2^32)
    int32_t d[L], e[L], f[L], g[L];
#pragma unroll
    for (int j = 0; j < L; ++j) {
        const int k = (30 * j) / 32, s = (30 * j) % 32;
        const uint64_t two = ((k + 1 < N ? (uint64_t)a.l[k + 1] : 0ull) << 32) | (k < N ? (uint64_t)a.l[k] : 0ull);
        g[j] = (int32_t)((two >> s) & 0x3fffffffull);
        f[j] = (int32_t)mod30<C>(j);
        d[j] = 0;
        e[j] = j == 0 ? 1 : 0;
    }
    int32_t eta = -1;
#pragma unroll 1
    for (int it = 0; it < GROUPS; ++it) {
        Trans t;
        eta = divsteps_30(eta, (uint32_t)f[0], (uint32_t)g[0], t);
        // [d, e] <- t [d, e] / 2^30 mod p, both kept in (-2p, p)
        {
            const int32_t sd = d[L - 1] >> 31, se = e[L - 1] >> 31;
            int32_t md = (t.u & sd) + (t.v & se), me = (t.q & sd) + (t.r & se);
            int64_t cd = (int64_t)t.u * d[0] + (int64_t)t.v * e[0];
            int64_t ce = (int64_t)t.q * d[0] + (int64_t)t.r * e[0];
            md -= (int32_t)((PINV30 * (uint32_t)cd + (uint32_t)md) & (uint32_t)M30);
            me -= (int32_t)((PINV30 * (uint32_t)ce + (uint32_t)me) & (uint32_t)M30);
            cd += (int64_t)(int32_t)mod30<C>(0) * md;
            ce += (int64_t)(int32_t)mod30<C>(0) * me;
            cd >>= 30;
            ce >>= 30;
#pragma unroll
            for (int j = 1; j < L; ++j) {
                cd += (int64_t)t.u * d[j] + (int64_t)t.v * e[j];
                ce += (int64_t)t.q * d[j] + (int64_t)t.r * e[j];
                cd += (int64_t)(int32_t)mod30<C>(j) * md;
                ce += (int64_t)(int32_t)mod30<C>(j) * me;
                d[j - 1] = (int32_t)cd & M30;
                e[j - 1] = (int32_t)ce & M30;
                cd >>= 30;
                ce >>= 30;
            }
            d[L - 1] = (int32_t)cd;
            e[L - 1] = (int32_t)ce;
        }
        // [f, g] <- t [f, g] / 2^30 (exact)
        {
            int64_t cf = (int64_t)t.u * f[0] + (int64_t)t.v * g[0];
            int64_t cg = (int64_t)t.q * f[0] + (int64_t)t.r * g[0];
            cf >>= 30;
            cg >>= 30;
#pragma unroll
            for (int j = 1; j < L; ++j) {
                cf += (int64_t)t.u * f[j] + (int64_t)t.v * g[j];
                cg += (int64_t)t.q * f[j] + (int64_t)t.r * g[j];
                f[j - 1] = (int32_t)cf & M30;
                g[j - 1] = (int32_t)cg & M30;
                cf >>= 30;
                cg >>= 30;
            }
            f[L - 1] = (int32_t)cf;
            g[L - 1] = (int32_t)cg;
        }
    }
    // g = 0 and f = +-1 now: d = +-(a R)^-1.  Bring d from (-2p, p) into [0, p), negating if f = -1.
    {
        int32_t cond = d[L - 1] >> 31;
#pragma unroll
        for (int j = 0; j < L; ++j) d[j] += (int32_t)mod30<C>(j) & cond;
        cond = f[L - 1] >> 31;
#pragma unroll
        for (int j = 0; j < L; ++j) d[j] = (d[j] ^ cond) - cond;
#pragma unroll
        for (int j = 0; j < L - 1; ++j) {
            d[j + 1] += d[j] >> 30;
            d[j] &= M30;
        }
        cond = d[L - 1] >> 31;
#pragma unroll
        for (int j = 0; j < L; ++j) d[j] += (int32_t)mod30<C>(j) & cond;
#pragma unroll
        for (int j = 0; j < L - 1; ++j) {
            d[j + 1] += d[j] >> 30;
            d[j] &= M30;
        }
    }
    Fp<C> r;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        // bits [32 k, 32 k + 32) out of the 30-bit limbs
        const int j = (32 * k) / 30, s = (32 * k) % 30;
        uint64_t w = (uint64_t)(uint32_t)d[j] >> s;
        if (j + 1 < L) w |= (uint64_t)(uint32_t)d[j + 1] << (30 - s);
        if (j + 2 < L && 60 - s < 32) w |= (uint64_t)(uint32_t)d[j + 2] << (60 - s);
        r.l[k] = (uint32_t)w;
    }
    // r = (a R)^-1 = a^-1 R^-1; one Montgomery product with R^3 puts it back in Montgomery form
    const Fp<C> r2 = Fp<C>::r2();
    return fp_mul(r, fp_mul(r2, r2));
}

using Fq = Fp<FqCfg>;
using Fr = Fp<FrCfg>;

// 16-byte vector loads/stores of whole elements (global or shared); p must be 16-byte aligned
template <class C>
VSP_DEV Fp<C> fp_load(const void* p) {
    Fp<C> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
    for (int i = 0; i < C::N / 4; ++i) {
        uint4 v = q[i];
        r.l[4 * i] = v.x;
        r.l[4 * i + 1] = v.y;
        r.l[4 * i + 2] = v.z;
        r.l[4 * i + 3] = v.w;
    }
    return r;
}
template <class C>
VSP_DEV void fp_store(void* p, const Fp<C>& a) {
    uint4* q = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < C::N / 4; ++i) q[i] = make_uint4(a.l[4 * i], a.l[4 * i + 1], a.l[4 * i + 2], a.l[4 * i + 3]);
}

}  // namespace vsp
