// ec.cuh — BLS12-381 G1 / G2 group arithmetic for sm_100a: Fq2, affine + XYZZ points (a = 0 short Weierstrass).
//
// Replaces crypto3-algebra's `curves::bls12<381>::g1_type<> / g2_type<>` value types (Jacobian a4 = 0 in the
// reference: /root/reference/bin/cli/include/nil/vote_saver/common.hpp:119; source in the un-vendored
// submodule crypto3-algebra @ e4fa2c5e).  The device path accumulates in XYZZ (x = X/ZZ, y = Y/ZZZ), whose mixed
// addition costs 8M + 2S instead of Jacobian's 7M + 4S; every result leaves the device as a canonical affine
// point, so the coordinate system is invisible to parity.
#pragma once
#include "fp.cuh"

namespace vsp {

// ---------------------------------------------------------------------------------------------
// field "ops" façades so curve code is written once for Fq and Fq2
// ---------------------------------------------------------------------------------------------
struct FqOps {
    using T = Fq;
    VSP_DEV static T zero() { return Fq::zero(); }
    VSP_DEV static T one() { return Fq::one(); }
    VSP_DEV static T add(const T& a, const T& b) { return fp_add(a, b); }
    VSP_DEV static T sub(const T& a, const T& b) { return fp_sub(a, b); }
#ifndef VSP_FQ_MUL_INLINE
    // ONE shared copy of each product per kernel (non-inlined; operands by value so that they travel in registers)
    // instead of ten inlined 6 KB copies per mixed addition.  With three resident blocks per SM the accumulation kernel
    // spent 1.2 warp-cycles per issue waiting for instructions (ncu: stalled_no_instruction; 0.34 at two blocks): twelve
    // warps at different places of a 60 KB loop body do not fit the instruction cache.  Measured on the ballot step:
    // G1 accumulation 109.5 -> 105.2 ms, bucket reduction 24.4 -> 19.9 ms, 1,300 -> 1,368 proofs/s.
    static __device__ __noinline__ T mul(T a, T b) { return fp_mul(a, b); }
    static __device__ __noinline__ T sqr(T a) { return fp_sqr(a); }
    static __device__ __noinline__ T mul_sub_mul(T a, T b, T c, T d) { return fp_mul_sub_mul(a, b, c, d); }
#else
    VSP_DEV static T mul(const T& a, const T& b) { return fp_mul(a, b); }
    VSP_DEV static T sqr(const T& a) { return fp_sqr(a); }
    // a b - c d on one reduction
    VSP_DEV static T mul_sub_mul(const T& a, const T& b, const T& c, const T& d) { return fp_mul_sub_mul(a, b, c, d); }
#endif
    VSP_DEV static T neg(const T& a) { return fp_neg(a); }
    VSP_DEV static T dbl(const T& a) { return fp_dbl(a); }
    VSP_DEV static bool is_zero(const T& a) { return a.is_zero(); }
    VSP_DEV static bool eq(const T& a, const T& b) { return a == b; }
    // the constant-time division-step inverse (fp.cuh): ~30x fewer multiplies than the power a^(p-2), same result
    static __device__ __noinline__ T inv(const T& a) { return fp_inv_safegcd(a); }
    static __device__ __noinline__ T inv_lowlat(const T& a) { return fp_inv_gcd(a); }
    // constant-time division-step inverse: ~30x fewer multiplies than the power, no divergence between lanes
    VSP_DEV static T inv_fast(const T& a) { return fp_inv_safegcd(a); }
    VSP_DEV static T to_mont(const T& a) { return fp_to_mont(a); }
    VSP_DEV static T from_mont(const T& a) { return fp_from_mont(a); }
    VSP_DEV static T load(const void* p) { return fp_load<FqCfg>(p); }
    VSP_DEV static void store(void* p, const T& a) { fp_store(p, a); }
    static constexpr int BYTES = 48;
};

struct alignas(16) Fq2 {
    Fq c0, c1;
};

struct Fq2Ops {
    using T = Fq2;
    VSP_DEV static T zero() { return T{Fq::zero(), Fq::zero()}; }
    VSP_DEV static T one() { return T{Fq::one(), Fq::zero()}; }
    VSP_DEV static T add(const T& a, const T& b) { return T{fp_add(a.c0, b.c0), fp_add(a.c1, b.c1)}; }
    VSP_DEV static T sub(const T& a, const T& b) { return T{fp_sub(a.c0, b.c0), fp_sub(a.c1, b.c1)}; }
    // (a0 + a1 u)(b0 + b1 u), u^2 = -1, Karatsuba: 3 Fq multiplications.  One shared copy (28 of them inlined per mixed
    // addition would not fit the instruction cache); operands BY VALUE so that they travel in registers — a reference
    // parameter of a non-inlined function makes the caller spill both operands to its stack frame
    // Lazy reduction: the three products stay double-width, the two results are formed there and reduced once each — two
    // Montgomery reductions instead of three (3 x 144 + 2 x 156 = 744 multiply-adds against 900).
    //   c1 = (a0 + a1)(b0 + b1) - a0 b0 - a1 b1 = a0 b1 + a1 b0 >= 0, < 2 p^2;   c0 = a0 b0 - a1 b1 + p^2 in (0, 2 p^2)
    static __device__ __noinline__ T mul(T a, T b) {
        uint32_t t0[24], t1[24], t2[24];
        fp_mul_wide<FqCfg>(a.c0, b.c0, t0);
        fp_mul_wide<FqCfg>(a.c1, b.c1, t1);
        fp_mul_wide<FqCfg>(fp_add_raw(a.c0, a.c1), fp_add_raw(b.c0, b.c1), t2);
        wide_sub(t2, t0);
        wide_sub(t2, t1);
        wide_add_p2<FqCfg>(t0);
        wide_sub(t0, t1);
        return T{fp_redc<FqCfg>(t0), fp_redc<FqCfg>(t2)};
    }
    // a b - c d on shared reductions (the Y3 = R (Q - X3) - Y1 PPP of the addition formulas): 6 wide products, 2 reductions
    static __device__ __noinline__ T mul_sub_mul(T a, T b, T c, T d) {
        uint32_t t0[24], t1[24], t2[24], u[24];
        fp_mul_wide<FqCfg>(a.c0, b.c0, t0);
        fp_mul_wide<FqCfg>(a.c1, b.c1, t1);
        fp_mul_wide<FqCfg>(fp_add_raw(a.c0, a.c1), fp_add_raw(b.c0, b.c1), t2);
        wide_sub(t2, t0);
        wide_sub(t2, t1);          // (a b).c1 = a0 b1 + a1 b0
        wide_add_p2<FqCfg>(t0);
        wide_sub(t0, t1);          // (a b).c0 + p^2
        // minus c d: c0 -= c0 d0 - c1 d1;  c1 -= c0 d1 + c1 d0
        fp_mul_wide<FqCfg>(c.c1, d.c1, u);
        wide_add(t0, u);           // + c1 d1
        fp_mul_wide<FqCfg>(c.c0, d.c0, t1);
        wide_add_p2<FqCfg>(t0);
        wide_sub(t0, t1);          // - c0 d0 (+ p^2): t0 in (0, 4 p^2)
        wide_add(u, t1);           // u = c0 d0 + c1 d1
        fp_mul_wide<FqCfg>(fp_add_raw(c.c0, c.c1), fp_add_raw(d.c0, d.c1), t1);
        wide_sub(t1, u);           // c0 d1 + c1 d0 < 2 p^2
        wide_add_p2<FqCfg>(t2);
        wide_add_p2<FqCfg>(t2);
        wide_sub(t2, t1);          // t2 in (0, 4 p^2)
        return T{fp_redc<FqCfg>(t0), fp_redc<FqCfg>(t2)};
    }
    static __device__ __noinline__ T sqr(T a) {
        Fq m = fp_mul(a.c0, a.c1);
        Fq r0 = fp_mul(fp_add(a.c0, a.c1), fp_sub(a.c0, a.c1));
        return T{r0, fp_dbl(m)};
    }
    VSP_DEV static T neg(const T& a) { return T{fp_neg(a.c0), fp_neg(a.c1)}; }
    VSP_DEV static T dbl(const T& a) { return T{fp_dbl(a.c0), fp_dbl(a.c1)}; }
    VSP_DEV static bool is_zero(const T& a) { return a.c0.is_zero() && a.c1.is_zero(); }
    VSP_DEV static bool eq(const T& a, const T& b) { return a.c0 == b.c0 && a.c1 == b.c1; }
    static __device__ __noinline__ T inv(const T& a) {
        Fq n = fp_inv_safegcd(fp_add(fp_sqr(a.c0), fp_sqr(a.c1)));
        return T{fp_mul(a.c0, n), fp_neg(fp_mul(a.c1, n))};
    }
    static __device__ __noinline__ T inv_lowlat(const T& a) {
        Fq n = fp_inv_gcd(fp_add(fp_sqr(a.c0), fp_sqr(a.c1)));
        return T{fp_mul(a.c0, n), fp_neg(fp_mul(a.c1, n))};
    }
    static __device__ __noinline__ T inv_fast(const T& a) {
        Fq n = fp_inv_safegcd(fp_add(fp_sqr(a.c0), fp_sqr(a.c1)));
        return T{fp_mul(a.c0, n), fp_neg(fp_mul(a.c1, n))};
    }
    VSP_DEV static T to_mont(const T& a) { return T{fp_to_mont(a.c0), fp_to_mont(a.c1)}; }
    VSP_DEV static T from_mont(const T& a) { return T{fp_from_mont(a.c0), fp_from_mont(a.c1)}; }
    VSP_DEV static T load(const void* p) {
        const char* q = reinterpret_cast<const char*>(p);
        return T{fp_load<FqCfg>(q), fp_load<FqCfg>(q + 48)};
    }
    VSP_DEV static void store(void* p, const T& a) {
        char* q = reinterpret_cast<char*>(p);
        fp_store(q, a.c0);
        fp_store(q + 48, a.c1);
    }
    static constexpr int BYTES = 96;
};

// ---------------------------------------------------------------------------------------------
// points
// ---------------------------------------------------------------------------------------------
template <class F>
struct Affine {
    typename F::T x, y;  // infinity = (0, 0): not on y^2 = x^3 + b, b != 0
    VSP_DEV bool is_inf() const { return F::is_zero(x) && F::is_zero(y); }
    VSP_DEV static Affine inf() { return Affine{F::zero(), F::zero()}; }
    VSP_DEV static Affine load(const void* p) {
        const char* q = reinterpret_cast<const char*>(p);
        return Affine{F::load(q), F::load(q + F::BYTES)};
    }
    VSP_DEV void store(void* p) const {
        char* q = reinterpret_cast<char*>(p);
        F::store(q, x);
        F::store(q + F::BYTES, y);
    }
    static constexpr int BYTES = 2 * F::BYTES;
};

template <class F>
struct XYZZ {
    using E = typename F::T;
    E X, Y, ZZ, ZZZ;  // infinity <=> ZZ == 0
    static constexpr int BYTES = 4 * F::BYTES;

    VSP_DEV bool is_inf() const { return F::is_zero(ZZ); }
    VSP_DEV static XYZZ inf() { return XYZZ{F::zero(), F::zero(), F::zero(), F::zero()}; }
    VSP_DEV static XYZZ from_affine(const Affine<F>& a) {
        if (a.is_inf()) return inf();
        return XYZZ{a.x, a.y, F::one(), F::one()};
    }
    VSP_DEV static XYZZ load(const void* p) {
        const char* q = reinterpret_cast<const char*>(p);
        return XYZZ{F::load(q), F::load(q + F::BYTES), F::load(q + 2 * F::BYTES), F::load(q + 3 * F::BYTES)};
    }
    VSP_DEV void store(void* p) const {
        char* q = reinterpret_cast<char*>(p);
        F::store(q, X);
        F::store(q + F::BYTES, Y);
        F::store(q + 2 * F::BYTES, ZZ);
        F::store(q + 3 * F::BYTES, ZZZ);
    }
    VSP_DEV XYZZ negated() const { return XYZZ{X, F::neg(Y), ZZ, ZZZ}; }
};

// 2 * (affine point)   (mdbl-2008-s-1, a = 0)
template <class F>
__device__ __noinline__ XYZZ<F> ec_dbl_affine(const Affine<F>& p) {
    using E = typename F::T;
    E U = F::dbl(p.y);
    E V = F::sqr(U);
    E W = F::mul(U, V);
    E S = F::mul(p.x, V);
    E xx = F::sqr(p.x);
    E M = F::add(F::dbl(xx), xx);
    E X3 = F::sub(F::sqr(M), F::dbl(S));
    E Y3 = F::mul_sub_mul(M, F::sub(S, X3), W, p.y);
    return XYZZ<F>{X3, Y3, V, W};
}

// acc += q  (q affine, optionally negated by the caller)   madd-2008-s: 8M + 2S
template <class F>
VSP_DEV void ec_madd(XYZZ<F>& acc, const Affine<F>& q) {
    using E = typename F::T;
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = XYZZ<F>{q.x, q.y, F::one(), F::one()};
        return;
    }
    E U2 = F::mul(q.x, acc.ZZ);
    E S2 = F::mul(q.y, acc.ZZZ);
    E P = F::sub(U2, acc.X);
    E R = F::sub(S2, acc.Y);
    if (F::is_zero(P)) {
        if (F::is_zero(R)) {
            acc = ec_dbl_affine<F>(q);
        } else {
            acc = XYZZ<F>::inf();
        }
        return;
    }
    // every value is consumed as early as the formulas allow (the products are calls, which the compiler does not
    // reorder): PP and P die before the second half, so the G2 kernel carries one Fq2 less across it
    E PP = F::sqr(P);
    E Q = F::mul(acc.X, PP);
    E PPP = F::mul(P, PP);
    acc.ZZ = F::mul(acc.ZZ, PP);
    acc.ZZZ = F::mul(acc.ZZZ, PPP);
    acc.X = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    acc.Y = F::mul_sub_mul(R, F::sub(Q, acc.X), acc.Y, PPP);
}

// ---------------------------------------------------------------------------------------------
// affine additions on a shared inversion (the bucket accumulation of msm_impl.cuh)
//   P + Q, P != +-Q:  lambda = (y2 - y1) / (x2 - x1);  x3 = lambda^2 - x1 - x2;  y3 = lambda (x1 - x3) - y1
// With 1 / (x2 - x1) supplied, that is 2 products + 1 square; Montgomery's trick supplies it for 3 more products per
// addition and one inversion per batch: 5M + 1S against the 8M + 2S of the XYZZ mixed addition.
// ---------------------------------------------------------------------------------------------
template <class F>
VSP_DEV void ec_affine_add_with_inverse(typename F::T& x1, typename F::T& y1, const typename F::T& x2, const typename F::T& y2,
                                        const typename F::T& dinv) {
    using E = typename F::T;
    const E lam = F::mul(F::sub(y2, y1), dinv);
    const E x3 = F::sub(F::sub(F::sqr(lam), x1), x2);
    y1 = F::sub(F::mul(lam, F::sub(x1, x3)), y1);
    x1 = x3;
}
// 2 P in affine coordinates with its own inversion (the rare "same point twice in one bucket" case); false: 2 P = O
template <class F>
__device__ __noinline__ bool ec_affine_dbl(typename F::T& x, typename F::T& y) {
    using E = typename F::T;
    if (F::is_zero(y)) return false;
    const E xx = F::sqr(x);
    const E lam = F::mul(F::add(F::dbl(xx), xx), F::inv_fast(F::dbl(y)));
    const E x3 = F::sub(F::sqr(lam), F::dbl(x));
    y = F::sub(F::mul(lam, F::sub(x, x3)), y);
    x = x3;
    return true;
}

// dbl-2008-s-1, a = 0   (not inlined: only the cold reduction kernels use it, and sharing one copy keeps ptxas fast)
template <class F>
__device__ __noinline__ XYZZ<F> ec_dbl(const XYZZ<F>& p) {
    using E = typename F::T;
    if (p.is_inf()) return p;
    E U = F::dbl(p.Y);
    E V = F::sqr(U);
    E W = F::mul(U, V);
    E S = F::mul(p.X, V);
    E xx = F::sqr(p.X);
    E M = F::add(F::dbl(xx), xx);
    E X3 = F::sub(F::sqr(M), F::dbl(S));
    E Y3 = F::mul_sub_mul(M, F::sub(S, X3), W, p.Y);
    return XYZZ<F>{X3, Y3, F::mul(V, p.ZZ), F::mul(W, p.ZZZ)};
}

// acc += q   add-2008-s: 12M + 2S   (not inlined, see ec_dbl)
template <class F>
__device__ __noinline__ void ec_add(XYZZ<F>& acc, const XYZZ<F>& q) {
    using E = typename F::T;
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = q;
        return;
    }
    E U1 = F::mul(acc.X, q.ZZ);
    E P = F::sub(F::mul(q.X, acc.ZZ), U1);
    E S1 = F::mul(acc.Y, q.ZZZ);
    E R = F::sub(F::mul(q.Y, acc.ZZZ), S1);
    if (F::is_zero(P)) {
        if (F::is_zero(R)) {
            acc = ec_dbl<F>(acc);
        } else {
            acc = XYZZ<F>::inf();
        }
        return;
    }
    // as in ec_madd: every value consumed as early as the formulas allow, results written back to `acc` (memory) at once
    E PP = F::sqr(P);
    E Q = F::mul(U1, PP);
    E PPP = F::mul(P, PP);
    acc.ZZ = F::mul(F::mul(acc.ZZ, q.ZZ), PP);
    acc.ZZZ = F::mul(F::mul(acc.ZZZ, q.ZZZ), PPP);
    acc.X = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    acc.Y = F::mul_sub_mul(R, F::sub(Q, acc.X), S1, PPP);
}

// canonical affine (Montgomery form) of an XYZZ point: one field inversion
template <class F>
__device__ __noinline__ Affine<F> ec_to_affine(const XYZZ<F>& p, bool lowlat = false) {
    using E = typename F::T;
    if (p.is_inf()) return Affine<F>::inf();
    // lowlat: the caller is the only working lane of its warp (fp_inv_gcd's branches cost nothing then)
    E i = lowlat ? F::inv_lowlat(F::mul(p.ZZ, p.ZZZ)) : F::inv(F::mul(p.ZZ, p.ZZZ));
    E zz_inv = F::mul(i, p.ZZZ);
    E zzz_inv = F::mul(i, p.ZZ);
    return Affine<F>{F::mul(p.X, zz_inv), F::mul(p.Y, zzz_inv)};
}

// k * p for a small non-negative k (bucket-segment weights), MSB-first double-and-add
template <class F>
VSP_DEV XYZZ<F> ec_mul_small(const XYZZ<F>& p, uint32_t k) {
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int bit = 31 - __clz(k | 1u); bit >= 0; --bit) {
        acc = ec_dbl<F>(acc);
        if ((k >> bit) & 1u) ec_add<F>(acc, p);
    }
    if (k == 0) return XYZZ<F>::inf();
    return acc;
}

using G1Affine = Affine<FqOps>;
using G2Affine = Affine<Fq2Ops>;
using G1XYZZ = XYZZ<FqOps>;
using G2XYZZ = XYZZ<Fq2Ops>;

}  // namespace vsp
