// emul.cpp — the device field / curve templates (csrc/fp.cuh, csrc/ec.cuh) compiled with g++ under VSP_HOST_EMUL,
// where the PTX carry-chain instructions are emulated bit for bit.  Lets the CPU test tier check the limb schedules
// (Montgomery multiplication on even/odd accumulators, XYZZ formulas, exceptional cases) against the oracle.
// Values cross this boundary as canonical little-endian u64 limbs, like the product ABI.
#define VSP_HOST_EMUL 1
#include <cstdint>
#include <cstring>

#include "../../vote-saver-protocol_b200/csrc/ec.cuh"
#include "../../vote-saver-protocol_b200/csrc/scratch_class.h"

using namespace vsp;

template <class C>
static Fp<C> ld(const uint64_t* p) {
    Fp<C> r;
    memcpy(r.l, p, 4 * C::N);
    return r;
}
template <class C>
static void st(uint64_t* p, const Fp<C>& a) {
    memcpy(p, a.l, 4 * C::N);
}

template <class C>
static void binop(int op, const uint64_t* a, const uint64_t* b, uint64_t* out) {
    Fp<C> x = fp_to_mont(ld<C>(a)), y = fp_to_mont(ld<C>(b)), r;
    switch (op) {
        case 0: r = fp_mul(x, y); break;
        case 1: r = fp_add(x, y); break;
        case 2: r = fp_sub(x, y); break;
        case 3: r = fp_neg(x); break;
        case 4: r = fp_inv(x); break;
        case 5: r = fp_sqr(x); break;
        case 6: r = fp_inv_gcd(x); break;
        case 7: r = fp_inv_safegcd(x); break;
        default: r = Fp<C>::zero();
    }
    st<C>(out, fp_from_mont(r));
}

template <class F>
static Affine<F> ld_aff(const uint64_t* p) {
    Affine<F> a = Affine<F>::load(p);
    if (!a.is_inf()) {
        a.x = F::to_mont(a.x);
        a.y = F::to_mont(a.y);
    }
    return a;
}
template <class F>
static void st_aff(uint64_t* out, const XYZZ<F>& p) {
    Affine<F> a = ec_to_affine<F>(p);
    a.x = F::from_mont(a.x);
    a.y = F::from_mont(a.y);
    a.store(out);
}

// op 0: madd(from_affine(a), b)   1: add(xyzz(a)*, xyzz(b)*) with non-trivial ZZ   2: dbl(xyzz(a)*)   3: k * a (small k)
template <class F>
static void ecop(int op, const uint64_t* a, const uint64_t* b, uint32_t k, uint64_t* out) {
    Affine<F> pa = ld_aff<F>(a), pb = ld_aff<F>(b);
    XYZZ<F> xa = XYZZ<F>::from_affine(pa), xb = XYZZ<F>::from_affine(pb);
    // give the XYZZ operands a non-trivial denominator: 2P - P computed through doubling and a negated mixed add
    auto denorm = [](const XYZZ<F>& p, const Affine<F>& aff) {
        if (p.is_inf()) return p;
        XYZZ<F> d = ec_dbl<F>(p);
        Affine<F> n = aff;
        n.y = F::neg(n.y);
        ec_madd<F>(d, n);
        return d;
    };
    XYZZ<F> r;
    switch (op) {
        case 0: r = xa; ec_madd<F>(r, pb); break;
        case 1: r = denorm(xa, pa); ec_add<F>(r, denorm(xb, pb)); break;
        case 2: r = ec_dbl<F>(denorm(xa, pa)); break;
        case 3: r = ec_mul_small<F>(denorm(xa, pa), k); break;
        case 4: r = ec_dbl_affine<F>(pa); break;
        default: r = XYZZ<F>::inf();
    }
    st_aff<F>(out, r);
}

// op 0: redc(mul_wide(a, b)) on raw limbs (a, b any N-limb values with a b < p 2^(32N));  1: fp_mul_sub_mul(a, b, c, d)
template <class C>
static void wideop(int op, const uint64_t* a, const uint64_t* b, const uint64_t* c, const uint64_t* d, uint64_t* out) {
    if (op == 0) {
        uint32_t T[2 * C::N];
        fp_mul_wide<C>(ld<C>(a), ld<C>(b), T);
        st<C>(out, fp_redc<C>(T));
    } else {
        st<C>(out, fp_mul_sub_mul<C>(ld<C>(a), ld<C>(b), ld<C>(c), ld<C>(d)));
    }
}

extern "C" {
void emul_fq_wide(int op, const uint64_t* a, const uint64_t* b, const uint64_t* c, const uint64_t* d, uint64_t* out) { wideop<FqCfg>(op, a, b, c, d, out); }
void emul_fr_wide(int op, const uint64_t* a, const uint64_t* b, const uint64_t* c, const uint64_t* d, uint64_t* out) { wideop<FrCfg>(op, a, b, c, d, out); }
// Fq2 on raw Montgomery-domain limbs (c0 | c1): op 0 mul, 1 sqr, 2 mul_sub_mul(a, b, c, d)
void emul_fq2_raw(int op, const uint64_t* a, const uint64_t* b, const uint64_t* c, const uint64_t* d, uint64_t* out) {
    auto L = [](const uint64_t* p) { return Fq2{ld<FqCfg>(p), ld<FqCfg>(p + 6)}; };
    Fq2 r = op == 0 ? Fq2Ops::mul(L(a), L(b)) : op == 1 ? Fq2Ops::sqr(L(a)) : Fq2Ops::mul_sub_mul(L(a), L(b), L(c), L(d));
    st<FqCfg>(out, r.c0);
    st<FqCfg>(out + 6, r.c1);
}
void emul_fr_op(int op, const uint64_t* a, const uint64_t* b, uint64_t* out) { binop<FrCfg>(op, a, b, out); }
void emul_fq_op(int op, const uint64_t* a, const uint64_t* b, uint64_t* out) { binop<FqCfg>(op, a, b, out); }
void emul_g1_op(int op, const uint64_t* a, const uint64_t* b, uint32_t k, uint64_t* out) { ecop<FqOps>(op, a, b, k, out); }
void emul_g2_op(int op, const uint64_t* a, const uint64_t* b, uint32_t k, uint64_t* out) { ecop<Fq2Ops>(op, a, b, k, out); }
// raw Montgomery product of unreduced-domain limbs (checks fp_mul on arbitrary < p inputs without conversions)
void emul_fq_montmul_raw(const uint64_t* a, const uint64_t* b, uint64_t* out) { st<FqCfg>(out, fp_mul(ld<FqCfg>(a), ld<FqCfg>(b))); }
void emul_fr_montmul_raw(const uint64_t* a, const uint64_t* b, uint64_t* out) { st<FrCfg>(out, fp_mul(ld<FrCfg>(a), ld<FrCfg>(b))); }
// raw Montgomery square (fp_sqr's own schedule: doubled cross products + diagonal, then the sliding-frame reduction)
void emul_fq_montsqr_raw(const uint64_t* a, uint64_t* out) { st<FqCfg>(out, fp_sqr(ld<FqCfg>(a))); }
void emul_fr_montsqr_raw(const uint64_t* a, uint64_t* out) { st<FrCfg>(out, fp_sqr(ld<FrCfg>(a))); }
}

// the scratch cache's size class of a request of n bytes (csrc/scratch_class.h)
extern "C" uint64_t emul_scratch_size_class(uint64_t n) { return (uint64_t)vsp::scratch_size_class((size_t)n); }
