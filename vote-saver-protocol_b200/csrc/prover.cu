// prover.cu — Groth16 (r1cs_gg_ppzksnark) prover for VoteSAVER's encrypted-input ballots and the SAVER
// ciphertext / rerandomisation, orchestrated on CUDA streams.
//
// Replaces (paths relative to /root/reference):
//   zk::prove<r1cs_gg_ppzksnark<bls12<381>, ..., proving_mode::encrypted_input>>   call site
//       libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:353-354; prover and r1cs_to_qap::witness_map live in
//       the un-vendored submodule crypto3-zk @ d335bb8b (.SUBMODULES.json:199) — restated per SURVEY.md 8 a-4 / a-7:
//         witness_map: evaluate A w, B w, C w; rows n_c .. n_c + n_in of A carry (1, x_1 .. x_nin);
//                      iFFT, coset FFT (x3), H = (A o B - C) / Z on the coset, coset iFFT
//         g_A = alpha + A(w) + r delta;  g_B = beta2 + B2(w) + s delta2;  g1_B = beta + B1(w) + s delta
//         g_C = H(h) + L(w_aux) + s g_A + r g1_B - (r s) delta   [+ r_saver * gamma_inverse_sum_s_g1]
//   encrypt_op<elgamal_verifiable>::process      elgamal_verifiable.hpp:322-357
//   rerandomize_op<elgamal_verifiable>::process  elgamal_verifiable.hpp:632-667
//
// Ballots are proved in chunks of up to `chunk` (VSP_BATCH_CHUNK, default 64) that share every launch: each kernel
// carries the ballot index in a grid dimension, so one chunk costs the launches of one proof and the serial tails
// (field inversions, the two variable-base multiples) run for all its ballots at once.
// Stream plan for one chunk (fork after the witness upload, join before assembly):
//   main   : A w / B w / C w  ->  3B batched iNTT+coset, NTT  ->  pointwise H  ->  coset iNTT  ->  multiexp H  ->  assembly
//   side 2 : multiexp A, multiexp L       side 3 : fixed-base r/s multiples, multiexp B (G2), multiexp B (G1)
#include "common.cuh"
#include <type_traits>
#include "ec.cuh"
#include "host_fr.h"
#include "msm.cuh"

#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

namespace vsp {
int domain_plan(uint64_t min_size, int* kind, uint64_t* m);
int fill_power_table(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_out, uint64_t n, const hostfr::Fr4& g,
                     const hostfr::Fr4& k, bool canonical_out = false);
}  // namespace vsp

using namespace vsp;

struct vsp_pk {
    vsp_ctx* ctx = nullptr;
    uint64_t ni = 0, na = 0, nc = 0, nv = 0, m = 0;
    int kind = 0;
    // point-range slice of the five queries held by this key (full key: lo = 0, n = whole query)
    bool sliced = false;
    uint64_t A_lo = 0, A_n = 0, B_lo = 0, H_lo = 0, H_n = 0, L_lo = 0, L_n = 0;
    bool short_rows = false;  // average row has <= 8 terms per matrix: thread-per-row R1CS evaluation
    // constraints with more than LONG_ROW_TERMS terms in one matrix (e.g. the closing "sum of everything, squared"
    // constraint of the reference's synthetic examples, r1cs_examples.hpp:116-126) are cut over many blocks
    uint32_t* d_long_rows = nullptr;
    unsigned n_long = 0;
    // CSR (device): rows u64, cols u32, values Montgomery (A, C: v R;  B: v R^2 so that B w comes out in Montgomery form)
    uint64_t* d_row[3] = {nullptr, nullptr, nullptr};
    uint32_t* d_col[3] = {nullptr, nullptr, nullptr};
    uint32_t* d_val[3] = {nullptr, nullptr, nullptr};
    BasesDev A, Bg1, Bg2, H, L;
    BasesDev fixed1;  // alpha_g1, beta_g1, delta_g1 (Montgomery affine)
    BasesDev fixed2;  // beta_g2, delta_g2
    // 8-bit fixed-base tables T[w][i] = 2^(8 w) P_i:  fx1tab = {delta_g1, gamma_inverse_sum_s_g1}, fx2tab = {delta_g2}
    BasesDev fx1tab, fx2tab;
    uint64_t gamma_key[12] = {0};  // the gamma_inverse_sum_s_g1 currently in fx1tab (all-zero: none)
    uint32_t* d_B_idx = nullptr;
    uint64_t B_count = 0;
    uint32_t* d_gpow = nullptr;     // g^i, i < m (Montgomery)
    uint32_t* d_ginvpow = nullptr;  // basic domain: Z(g)^-1 g^-i ; step domain: g^-i
    uint32_t* d_zinv = nullptr;     // step domain only: 1 / Z on the coset, per element (Montgomery)
    uint64_t h_delta_g1[12];        // canonical copies
    uint64_t h_delta_g2[24];
};

// SAVER public key resident on the GPU: 8-bit fixed-base tables of its 3 msg_size + 2 G1 bases
//   order: delta | delta_s[0..n) | gamma_abc_rest[0..n) | t[0..n) | delta_sum_s
struct vsp_saver_pub {
    vsp_ctx* ctx = nullptr;
    unsigned n = 0;  // msg_size
    BasesDev tab;
    uint64_t gamma_inv[12];  // canonical gamma_inverse_sum_s_g1
};

namespace vsp {

// ---------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------
__global__ void k_vals_to_mont(uint32_t* v, uint64_t n, int squared) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr x = fp_load<FrCfg>(v + 8 * i);
    x = fp_mul(x, Fr::r2());                 // v R
    if (squared) x = fp_mul(x, Fr::r2());    // v R^2
    fp_store(v + 8 * i, x);
}

VSP_DEV Fr warp_sum(Fr a) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        Fr b;
#pragma unroll
        for (int k = 0; k < 8; ++k) b.l[k] = __shfl_down_sync(0xffffffffu, a.l[k], off);
        a = fp_add(a, b);
    }
    return a;
}

struct CsrDev {
    const uint64_t* row[3];
    const uint32_t* col[3];
    const uint32_t* val[3];
};

constexpr uint64_t LONG_ROW_TERMS = 4096;
constexpr unsigned LONG_SLICES = 512;

// one warp per (domain row, ballot = blockIdx.y): aA / aB / aC[row] = <matrix row, (1 || primary || aux)>; rows
// n_c .. n_c + n_in of aA hold the input-consistency values, everything above is zero.  aB comes out in Montgomery form
// (its values carry R^2).  abc = [aA: B x m | aB: B x m | aC: B x m]
__global__ void __launch_bounds__(256) k_r1cs_eval(CsrDev csr, const uint32_t* __restrict__ w_all, uint64_t w_stride,
                                                   uint64_t nc, uint64_t ni, uint64_t m, uint32_t* abc) {
    const uint64_t row = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (row >= m) return;
    const uint64_t B = gridDim.y, b = blockIdx.y;
    const uint32_t* w = w_all + 8 * b * w_stride;
    uint32_t* outs[3] = {abc + 8 * (b * m), abc + 8 * ((B + b) * m), abc + 8 * ((2 * B + b) * m)};
    if (row >= nc) {
        if (lane == 0) {
            Fr z = Fr::zero();
            Fr a = (row - nc <= ni) ? fp_load<FrCfg>(w + 8 * (row - nc)) : z;
            fp_store(outs[0] + 8 * row, a);
            fp_store(outs[1] + 8 * row, z);
            fp_store(outs[2] + 8 * row, z);
        }
        return;
    }
#pragma unroll
    for (int mtx = 0; mtx < 3; ++mtx) {
        const uint64_t lo = csr.row[mtx][row], hi = csr.row[mtx][row + 1];
        if (hi - lo > LONG_ROW_TERMS) continue;  // k_r1cs_long_* writes it (warp-uniform)
        Fr acc = Fr::zero();
        for (uint64_t k = lo + lane; k < hi; k += 32) {
            Fr v = fp_load<FrCfg>(csr.val[mtx] + 8 * k);
            Fr x = fp_load<FrCfg>(w + 8 * uint64_t(csr.col[mtx][k]));
            acc = fp_add(acc, fp_mul(v, x));
        }
        acc = warp_sum(acc);
        if (lane == 0) fp_store(outs[mtx] + 8 * row, acc);
    }
}

// the same with one thread per (row, ballot): the ballot circuits' rows hold a handful of terms (booleanity, XOR and
// packing constraints), where a warp per row leaves 29+ lanes idle
__global__ void __launch_bounds__(128) k_r1cs_eval_rows(CsrDev csr, const uint32_t* __restrict__ w_all, uint64_t w_stride,
                                                        uint64_t nc, uint64_t ni, uint64_t m, uint32_t* abc) {
    const uint64_t row = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= m) return;
    const uint64_t B = gridDim.y, b = blockIdx.y;
    const uint32_t* w = w_all + 8 * b * w_stride;
    uint32_t* outs[3] = {abc + 8 * (b * m), abc + 8 * ((B + b) * m), abc + 8 * ((2 * B + b) * m)};
    if (row >= nc) {
        Fr z = Fr::zero();
        Fr a = (row - nc <= ni) ? fp_load<FrCfg>(w + 8 * (row - nc)) : z;
        fp_store(outs[0] + 8 * row, a);
        fp_store(outs[1] + 8 * row, z);
        fp_store(outs[2] + 8 * row, z);
        return;
    }
#pragma unroll 1
    for (int mtx = 0; mtx < 3; ++mtx) {
        const uint64_t lo = csr.row[mtx][row], hi = csr.row[mtx][row + 1];
        if (hi - lo > LONG_ROW_TERMS) continue;  // k_r1cs_long_* writes it
        Fr acc = Fr::zero();
        for (uint64_t k = lo; k < hi; ++k) {
            Fr v = fp_load<FrCfg>(csr.val[mtx] + 8 * k);
            Fr x = fp_load<FrCfg>(w + 8 * uint64_t(csr.col[mtx][k]));
            acc = fp_add(acc, fp_mul(v, x));
        }
        fp_store(outs[mtx] + 8 * row, acc);
    }
}

// long rows: block (slice, long row * 3 + matrix, ballot) sums its slice of the row's terms into part[...][slice];
// k_r1cs_long_final adds the slices up and writes the value (rows short in this matrix are left to the row kernels)
__global__ void __launch_bounds__(256) k_r1cs_long_partial(CsrDev csr, const uint32_t* __restrict__ long_rows,
                                                           const uint32_t* __restrict__ w_all, uint64_t w_stride,
                                                           uint32_t* __restrict__ part) {
    __shared__ uint32_t sm[8 * 8];
    const unsigned lr = blockIdx.y / 3, mtx = blockIdx.y % 3, b = blockIdx.z;
    const uint64_t row = long_rows[lr];
    const uint64_t lo = csr.row[mtx][row], hi = csr.row[mtx][row + 1];
    Fr acc = Fr::zero();
    if (hi - lo > LONG_ROW_TERMS) {
        const uint64_t per = (hi - lo + LONG_SLICES - 1) / LONG_SLICES;
        const uint64_t s0 = lo + per * blockIdx.x, s1 = s0 + per < hi ? s0 + per : hi;
        const uint32_t* w = w_all + 8 * uint64_t(b) * w_stride;
        for (uint64_t k = s0 + threadIdx.x; k < s1; k += blockDim.x) {
            Fr v = fp_load<FrCfg>(csr.val[mtx] + 8 * k);
            Fr x = fp_load<FrCfg>(w + 8 * uint64_t(csr.col[mtx][k]));
            acc = fp_add(acc, fp_mul(v, x));
        }
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) fp_store(sm + 8 * (threadIdx.x >> 5), acc);
    __syncthreads();
    if (threadIdx.x == 0) {
        for (unsigned q = 1; q < blockDim.x / 32; ++q) acc = fp_add(acc, fp_load<FrCfg>(sm + 8 * q));
        fp_store(part + 8 * ((uint64_t(b) * gridDim.y + blockIdx.y) * LONG_SLICES + blockIdx.x), acc);
    }
}
__global__ void k_r1cs_long_final(CsrDev csr, const uint32_t* __restrict__ long_rows, unsigned n_long, unsigned B, uint64_t m,
                                  const uint32_t* __restrict__ part, uint32_t* abc) {
    const unsigned q = blockIdx.x * blockDim.x + threadIdx.x;  // (ballot, long row, matrix)
    if (q >= B * n_long * 3) return;
    const unsigned b = q / (n_long * 3), rm = q % (n_long * 3), lr = rm / 3, mtx = rm % 3;
    const uint64_t row = long_rows[lr];
    if (csr.row[mtx][row + 1] - csr.row[mtx][row] <= LONG_ROW_TERMS) return;
    Fr acc = Fr::zero();
    for (unsigned sl = 0; sl < LONG_SLICES; ++sl) acc = fp_add(acc, fp_load<FrCfg>(part + 8 * (uint64_t(q) * LONG_SLICES + sl)));
    fp_store(abc + 8 * ((uint64_t(mtx) * B + b) * m + row), acc);
}

// h[b][i] = (a[b][i] * b_mont[b][i] - c[b][i]) [* zinv[i]]      (ballot b = blockIdx.y, vectors packed B x m)
__global__ void k_h_pointwise(uint32_t* a, const uint32_t* b, const uint32_t* c, const uint32_t* zinv, uint64_t m) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint64_t o = 8 * (uint64_t(blockIdx.y) * m + i);
    Fr x = fp_sub(fp_mul(fp_load<FrCfg>(a + o), fp_load<FrCfg>(b + o)), fp_load<FrCfg>(c + o));
    if (zinv) x = fp_mul(x, fp_load<FrCfg>(zinv + 8 * i));
    fp_store(a + o, x);
}
// a[v][i] *= tbl[i]   (v = blockIdx.y, vectors packed)
__global__ void k_mul_table2(uint32_t* a, uint64_t n, const uint32_t* tbl) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t o = 8 * (uint64_t(blockIdx.y) * n + i);
    fp_store(a + o, fp_mul(fp_load<FrCfg>(a + o), fp_load<FrCfg>(tbl + 8 * i)));
}
__global__ void k_fill_const(uint32_t* a, uint64_t n, Fr v) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, v);
}
// out[b][i] = w[b][idx[i]]
__global__ void k_gather_fr(const uint32_t* __restrict__ w, uint64_t w_stride, const uint32_t* __restrict__ idx, uint64_t n,
                            uint32_t* out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t b = blockIdx.y;
    fp_store(out + 8 * (b * n + i), fp_load<FrCfg>(w + 8 * (b * w_stride + idx[i])));
}
// w[b][0] = 1 (the constant-one variable of every assignment)
__global__ void k_set_one(uint32_t* w, uint64_t w_stride, unsigned B) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    Fr one = Fr::zero();
    one.l[0] = 1;
    fp_store(w + 8 * (uint64_t(b) * w_stride), one);
}

// k * P, P affine (Montgomery), k canonical 8 x 32-bit: MSB-first double-and-add
template <class F>
VSP_DEV XYZZ<F> ec_scalar_mul(const Affine<F>& p, const uint32_t* k) {
    XYZZ<F> acc = XYZZ<F>::inf();
    int top = 255;
    while (top >= 0 && !((k[top >> 5] >> (top & 31)) & 1u)) --top;
    for (int bit = top; bit >= 0; --bit) {
        acc = ec_dbl<F>(acc);
        if ((k[bit >> 5] >> (bit & 31)) & 1u) ec_madd<F>(acc, p);
    }
    return acc;
}

// G1 endomorphism (GLV): phi(x, y) = (beta x, y) = lambda (x, y) with lambda = z^2 - 1 (z the BLS parameter), since
// lambda^2 + lambda + 1 = z^4 - z^2 + 1 = r.  A scalar k < r splits as k = q (lambda + 1) + rem = (q + rem) + q lambda
// with q = floor(k / z^2) < 2^128, so k P = k1 P + k2 phi(P) with k1 = q + rem < 2^129, k2 = q: one joint
// double-and-add over 129 bits with the table {P, phi P, P + phi P} instead of 255 doublings.  beta is the cube root of
// unity matching this lambda (checked against the CPU restatement's lambda * G in tests/test_gpu_keygen.py through
// vsp_g1_mul_batch); canonical limbs here, converted once per thread.
__device__ const uint32_t GLV_BETA[12] = {0x0000aaacu, 0x8bfd0000u, 0x4f49fffdu, 0x409427ebu, 0x0fb85f9bu, 0x897d2965u, 0x89759ad4u, 0xaa0d857du, 0x63d4de85u, 0xec024086u, 0x397fe699u, 0x1a0111eau};
VSP_DEV void glv_split(const uint32_t* k, uint32_t (&k1)[5], uint32_t (&k2)[4]) {
    const uint64_t Lhi = 0xac45a4010001a402ULL, Llo = 0x0000000100000000ULL;  // z^2
    uint64_t rh = 0, rl = 0, qh = 0, ql = 0;
    for (int bit = 255; bit >= 0; --bit) {
        const uint64_t carry = rh >> 63;
        rh = (rh << 1) | (rl >> 63);
        rl = (rl << 1) | ((k[bit >> 5] >> (bit & 31)) & 1u);
        if (carry || rh > Lhi || (rh == Lhi && rl >= Llo)) {
            const uint64_t borrow = rl < Llo ? 1u : 0u;
            rl -= Llo;
            rh = rh - Lhi - borrow;
            if (bit < 64)
                ql |= uint64_t(1) << bit;
            else if (bit < 128)
                qh |= uint64_t(1) << (bit - 64);
        }
    }
    const uint64_t sl = ql + rl, c0 = sl < ql ? 1u : 0u;
    const uint64_t t = qh + rh, c1 = t < qh ? 1u : 0u;
    const uint64_t sh = t + c0, c2 = sh < t ? 1u : 0u;
    k1[0] = (uint32_t)sl; k1[1] = (uint32_t)(sl >> 32); k1[2] = (uint32_t)sh; k1[3] = (uint32_t)(sh >> 32);
    k1[4] = (uint32_t)(c1 + c2);
    k2[0] = (uint32_t)ql; k2[1] = (uint32_t)(ql >> 32); k2[2] = (uint32_t)qh; k2[3] = (uint32_t)(qh >> 32);
}
// k * P on G1 by the split above; P Montgomery affine, not infinity; k canonical < r
VSP_DEV G1XYZZ ec_scalar_mul_glv(const G1Affine& p, const uint32_t* k) {
    uint32_t k1[5], k2[4];
    glv_split(k, k1, k2);
    Fq beta;
#pragma unroll
    for (int i = 0; i < 12; ++i) beta.l[i] = GLV_BETA[i];
    G1Affine ph = p;
    ph.x = fp_mul(p.x, fp_to_mont(beta));
    G1XYZZ both = G1XYZZ::from_affine(p);
    ec_madd<FqOps>(both, ph);  // P + phi(P): distinct points, neither the other's negative (lambda != +-1)
    G1XYZZ acc = G1XYZZ::inf();
    int top = 128;
    while (top >= 0 && !(((k1[top >> 5] >> (top & 31)) | (top < 128 ? k2[top >> 5] >> (top & 31) : 0u)) & 1u)) --top;
    for (int bit = top; bit >= 0; --bit) {
        acc = ec_dbl<FqOps>(acc);
        const uint32_t b1 = (k1[bit >> 5] >> (bit & 31)) & 1u;
        const uint32_t b2 = bit < 128 ? (k2[bit >> 5] >> (bit & 31)) & 1u : 0u;
        if (b1 & b2)
            ec_add<FqOps>(acc, both);
        else if (b1)
            ec_madd<FqOps>(acc, p);
        else if (b2)
            ec_madd<FqOps>(acc, ph);
    }
    return acc;
}

// out[i] (XYZZ, Montgomery) = scalars[i] * points[i];  points canonical affine (mont == 0) or Montgomery affine.
// Variable-base multiples (s g_A, r g1_B; rerandomisation): one job per block, lane 0 of one warp — independent
// scalars diverge at every add step, so jobs sharing a warp would serialise; separate warps run concurrently.
// G1 jobs take the endomorphism split above; G2 jobs the plain double-and-add.
template <class F>
__global__ void k_scalar_mul_batch(const uint64_t* __restrict__ points, const uint32_t* __restrict__ scalars, unsigned n,
                                   int mont, uint32_t* __restrict__ out) {
    const unsigned i = blockIdx.x;
    if (i >= n || threadIdx.x != 0) return;
    Affine<F> p = Affine<F>::load(points + uint64_t(i) * (Affine<F>::BYTES / 8));
    if (!mont && !p.is_inf()) {
        p.x = F::to_mont(p.x);
        p.y = F::to_mont(p.y);
    }
    uint32_t k[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) k[q] = scalars[8 * i + q];
    XYZZ<F> r;
    if constexpr (std::is_same<F, FqOps>::value) {
        r = p.is_inf() ? XYZZ<F>::inf() : ec_scalar_mul_glv(p, k);
    } else {
        r = ec_scalar_mul<F>(p, k);
    }
    r.store(out + uint64_t(i) * (XYZZ<F>::BYTES / 4));
}

// Fixed-base multiples from an 8-bit window table T[w][i] = 2^(8 w) P_i (32 levels of n_bases affine points):
// one warp per job, lane w computes  digit_w * T[w][base]  by an 8-step double-and-add, then the 32 partial results
// are tree-summed through shared memory: 8 doublings + 5 additions deep instead of 255 + 127.
//   job j of ballot b:  base = spec[j].base;  scalar = spec[j].sel >= 0 ? host_sc[b][sel] : w[b][-sel]  (canonical Fr)
struct FixedJob {
    int base;
    int sel;
};
constexpr int FIXED_WARPS = 2;
template <class F>
__global__ void __launch_bounds__(32 * FIXED_WARPS) k_fixed_mul(const uint32_t* __restrict__ table, uint64_t n_bases,
                                                                const FixedJob* __restrict__ spec, unsigned jobs_per_ballot,
                                                                unsigned n_jobs, const uint32_t* __restrict__ host_sc,
                                                                unsigned sc_per_ballot, const uint32_t* __restrict__ w,
                                                                uint64_t w_stride, uint32_t* __restrict__ out) {
    extern __shared__ uint4 sm4[];
    constexpr int L = XYZZ<F>::BYTES / 4, AL = Affine<F>::BYTES / 4;
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* sm = reinterpret_cast<uint32_t*>(sm4) + warp * 32 * L;
    const unsigned job = blockIdx.x * FIXED_WARPS + warp;
    const bool live = job < n_jobs;
    XYZZ<F> acc = XYZZ<F>::inf();
    if (live) {
        const unsigned b = job / jobs_per_ballot, j = job - b * jobs_per_ballot;
        const FixedJob fj = spec[j];
        const uint32_t* sp = fj.sel >= 0 ? host_sc + 8 * (uint64_t(b) * sc_per_ballot + fj.sel)
                                         : w + 8 * (uint64_t(b) * w_stride + uint64_t(-fj.sel));
        const uint32_t d = (sp[lane >> 2] >> ((lane & 3) * 8)) & 0xffu;
        if (d) {
            const Affine<F> p = Affine<F>::load(table + (uint64_t(lane) * n_bases + fj.base) * AL);
            for (int bit = 31 - __clz(d); bit >= 0; --bit) {
                acc = ec_dbl<F>(acc);
                if ((d >> bit) & 1u) ec_madd<F>(acc, p);
            }
        }
    }
    acc.store(sm + lane * L);
    __syncwarp();
    for (unsigned st = 16; st > 0; st >>= 1) {
        if (lane < st) {
            ec_add<F>(acc, XYZZ<F>::load(sm + (lane + st) * L));
            acc.store(sm + lane * L);
        }
        __syncwarp();
    }
    if (live && lane == 0) acc.store(out + uint64_t(job) * L);
}

// out[j] (canonical affine) = sum of xyzz[i] with group[i] == j  (+ addend[j], canonical affine, optional)
template <class F>
__global__ void k_group_sum(const uint32_t* __restrict__ xyzz, const uint32_t* __restrict__ group, unsigned n,
                            const uint64_t* __restrict__ addend, unsigned n_groups, uint64_t* __restrict__ out) {
    const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_groups) return;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (unsigned i = 0; i < n; ++i) {
        if (group[i] != j) continue;
        XYZZ<F> q = XYZZ<F>::load(xyzz + uint64_t(i) * (XYZZ<F>::BYTES / 4));
        ec_add<F>(acc, q);
    }
    if (addend) {
        Affine<F> a = Affine<F>::load(addend + uint64_t(j) * (Affine<F>::BYTES / 8));
        if (!a.is_inf()) {
            a.x = F::to_mont(a.x);
            a.y = F::to_mont(a.y);
            ec_madd<F>(acc, a);
        }
    }
    Affine<F> r = ec_to_affine<F>(acc);
    r.x = F::from_mont(r.x);
    r.y = F::from_mont(r.y);
    r.store(out + uint64_t(j) * (Affine<F>::BYTES / 8));
}

template <class F>
VSP_DEV void write_canonical(uint64_t* out, const XYZZ<F>& p, bool lowlat = false) {
    Affine<F> a = ec_to_affine<F>(p, lowlat);
    a.x = F::from_mont(a.x);
    a.y = F::from_mont(a.y);
    a.store(out);
}

// SAVER ciphertext of ballot b from its 3 n + 2 fixed-base multiples (job order of vsp_saver_pub):
//   ct[0] = r delta;  ct[1 + i] = r delta_s[i] + m_i gamma_abc[i];  ct[n + 1] = r delta_sum_s + sum_i m_i t[i]
// thread = (ballot, slot); elgamal_verifiable.hpp:341-352
__global__ void k_ct_sum(const uint32_t* __restrict__ mul, unsigned n, unsigned B, uint64_t* __restrict__ ct) {
    const unsigned q = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned slots = n + 2;
    if (q >= B * slots) return;
    const unsigned b = q / slots, slot = q - b * slots;
    constexpr int L = G1XYZZ::BYTES / 4;
    const uint32_t* jb = mul + uint64_t(b) * (3 * n + 2) * L;
    G1XYZZ acc;
    if (slot == 0) {
        acc = G1XYZZ::load(jb);
    } else if (slot <= n) {
        acc = G1XYZZ::load(jb + uint64_t(slot) * L);
        ec_add<FqOps>(acc, G1XYZZ::load(jb + uint64_t(n + slot) * L));
    } else {
        acc = G1XYZZ::load(jb + uint64_t(3 * n + 1) * L);
        for (unsigned i = 0; i < n; ++i) ec_add<FqOps>(acc, G1XYZZ::load(jb + uint64_t(2 * n + 1 + i) * L));
    }
    write_canonical<FqOps>(ct + uint64_t(q) * 12, acc);
}

// ---------------------------------------------------------------------------------------------
// proof assembly for a chunk of B ballots; lane = ballot, so a warp runs one role in lock step
// ---------------------------------------------------------------------------------------------
struct AsmPtrs {
    const uint32_t* evA;    // B x XYZZ G1
    const uint32_t* evB1;   // B x XYZZ G1
    const uint32_t* evB2;   // B x XYZZ G2
    const uint32_t* evH;    // B x XYZZ G1
    const uint32_t* evL;    // B x XYZZ G1
    const uint32_t* fx1;    // B x 4 XYZZ G1: r d1, s d1, (-rs) d1, r_saver gamma_inv (or infinity)
    const uint32_t* fx2;    // B x XYZZ G2: s d2
    const uint32_t* fixed1; // affine Montgomery: alpha, beta1, delta1
    const uint32_t* fixed2; // affine Montgomery: beta2, delta2
    uint64_t* gab;          // scratch: B x (canonical affine g_A, g1_B) (2 x 12 limbs)
    const uint32_t* sg;     // B x 2 XYZZ G1: s g_A, r g1_B
    uint64_t* proof;        // B x 48 limbs
    unsigned B;
};

// phase 1, role = blockIdx.y:
//   0: g_A  = alpha + evA  + r delta1      -> proof[b][0..12)  and gab[b][0]
//   1: g1_B = beta1 + evB1 + s delta1      -> gab[b][1]
//   2: g_B  = beta2 + evB2 + s delta2      -> proof[b][12..36)
__global__ void __launch_bounds__(32) k_prove_phase1(AsmPtrs p) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    constexpr int L1 = G1XYZZ::BYTES / 4, L2 = G2XYZZ::BYTES / 4, A1 = G1Affine::BYTES / 4;
    if (blockIdx.y == 0) {
        G1XYZZ acc = G1XYZZ::load(p.evA + uint64_t(b) * L1);
        ec_madd<FqOps>(acc, G1Affine::load(p.fixed1));
        ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + uint64_t(4 * b) * L1));
        write_canonical<FqOps>(p.gab + uint64_t(b) * 24, acc, p.B == 1);
        for (int i = 0; i < 12; ++i) p.proof[uint64_t(b) * 48 + i] = p.gab[uint64_t(b) * 24 + i];
    } else if (blockIdx.y == 1) {
        G1XYZZ acc = G1XYZZ::load(p.evB1 + uint64_t(b) * L1);
        ec_madd<FqOps>(acc, G1Affine::load(p.fixed1 + A1));
        ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + uint64_t(4 * b + 1) * L1));
        write_canonical<FqOps>(p.gab + uint64_t(b) * 24 + 12, acc, p.B == 1);
    } else {
        G2XYZZ acc = G2XYZZ::load(p.evB2 + uint64_t(b) * L2);
        ec_madd<Fq2Ops>(acc, G2Affine::load(p.fixed2));
        ec_add<Fq2Ops>(acc, G2XYZZ::load(p.fx2 + uint64_t(b) * L2));
        write_canonical<Fq2Ops>(p.proof + uint64_t(b) * 48 + 12, acc, p.B == 1);
    }
}

// phase 3: g_C = evH + evL + s g_A + r g1_B + (-rs) delta1 + r_saver gamma_inv
__global__ void __launch_bounds__(32) k_prove_phase3(AsmPtrs p) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    constexpr int L1 = G1XYZZ::BYTES / 4;
    G1XYZZ acc = G1XYZZ::load(p.evH + uint64_t(b) * L1);
    ec_add<FqOps>(acc, G1XYZZ::load(p.evL + uint64_t(b) * L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.sg + uint64_t(2 * b) * L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.sg + uint64_t(2 * b + 1) * L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + uint64_t(4 * b + 2) * L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + uint64_t(4 * b + 3) * L1));
    write_canonical<FqOps>(p.proof + uint64_t(b) * 48 + 36, acc, p.B == 1);
}

static void free_dev(void* p) {
    if (p) cudaFree(p);
}

}  // namespace vsp

// ---------------------------------------------------------------------------------------------
// proving key
// ---------------------------------------------------------------------------------------------
extern "C" void vsp_pk_destroy(vsp_pk* pk) {
    if (!pk) return;
    cudaSetDevice(pk->ctx->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 3; ++i) {
        free_dev(pk->d_row[i]);
        free_dev(pk->d_col[i]);
        free_dev(pk->d_val[i]);
    }
    bases_free(&pk->A);
    bases_free(&pk->Bg1);
    bases_free(&pk->Bg2);
    bases_free(&pk->H);
    bases_free(&pk->L);
    bases_free(&pk->fixed1);
    bases_free(&pk->fixed2);
    bases_free(&pk->fx1tab);
    bases_free(&pk->fx2tab);
    free_dev(pk->d_B_idx);
    free_dev(pk->d_long_rows);
    free_dev(pk->d_gpow);
    free_dev(pk->d_ginvpow);
    free_dev(pk->d_zinv);
    delete pk;
}

extern "C" uint64_t vsp_pk_domain_size(const vsp_pk* pk) { return pk ? pk->m : 0; }

static int pk_load_impl(vsp_ctx* ctx, const vsp_proving_key_t* d, const vsp_pk_slice_t* sl, vsp_pk* pk) {
    using namespace hostfr;
    const vsp_r1cs_t& cs = d->cs;
    pk->ctx = ctx;
    pk->ni = cs.num_inputs;
    pk->na = cs.num_aux;
    pk->nc = cs.num_constraints;
    pk->nv = pk->ni + pk->na;
    if (pk->nv + 1 >= (uint64_t(1) << 31)) return set_error(ctx, VSP_E_INVALID, "too many variables");
    if (domain_plan(pk->nc + pk->ni + 1, &pk->kind, &pk->m) != VSP_OK || pk->kind == VSP_DOMAIN_EXTENDED_RADIX2)
        return set_error(ctx, VSP_E_DOMAIN, "no executable evaluation domain for %llu constraints",
                         (unsigned long long)pk->nc);
    // the slice of each query this key holds: everything for a full key
    pk->sliced = sl != nullptr;
    pk->A_lo = sl ? sl->A_lo : 0;
    pk->A_n = sl ? sl->A_hi - sl->A_lo : pk->nv + 1;
    pk->B_lo = sl ? sl->B_lo : 0;
    pk->H_lo = sl ? sl->H_lo : 0;
    pk->H_n = sl ? sl->H_hi - sl->H_lo : pk->m - 1;
    pk->L_lo = sl ? sl->L_lo : 0;
    pk->L_n = sl ? sl->L_hi - sl->L_lo : pk->na;
    if (sl) {
        if (sl->A_lo > sl->A_hi || sl->A_hi > pk->nv + 1 || sl->H_lo > sl->H_hi || sl->H_hi > pk->m - 1 ||
            sl->L_lo > sl->L_hi || sl->L_hi > pk->na || sl->B_lo > sl->B_hi || sl->B_hi - sl->B_lo != d->B_count)
            return set_error(ctx, VSP_E_INVALID, "proving-key slice out of range");
        if (d->H_count != pk->H_n) return set_error(ctx, VSP_E_INVALID, "H_query slice length differs from H_hi - H_lo");
    } else if (d->H_count + 1 < pk->m) {
        return set_error(ctx, VSP_E_INVALID, "H_query shorter than m - 1");
    }
    cudaStream_t s = ctx->stream;
    const uint64_t* rows[3] = {cs.a_row, cs.b_row, cs.c_row};
    const uint32_t* cols[3] = {cs.a_col, cs.b_col, cs.c_col};
    const uint64_t* vals[3] = {cs.a_val, cs.b_val, cs.c_val};
    uint64_t nnz_total = 0;
    for (int i = 0; i < 3; ++i) {
        if (!rows[i]) return set_error(ctx, VSP_E_INVALID, "null CSR row pointer");
        const uint64_t nnz = rows[i][pk->nc];
        nnz_total += nnz;
        for (uint64_t k = 0; k < nnz; ++k)
            if (cols[i][k] > pk->nv) return set_error(ctx, VSP_E_INVALID, "constraint refers to variable %u > %llu",
                                                      cols[i][k], (unsigned long long)pk->nv);
        VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_row[i], (pk->nc + 1) * 8));
        VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_col[i], (nnz ? nnz : 1) * 4));
        VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_val[i], (nnz ? nnz : 1) * 32));
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(pk->d_row[i], rows[i], (pk->nc + 1) * 8, cudaMemcpyHostToDevice, s));
        if (nnz) {
            VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(pk->d_col[i], cols[i], nnz * 4, cudaMemcpyHostToDevice, s));
            VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(pk->d_val[i], vals[i], nnz * 32, cudaMemcpyHostToDevice, s));
            VSP_LAUNCH_K(ctx, KC_QAP, k_vals_to_mont, (unsigned)((nnz + 127) / 128), 128, 0, s, pk->d_val[i], nnz, i == 1 ? 1 : 0);
            VSP_CUDA_CHECK(ctx, cudaGetLastError());
        }
    }
    {
        std::vector<uint32_t> lr;
        for (uint64_t r = 0; r < pk->nc; ++r)
            for (int i = 0; i < 3; ++i)
                if (rows[i][r + 1] - rows[i][r] > LONG_ROW_TERMS) {
                    lr.push_back((uint32_t)r);
                    break;
                }
        pk->n_long = (unsigned)lr.size();
        if (pk->n_long > 4096) return set_error(ctx, VSP_E_INVALID, "too many constraints with more than %llu terms",
                                                (unsigned long long)LONG_ROW_TERMS);
        if (pk->n_long) {
            VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_long_rows, lr.size() * 4));
            VSP_CUDA_CHECK(ctx, cudaMemcpy(pk->d_long_rows, lr.data(), lr.size() * 4, cudaMemcpyHostToDevice));
        }
    }
    pk->short_rows = nnz_total <= 24 * pk->nc && !getenv("VSP_R1CS_WARP_ROWS");  // env: force the warp-per-row kernel (tests)
    // window tables pay while a query is small enough that the bucket reduction would otherwise rival the bucket
    // additions (a ballot circuit: 2^16 points); from VSP_TABLE_MAX_LOG (default 2^22) points up a query stays table-less:
    // 13+ levels of a 2^24-point table are 20+ GB and half a minute of doublings for a few percent of multiexp time
    unsigned table_max_log = 22;
    if (const char* e = getenv("VSP_TABLE_MAX_LOG")) table_max_log = (unsigned)atoi(e);
    auto want_table = [&](uint64_t n) { return table_max_log < 63 && n <= (uint64_t(1) << table_max_log); };
    VSP_TRY(bases_upload(ctx, false, pk->A_n, d->A_query, &pk->A, want_table(pk->A_n)));
    VSP_TRY(bases_upload(ctx, false, d->B_count, d->B_g1, &pk->Bg1, want_table(d->B_count)));
    VSP_TRY(bases_upload(ctx, true, d->B_count, d->B_g2, &pk->Bg2, want_table(d->B_count)));
    VSP_TRY(bases_upload(ctx, false, pk->H_n, d->H_query, &pk->H, want_table(pk->H_n)));
    VSP_TRY(bases_upload(ctx, false, pk->L_n, d->L_query, &pk->L, want_table(pk->L_n)));
    pk->B_count = d->B_count;
    for (uint64_t k = 0; k < d->B_count; ++k)
        if (d->B_idx[k] > pk->nv) return set_error(ctx, VSP_E_INVALID, "B_query index out of range");
    VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_B_idx, (d->B_count ? d->B_count : 1) * 4));
    if (d->B_count)
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(pk->d_B_idx, d->B_idx, d->B_count * 4, cudaMemcpyHostToDevice, s));
    uint64_t f1[36], f2[48];
    memcpy(f1, d->alpha_g1, 96);
    memcpy(f1 + 12, d->beta_g1, 96);
    memcpy(f1 + 24, d->delta_g1, 96);
    memcpy(f2, d->beta_g2, 192);
    memcpy(f2 + 24, d->delta_g2, 192);
    memcpy(pk->h_delta_g1, d->delta_g1, 96);
    memcpy(pk->h_delta_g2, d->delta_g2, 192);
    VSP_TRY(bases_upload(ctx, false, 3, f1, &pk->fixed1, false));
    VSP_TRY(bases_upload(ctx, true, 2, f2, &pk->fixed2, false));
    // 8-bit fixed-base tables for the r / s multiples of delta (gamma_inverse_sum_s_g1 joins fx1tab on first use)
    uint64_t t1[24];
    memcpy(t1, d->delta_g1, 96);
    memset(t1 + 12, 0, 96);
    pk->fx1tab.force_c = 8;
    pk->fx2tab.force_c = 8;
    VSP_TRY(bases_upload(ctx, false, 2, t1, &pk->fx1tab, true));
    VSP_TRY(bases_upload(ctx, true, 1, d->delta_g2, &pk->fx2tab, true));
    if (!pk->fx1tab.table || !pk->fx2tab.table) return set_error(ctx, VSP_E_NOMEM, "fixed-base tables not built");

    // coset tables (multiplicative_generator = 7)
    const Fr4 g = from_u64(7), ginv = inv(g);
    VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_gpow, pk->m * 32));
    VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_ginvpow, pk->m * 32));
    VSP_TRY(fill_power_table(ctx, s, pk->d_gpow, pk->m, g, from_u64(1)));
    if (pk->kind == VSP_DOMAIN_BASIC_RADIX2) {
        const Fr4 zinv = inv(sub(pow_u64(g, pk->m), from_u64(1)));
        VSP_TRY(fill_power_table(ctx, s, pk->d_ginvpow, pk->m, ginv, zinv));
    } else {
        VSP_TRY(fill_power_table(ctx, s, pk->d_ginvpow, pk->m, ginv, from_u64(1)));
        // per-element 1/Z on the coset: run divide_by_z_on_coset over a vector of Montgomery ones
        VSP_CUDA_CHECK(ctx, cudaMalloc(&pk->d_zinv, pk->m * 32));
        Fr one_mont;
        memcpy(one_mont.l, to_mont(from_u64(1)).v, 32);
        VSP_LAUNCH_K(ctx, KC_QAP, k_fill_const, (unsigned)((pk->m + 127) / 128), 128, 0, s, pk->d_zinv, pk->m, one_mont);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        VSP_TRY(divide_by_z_on_coset_dev(ctx, s, pk->m, pk->d_zinv));
    }
    VSP_TRY(ntt_tables_ensure(ctx, ceil_log2_u64(pk->m)));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

extern "C" int vsp_pk_load(vsp_ctx* ctx, const vsp_proving_key_t* desc, vsp_pk** out) {
    if (!ctx || !desc || !out) return VSP_E_INVALID;
    *out = nullptr;
    cudaSetDevice(ctx->device);
    vsp_pk* pk = new (std::nothrow) vsp_pk();
    if (!pk) return VSP_E_NOMEM;
    int rc = pk_load_impl(ctx, desc, nullptr, pk);
    if (rc) {
        pk->ctx = ctx;
        vsp_pk_destroy(pk);
        return rc;
    }
    *out = pk;
    return VSP_OK;
}

extern "C" int vsp_pk_load_slice(vsp_ctx* ctx, const vsp_proving_key_t* desc, const vsp_pk_slice_t* slice, vsp_pk** out) {
    if (!ctx || !desc || !slice || !out) return VSP_E_INVALID;
    *out = nullptr;
    cudaSetDevice(ctx->device);
    vsp_pk* pk = new (std::nothrow) vsp_pk();
    if (!pk) return VSP_E_NOMEM;
    int rc = pk_load_impl(ctx, desc, slice, pk);
    if (rc) {
        pk->ctx = ctx;
        vsp_pk_destroy(pk);
        return rc;
    }
    *out = pk;
    return VSP_OK;
}


// ---------------------------------------------------------------------------------------------
// witness map
// ---------------------------------------------------------------------------------------------
namespace vsp {

// `nvec` packed vectors of m values: inverse_fft, multiply_by_coset(g), fft
static int coset_lde(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, uint32_t* d_a, unsigned nvec) {
    const uint64_t m = pk->m;
    if (pk->kind == VSP_DOMAIN_BASIC_RADIX2) {
        const unsigned L = ceil_log2_u64(m);
        VSP_TRY(ntt_pow2(ctx, s, d_a, d_a, L, true, nullptr, pk->d_gpow, nvec));  // inverse_fft then multiply_by_coset(g)
        VSP_TRY(ntt_pow2(ctx, s, d_a, d_a, L, false, nullptr, nullptr, nvec));    // fft
    } else {
        VSP_TRY(domain_fft_dev(ctx, s, m, true, d_a, nvec));
        VSP_LAUNCH_K(ctx, KC_QAP, k_mul_table2, dim3((unsigned)((m + 127) / 128), nvec), 128, 0, s, d_a, m, pk->d_gpow);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        VSP_TRY(domain_fft_dev(ctx, s, m, false, d_a, nvec));
    }
    return VSP_OK;
}

// d_w: B assignments (1 || primary || aux), canonical, w_stride elements apart.  abc: 3 B m Fr of scratch; on return
// abc[0 .. B m) holds coefficients_for_H[0..m) of every ballot (canonical).  Everything on stream s.
// The map in two halves, so that a proof sharded over several GPUs can split the three independent coset extensions
// (inverse_fft, multiply_by_coset, fft of aA, aB, aC) between ranks and exchange the results:
//   witness_map_evaluate  constraint evaluation for all three, coset extension of the vectors in `mask` (bit v: vector v)
//   witness_map_finish    H = (A o B - C) / Z on the coset, back to coefficients
static int witness_map_evaluate(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, const uint32_t* d_w, uint64_t w_stride,
                                unsigned B, uint32_t* abc, unsigned mask);
static int witness_map_finish(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, unsigned B, uint32_t* abc);
static int witness_map_dev(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, const uint32_t* d_w, uint64_t w_stride,
                           unsigned B, uint32_t* abc) {
    VSP_TRY(witness_map_evaluate(ctx, s, pk, d_w, w_stride, B, abc, 7u));
    return witness_map_finish(ctx, s, pk, B, abc);
}
static int witness_map_evaluate(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, const uint32_t* d_w, uint64_t w_stride,
                                unsigned B, uint32_t* abc, unsigned mask) {
    const uint64_t m = pk->m;
    CsrDev csr;
    for (int i = 0; i < 3; ++i) {
        csr.row[i] = pk->d_row[i];
        csr.col[i] = pk->d_col[i];
        csr.val[i] = pk->d_val[i];
    }
    if (pk->short_rows)
        VSP_LAUNCH_K(ctx, KC_QAP, k_r1cs_eval_rows, dim3((unsigned)((m + 127) / 128), B), 128, 0, s, csr, d_w, w_stride,
                     pk->nc, pk->ni, m, abc);
    else
        VSP_LAUNCH_K(ctx, KC_QAP, k_r1cs_eval, dim3((unsigned)((m * 32 + 255) / 256), B), 256, 0, s, csr, d_w, w_stride,
                     pk->nc, pk->ni, m, abc);
    DevBuf long_part;
    if (pk->n_long) {
        VSP_TRY(long_part.alloc(ctx, size_t(B) * pk->n_long * 3 * LONG_SLICES * 32, s));
        VSP_LAUNCH_K(ctx, KC_QAP, k_r1cs_long_partial, dim3(LONG_SLICES, pk->n_long * 3, B), 256, 0, s, csr, pk->d_long_rows, d_w,
                     w_stride, long_part.as<uint32_t>());
        VSP_LAUNCH_K(ctx, KC_QAP, k_r1cs_long_final, (B * pk->n_long * 3 + 63) / 64, 64, 0, s, csr, pk->d_long_rows, pk->n_long, B,
                     m, long_part.as<uint32_t>(), abc);
    }
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    if ((mask & 7u) == 7u) return coset_lde(ctx, s, pk, abc, 3 * B);
    for (unsigned v = 0; v < 3; ++v)
        if (mask & (1u << v)) VSP_TRY(coset_lde(ctx, s, pk, abc + 8 * (uint64_t(v) * B * m), B));
    return VSP_OK;
}
static int witness_map_finish(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, unsigned B, uint32_t* abc) {
    const uint64_t m = pk->m;
    uint32_t *aA = abc, *aB = abc + 8 * (uint64_t(B) * m), *aC = abc + 8 * (uint64_t(2 * B) * m);
    VSP_LAUNCH_K(ctx, KC_QAP, k_h_pointwise, dim3((unsigned)((m + 127) / 128), B), 128, 0, s, aA, aB, aC, pk->d_zinv, m);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    if (pk->kind == VSP_DOMAIN_BASIC_RADIX2) {
        // inverse_fft, multiply_by_coset(g^-1) and divide_by_z_on_coset in one pass chain (table = Z^-1 g^-i)
        VSP_TRY(ntt_pow2(ctx, s, aA, aA, ceil_log2_u64(m), true, nullptr, pk->d_ginvpow, B));
    } else {
        VSP_TRY(domain_fft_dev(ctx, s, m, true, aA, B));
        VSP_LAUNCH_K(ctx, KC_QAP, k_mul_table2, dim3((unsigned)((m + 127) / 128), B), 128, 0, s, aA, m, pk->d_ginvpow);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
    }
    return VSP_OK;
}

// B host witnesses -> packed device assignments (1 || primary || aux), stride nv + 1
static int upload_witness(vsp_ctx* ctx, const vsp_pk* pk, unsigned B, const uint64_t* primary, const uint64_t* aux,
                          DevBuf& d_w) {
    cudaStream_t s = ctx->stream;
    const uint64_t stride = pk->nv + 1;
    VSP_TRY(d_w.alloc(ctx, uint64_t(B) * stride * 32, s));
    VSP_LAUNCH_K(ctx, KC_QAP, k_set_one, (B + 127) / 128, 128, 0, s, d_w.as<uint32_t>(), stride, B);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    if (pk->ni)
        VSP_CUDA_CHECK(ctx, cudaMemcpy2DAsync(d_w.as<char>() + 32, stride * 32, primary, pk->ni * 32, pk->ni * 32, B,
                                              cudaMemcpyHostToDevice, s));
    if (pk->na)
        VSP_CUDA_CHECK(ctx, cudaMemcpy2DAsync(d_w.as<char>() + 32 + pk->ni * 32, stride * 32, aux, pk->na * 32, pk->na * 32,
                                              B, cudaMemcpyHostToDevice, s));
    return VSP_OK;
}

// Host witnesses of a batch, one chunk ahead: chunk c + 1 is copied on the context's copy stream into the other of two
// device buffers while chunk c is being proved, so the H2D transfer (2 MB per ballot at m = 2^16) leaves the critical
// path of the host-buffer entry points.  acquire(c) returns chunk c's assignments once main-stream work may read them
// (the wait is a stream dependency, not a host block); release(c) marks the buffer reusable after main-stream order.
struct WitnessPrefetch {
    vsp_ctx* ctx;
    const vsp_pk* pk;
    const uint64_t *primary, *aux;
    uint64_t k;
    unsigned chunk;
    DevBuf buf[2];
    cudaEvent_t up[2] = {nullptr, nullptr}, freed[2] = {nullptr, nullptr};
    bool used[2] = {false, false};
    cudaStream_t cs = nullptr;
    uint64_t issued = 0;  // chunks whose upload has been enqueued

    WitnessPrefetch(vsp_ctx* c, const vsp_pk* p, const uint64_t* pr, const uint64_t* ax, uint64_t kk, unsigned ch)
        : ctx(c), pk(p), primary(pr), aux(ax), k(kk), chunk(ch) {}
    ~WitnessPrefetch() {
        for (int i = 0; i < 2; ++i) {
            if (up[i]) cudaEventDestroy(up[i]);
            if (freed[i]) cudaEventDestroy(freed[i]);
        }
    }
    uint64_t n_chunks() const { return (k + chunk - 1) / chunk; }
    int init() {
        if (!ctx->copy && cudaStreamCreateWithFlags(&ctx->copy, cudaStreamNonBlocking) != cudaSuccess)
            return set_error(ctx, VSP_E_CUDA, "cudaStreamCreate(copy) failed");
        cs = ctx->prof.on ? ctx->stream : ctx->copy;
        const uint64_t stride = pk->nv + 1;
        const unsigned nbuf = n_chunks() > 1 ? 2 : 1;
        for (unsigned i = 0; i < nbuf; ++i) {
            const uint64_t B = (i == 0) ? (k < chunk ? k : chunk) : chunk;
            VSP_TRY(buf[i].alloc(ctx, B * stride * 32, ctx->stream));
            VSP_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&up[i], cudaEventDisableTiming));
            VSP_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&freed[i], cudaEventDisableTiming));
        }
        // the buffers come from the main stream's pool: order the copy stream after the allocation
        VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_main, ctx->stream));
        VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(cs, ctx->ev_main, 0));
        return issue();
    }
    int issue() {
        if (issued >= n_chunks()) return VSP_OK;
        const uint64_t c = issued, i0 = c * chunk;
        const unsigned slot = (unsigned)(c & 1), B = (unsigned)((k - i0 < chunk) ? (k - i0) : chunk);
        const uint64_t stride = pk->nv + 1;
        if (used[slot]) VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(cs, freed[slot], 0));
        char* d = buf[slot].as<char>();
        VSP_LAUNCH_K(ctx, KC_QAP, k_set_one, (B + 127) / 128, 128, 0, cs, buf[slot].as<uint32_t>(), stride, B);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        if (pk->ni)
            VSP_CUDA_CHECK(ctx, cudaMemcpy2DAsync(d + 32, stride * 32, primary + 4 * pk->ni * i0, pk->ni * 32, pk->ni * 32, B,
                                                  cudaMemcpyHostToDevice, cs));
        if (pk->na)
            VSP_CUDA_CHECK(ctx, cudaMemcpy2DAsync(d + 32 + pk->ni * 32, stride * 32, aux + 4 * pk->na * i0, pk->na * 32,
                                                  pk->na * 32, B, cudaMemcpyHostToDevice, cs));
        VSP_CUDA_CHECK(ctx, cudaEventRecord(up[slot], cs));
        used[slot] = true;
        ++issued;
        return VSP_OK;
    }
    // chunk c's assignments; also enqueues the upload of chunk c + 1
    int acquire(uint64_t c, const uint32_t** w) {
        const unsigned slot = (unsigned)(c & 1);
        VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, up[slot], 0));
        *w = buf[slot].as<uint32_t>();
        return VSP_OK;
    }
    int release(uint64_t c) {
        VSP_CUDA_CHECK(ctx, cudaEventRecord(freed[c & 1], ctx->stream));
        return issue();  // the next-but-one chunk reuses this buffer
    }
};

static bool fr_reduced(const uint64_t* x) { return !hostfr::geq(x, hostfr::MOD); }

// make fx1tab = {delta_g1, gamma_inv}: rebuilt only when a different gamma_inverse_sum_s_g1 arrives
static int ensure_gamma_table(vsp_ctx* ctx, vsp_pk* pk, const uint64_t* gamma_inv) {
    if (memcmp(pk->gamma_key, gamma_inv, 96) == 0) return VSP_OK;
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 4; ++i) VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->side[i]));
    uint64_t t1[24];
    memcpy(t1, pk->h_delta_g1, 96);
    memcpy(t1 + 12, gamma_inv, 96);
    BasesDev nb;
    nb.force_c = 8;
    VSP_TRY(bases_upload(ctx, false, 2, t1, &nb, true));
    if (!nb.table) {
        bases_free(&nb);
        return set_error(ctx, VSP_E_NOMEM, "fixed-base table not built");
    }
    bases_free(&pk->fx1tab);
    pk->fx1tab = nb;
    memcpy(pk->gamma_key, gamma_inv, 96);
    return VSP_OK;
}

// side stream i — or the main stream while per-launch profiling is on, so that timed launches do not overlap
static cudaStream_t side_stream(vsp_ctx* ctx, int i) { return ctx->prof.on ? ctx->stream : ctx->side[i]; }

// device constants of the proof's fixed-base jobs: (base, scalar slot) x 4 over fx1tab, x 1 over fx2tab.
// per-ballot host scalars: 0 r | 1 s | 2 -rs | 3 r_saver (0 without encrypted_input)
static const FixedJob PROOF_JOBS_G1[4] = {{0, 0}, {0, 1}, {0, 2}, {1, 3}};
static const FixedJob PROOF_JOBS_G2[1] = {{0, 1}};

// pinned staging slot of at least `bytes` for host data about to be uploaded on a stream (vsp_ctx::StageSlot)
static int stage_acquire(vsp_ctx* ctx, size_t bytes, vsp_ctx::StageSlot** out) {
    vsp_ctx::StageSlot& sl = ctx->stage[ctx->stage_next++ & 3];
    if (sl.pending) {
        VSP_CUDA_CHECK(ctx, cudaEventSynchronize(sl.done));
        sl.pending = false;
    }
    if (sl.cap < bytes) {
        if (sl.p) cudaFreeHost(sl.p);
        sl.p = nullptr;
        sl.cap = 0;
        VSP_CUDA_CHECK(ctx, cudaMallocHost(&sl.p, bytes));
        sl.cap = bytes;
    }
    if (!sl.done) VSP_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    *out = &sl;
    return VSP_OK;
}
// the copies out of the slot have been enqueued on `s`
static int stage_commit(vsp_ctx* ctx, vsp_ctx::StageSlot* sl, cudaStream_t s) {
    VSP_CUDA_CHECK(ctx, cudaEventRecord(sl->done, s));
    sl->pending = true;
    return VSP_OK;
}

// One chunk of B proofs, fully asynchronous: on return everything is enqueued on ctx->stream / side streams, joined
// back into ctx->stream, and d_proofs (B x 48 limbs) holds the results once ctx->stream completes.
//   d_w: B assignments, w_stride apart.  rs: B x (r, s, r_saver) host scalars.
static int prove_chunk(vsp_ctx* ctx, vsp_pk* pk, unsigned B, const uint32_t* d_w, uint64_t w_stride, const uint64_t* rs,
                       const uint64_t* gamma_inv, uint64_t* d_proofs) {
    using namespace hostfr;
    cudaStream_t s = ctx->stream;
    const uint64_t m = pk->m;
    constexpr size_t X1 = G1XYZZ::BYTES, X2 = G2XYZZ::BYTES;
    if (pk->sliced)
        return set_error(ctx, VSP_E_INVALID, "this key holds a point-range slice: use vsp_prove_partial / vsp_prove_assemble");
    if (gamma_inv) VSP_TRY(ensure_gamma_table(ctx, pk, gamma_inv));

    // host scalars of the chunk: fixed-base slots (4 per ballot), the variable-base pair (s, r) and the two job tables,
    // written straight into pinned memory
    const size_t fsc_n = size_t(B) * 16, vsc_n = size_t(B) * 8;
    vsp_ctx::StageSlot* stage = nullptr;
    VSP_TRY(stage_acquire(ctx, (fsc_n + vsc_n) * 8 + sizeof(PROOF_JOBS_G1) + sizeof(PROOF_JOBS_G2), &stage));
    uint64_t* fsc = static_cast<uint64_t*>(stage->p);
    uint64_t* vsc = fsc + fsc_n;
    char* host_jobs = reinterpret_cast<char*>(vsc + vsc_n);
    memcpy(host_jobs, PROOF_JOBS_G1, sizeof(PROOF_JOBS_G1));
    memcpy(host_jobs + sizeof(PROOF_JOBS_G1), PROOF_JOBS_G2, sizeof(PROOF_JOBS_G2));
    for (unsigned b = 0; b < B; ++b) {
        const uint64_t *r = rs + 12 * b, *sv = r + 4, *rsv = r + 8;
        if (!fr_reduced(r) || !fr_reduced(sv) || (gamma_inv && !fr_reduced(rsv)))
            return set_error(ctx, VSP_E_INVALID, "proving randomness not reduced mod r");
        Fr4 fr, fs;
        memcpy(fr.v, r, 32);
        memcpy(fs.v, sv, 32);
        const Fr4 frs = sub(from_u64(0), mul(fr, fs));
        uint64_t* f = &fsc[size_t(b) * 16];
        memcpy(f, r, 32);
        memcpy(f + 4, sv, 32);
        memcpy(f + 8, frs.v, 32);
        if (gamma_inv) memcpy(f + 12, rsv, 32); else memset(f + 12, 0, 32);
        memcpy(&vsc[size_t(b) * 8], sv, 32);     // s * g_A
        memcpy(&vsc[size_t(b) * 8 + 4], r, 32);  // r * g1_B
    }
    // tail scratch (MSM results, fixed-base multiples, g_A / g1_B, their multiples, the variable-base scalars): one of
    // two persistent slots, so this chunk's assembly tail can still be running while the next chunk is enqueued
    vsp_ctx::TailSlot& slot = ctx->tail[ctx->tail_next++ & 1];
    const size_t per_ballot = (4 * X1 + X2) + 4 * X1 + X2 + 192 + 2 * X1 + 64;
    if (slot.pending) VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, slot.done, 0));  // the slot's previous tail
    if (slot.cap < per_ballot * B) {
        if (slot.pending) VSP_CUDA_CHECK(ctx, cudaEventSynchronize(slot.done));
        if (slot.p) cudaFree(slot.p);
        slot.p = nullptr;
        slot.cap = 0;
        VSP_CUDA_CHECK(ctx, cudaMalloc(&slot.p, per_ballot * B));
        slot.cap = per_ballot * B;
    }
    char* tp = reinterpret_cast<char*>(slot.p);
    uint64_t* evA = reinterpret_cast<uint64_t*>(tp);  // evA | evB1 | evH | evL (G1) | evB2 (G2), B each
    uint64_t* evB1 = evA + size_t(B) * X1 / 8;
    uint64_t* evH = evB1 + size_t(B) * X1 / 8;
    uint64_t* evL = evH + size_t(B) * X1 / 8;
    uint64_t* evB2 = evL + size_t(B) * X1 / 8;
    tp += size_t(B) * (4 * X1 + X2);
    uint32_t* fx1 = reinterpret_cast<uint32_t*>(tp);
    tp += size_t(B) * 4 * X1;
    uint32_t* fx2 = reinterpret_cast<uint32_t*>(tp);
    tp += size_t(B) * X2;
    uint64_t* gab = reinterpret_cast<uint64_t*>(tp);
    tp += size_t(B) * 192;
    uint32_t* sg = reinterpret_cast<uint32_t*>(tp);
    tp += size_t(B) * 2 * X1;
    uint32_t* d_vsc = reinterpret_cast<uint32_t*>(tp);
    DevBuf abc, bsc, d_fsc, jobs;
    VSP_TRY(bsc.alloc(ctx, size_t(B) * (pk->B_count ? pk->B_count : 1) * 32, s));
    VSP_TRY(d_fsc.alloc(ctx, fsc_n * 8, s));
    VSP_TRY(jobs.alloc(ctx, sizeof(PROOF_JOBS_G1) + sizeof(PROOF_JOBS_G2), s));
    VSP_TRY(abc.alloc(ctx, size_t(3) * B * m * 32, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_fsc.p, fsc, fsc_n * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_vsc, vsc, vsc_n * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(jobs.p, host_jobs, sizeof(PROOF_JOBS_G1) + sizeof(PROOF_JOBS_G2), cudaMemcpyHostToDevice, s));
    VSP_TRY(stage_commit(ctx, stage, s));
    const FixedJob* jobs1 = jobs.as<FixedJob>();
    const FixedJob* jobs2 = jobs1 + 4;

    // fork: everything that only needs the witness and the scalars
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_fork, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 2), ctx->ev_fork, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 3), ctx->ev_fork, 0));
    // After the fork a failure must not return before the main stream has joined the side streams: the DevBuf
    // destructors release abc / bsc / d_fsc / jobs in main-stream order, and a side stream may still be reading them.
    struct ForkGuard {
        vsp_ctx* c;
        bool armed = true;
        ~ForkGuard() {
            if (!armed) return;
            for (int i = 2; i <= 3; ++i) {
                cudaStream_t ss = side_stream(c, i);
                if (ss == c->stream) continue;
                if (cudaEventRecord(c->ev_join[i], ss) == cudaSuccess) cudaStreamWaitEvent(c->stream, c->ev_join[i], 0);
            }
        }
    } guard{ctx};
    // side 2: A, L
    VSP_TRY(msm_run(ctx, side_stream(ctx, 2), pk->A, pk->nv + 1, d_w, evA, true, B, w_stride, X1 / 8));
    VSP_TRY(msm_run(ctx, side_stream(ctx, 2), pk->L, pk->na, d_w + 8 * (pk->ni + 1), evL, true, B, w_stride, X1 / 8));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[2], side_stream(ctx, 2)));
    // side 3: fixed-base multiples, B (G2), B (G1)
    cudaStream_t s3 = side_stream(ctx, 3);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<Fq2Ops>, (B + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS,
                 FIXED_WARPS * 32 * X2, s3, pk->fx2tab.d_pts, pk->fx2tab.n, jobs2, 1u, B, d_fsc.as<uint32_t>(), 4u,
                 (const uint32_t*)nullptr, uint64_t(0), fx2);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<FqOps>, (4 * B + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS,
                 FIXED_WARPS * 32 * X1, s3, pk->fx1tab.d_pts, pk->fx1tab.n, jobs1, 4u, 4 * B, d_fsc.as<uint32_t>(), 4u,
                 (const uint32_t*)nullptr, uint64_t(0), fx1);
    if (pk->B_count)
        VSP_LAUNCH_K(ctx, KC_QAP, k_gather_fr, dim3((unsigned)((pk->B_count + 127) / 128), B), 128, 0, s3, d_w, w_stride,
                     pk->d_B_idx, pk->B_count, bsc.as<uint32_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_TRY(msm_run(ctx, s3, pk->Bg2, pk->B_count, bsc.as<uint32_t>(), evB2, true, B, pk->B_count, X2 / 8));
    VSP_TRY(msm_run(ctx, s3, pk->Bg1, pk->B_count, bsc.as<uint32_t>(), evB1, true, B, pk->B_count, X1 / 8));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[3], s3));

    // main: witness map, H multiexp
    VSP_TRY(witness_map_dev(ctx, s, pk, d_w, w_stride, B, abc.as<uint32_t>()));
    VSP_TRY(msm_run(ctx, s, pk->H, m - 1, abc.as<uint32_t>(), evH, true, B, m, X1 / 8));
    // the assembly tail runs on side[1] once the five multiexps are done; the main stream only joins the side
    // streams (whose scratch it owns) and moves on to the next chunk
    cudaStream_t st = side_stream(ctx, 1);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_main, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(st, ctx->ev_main, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(st, ctx->ev_join[2], 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(st, ctx->ev_join[3], 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[2], 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[3], 0));
    guard.armed = false;  // joined

    AsmPtrs ap;
    ap.evA = reinterpret_cast<uint32_t*>(evA);
    ap.evB1 = reinterpret_cast<uint32_t*>(evB1);
    ap.evB2 = reinterpret_cast<uint32_t*>(evB2);
    ap.evH = reinterpret_cast<uint32_t*>(evH);
    ap.evL = reinterpret_cast<uint32_t*>(evL);
    ap.fx1 = fx1;
    ap.fx2 = fx2;
    ap.fixed1 = pk->fixed1.d_pts;
    ap.fixed2 = pk->fixed2.d_pts;
    ap.gab = gab;
    ap.sg = sg;
    ap.proof = d_proofs;
    ap.B = B;
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase1, dim3((B + 31) / 32, 3), 32, 0, st, ap);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, 2 * B, 32, 0, st, gab, d_vsc, 2 * B, 0, sg);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase3, (B + 31) / 32, 32, 0, st, ap);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaEventRecord(slot.done, st));
    slot.pending = true;
    return VSP_OK;  // the caller joins the tails (join_tails) before it hands results back
}

// make ctx->stream wait for every assembly tail still in flight
static int join_tails(vsp_ctx* ctx) {
    for (int i = 0; i < 2; ++i)
        if (ctx->tail[i].pending) VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->tail[i].done, 0));
    return VSP_OK;
}

// ballots per chunk: vsp_ctx_set_batch_chunk, else VSP_BATCH_CHUNK (read on every call), else 256 (measured on the ballot
// circuit: 1,384 / 1,417 / 1,443 ballots/s at 64 / 128 / 256 per chunk, 1,360 / 1,393 / 1,403 from host buffers); then clamped by the
// domain size so that a chunk's scratch (about 100 B per ballot and domain element, plus the multiexp keys) stays a few
// GB whatever the circuit: 256 ballots at m = 2^15, 128 at 2^16, 8 at 2^20, 1 from 2^23 up
static unsigned batch_chunk(const vsp_ctx* ctx, const vsp_pk* pk) {
    long x = ctx->batch_chunk;
    if (x <= 0) {
        const char* e = getenv("VSP_BATCH_CHUNK");
        x = e ? atol(e) : 256;
    }
    x = x < 1 ? 1 : (x > 1024 ? 1024 : x);
    const uint64_t cap = std::max<uint64_t>(1, (uint64_t(1) << 23) / (pk->m ? pk->m : 1));
    return (unsigned)std::min<uint64_t>((uint64_t)x, cap);
}

}  // namespace vsp

extern "C" int vsp_witness_map(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux,
                               uint64_t* h_out) {
    if (!ctx || !pk || !h_out || (!primary && pk->ni) || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf d_w, abc;
    VSP_TRY(upload_witness(ctx, pk, 1, primary, aux, d_w));
    VSP_TRY(abc.alloc(ctx, 3 * pk->m * 32, ctx->stream));
    VSP_TRY(witness_map_dev(ctx, ctx->stream, pk, d_w.as<uint32_t>(), pk->nv + 1, 1, abc.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(h_out, abc.p, pk->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    memset(h_out + 4 * pk->m, 0, 32);  // coefficients_for_H[m] = 0 for d1 = d2 = d3 = 0
    return VSP_OK;
}

// k ballots in chunks; witnesses either on the host (primary / aux, ballot-major) or on the device (d_assignments)
static int prove_batch_impl(vsp_ctx* ctx, const vsp_pk* cpk, uint64_t k, const uint64_t* primary, const uint64_t* aux,
                            const uint64_t* d_assignments, const uint64_t* rs, const uint64_t* gamma_inverse_sum_s_g1,
                            uint64_t* d_proofs) {
    vsp_pk* pk = const_cast<vsp_pk*>(cpk);  // the gamma_inv table cache is the only mutable part
    const uint64_t stride = pk->nv + 1;
    const unsigned chunk = batch_chunk(ctx, pk);
    WitnessPrefetch pre(ctx, pk, primary, aux, k, chunk);
    if (!d_assignments) VSP_TRY(pre.init());
    for (uint64_t i = 0, c = 0; i < k; i += chunk, ++c) {
        const unsigned B = (unsigned)((k - i < chunk) ? (k - i) : chunk);
        const uint32_t* w;
        if (d_assignments)
            w = reinterpret_cast<const uint32_t*>(d_assignments + 4 * stride * i);
        else
            VSP_TRY(pre.acquire(c, &w));
        VSP_TRY(prove_chunk(ctx, pk, B, w, stride, rs + 12 * i, gamma_inverse_sum_s_g1, d_proofs + 48 * i));
        if (!d_assignments) VSP_TRY(pre.release(c));
    }
    return join_tails(ctx);
}

extern "C" int vsp_prove_batch(vsp_ctx* ctx, const vsp_pk* pk, uint64_t k, const uint64_t* primary, const uint64_t* aux,
                               const uint64_t* rs, const uint64_t* gamma_inverse_sum_s_g1, uint64_t* out_proofs) {
    if (!ctx || !pk || !rs || !out_proofs || (!primary && pk->ni) || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    if (k == 0) return VSP_OK;
    DevBuf d_proofs;
    VSP_TRY(d_proofs.alloc(ctx, k * 384, ctx->stream));
    VSP_TRY(prove_batch_impl(ctx, pk, k, primary, aux, nullptr, rs, gamma_inverse_sum_s_g1, d_proofs.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_proofs, d_proofs.p, k * 384, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

extern "C" int vsp_prove(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux,
                         const uint64_t r[4], const uint64_t s[4], const uint64_t* r_saver,
                         const uint64_t* gamma_inverse_sum_s_g1, uint64_t* out_proof) {
    if (!ctx || !pk || !r || !s || !out_proof) return VSP_E_INVALID;
    const bool enc = r_saver && gamma_inverse_sum_s_g1;
    uint64_t rs[12] = {0};
    memcpy(rs, r, 32);
    memcpy(rs + 4, s, 32);
    if (enc) memcpy(rs + 8, r_saver, 32);
    return vsp_prove_batch(ctx, pk, 1, primary, aux, rs, enc ? gamma_inverse_sum_s_g1 : nullptr, out_proof);
}

// device-resident variant: d_assignments = k full assignments (1 || primary || auxiliary), (num_inputs + num_aux + 1)
// canonical Fr each; d_proofs = k * 48 limbs on the device.  Stream-ordered: returns once everything is enqueued.
extern "C" int vsp_prove_batch_dev(vsp_ctx* ctx, const vsp_pk* pk, uint64_t k, const uint64_t* d_assignments,
                                   const uint64_t* rs, const uint64_t* gamma_inverse_sum_s_g1, uint64_t* d_proofs) {
    if (!ctx || !pk || !rs || !d_proofs || !d_assignments) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return prove_batch_impl(ctx, pk, k, nullptr, nullptr, d_assignments, rs, gamma_inverse_sum_s_g1, d_proofs);
}
// ---------------------------------------------------------------------------------------------
// point-range sharded proving (BASELINE configs[4]): every rank holds a contiguous slice of each query,
// computes its share of the five multiexps (vsp_prove_partial), the shares are summed after one all_gather
// (sharded.py), and the proof is assembled from the sums (vsp_prove_assemble)
// ---------------------------------------------------------------------------------------------
namespace vsp {
template <class F>
__global__ void k_affine_to_xyzz(const uint64_t* __restrict__ in, unsigned n, uint32_t* __restrict__ out) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Affine<F> a = Affine<F>::load(in + uint64_t(i) * (Affine<F>::BYTES / 8));
    if (!a.is_inf()) {
        a.x = F::to_mont(a.x);
        a.y = F::to_mont(a.y);
    }
    XYZZ<F>::from_affine(a).store(out + uint64_t(i) * (XYZZ<F>::BYTES / 4));
}
}  // namespace vsp

namespace vsp {
// the five multiexp shares of this key (slice) for ONE assignment already on the device, as raw Montgomery XYZZ:
//   d_o (144 u64) = evA | evB1 | evH | evL (G1, 24 limbs each) | evB2 (G2, 48 limbs).  Stream-ordered, no host wait.
constexpr int PARTIAL_LIMBS = 4 * 24 + 48;
static int prove_partial_impl(vsp_ctx* ctx, const vsp_pk* pk, const uint32_t* w, uint64_t* o) {
    cudaStream_t s = ctx->stream;
    const uint64_t m = pk->m, stride = pk->nv + 1;
    DevBuf abc, bsc;
    VSP_TRY(bsc.alloc(ctx, (pk->B_count ? pk->B_count : 1) * 32, s));
    VSP_TRY(abc.alloc(ctx, size_t(3) * m * 32, s));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_fork, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 2), ctx->ev_fork, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 3), ctx->ev_fork, 0));
    int rc = msm_run(ctx, side_stream(ctx, 2), pk->A, pk->A_n, w + 8 * pk->A_lo, o, true);
    if (rc == VSP_OK) rc = msm_run(ctx, side_stream(ctx, 2), pk->L, pk->L_n, w + 8 * (pk->ni + 1 + pk->L_lo), o + 72, true);
    cudaEventRecord(ctx->ev_join[2], side_stream(ctx, 2));
    cudaStream_t s3 = side_stream(ctx, 3);
    if (rc == VSP_OK && pk->B_count)
        VSP_LAUNCH_K(ctx, KC_QAP, k_gather_fr, dim3((unsigned)((pk->B_count + 127) / 128), 1), 128, 0, s3, w, stride,
                     pk->d_B_idx, pk->B_count, bsc.as<uint32_t>());
    if (rc == VSP_OK) rc = msm_run(ctx, s3, pk->Bg2, pk->B_count, bsc.as<uint32_t>(), o + 96, true);
    if (rc == VSP_OK) rc = msm_run(ctx, s3, pk->Bg1, pk->B_count, bsc.as<uint32_t>(), o + 24, true);
    cudaEventRecord(ctx->ev_join[3], s3);
    // the witness map is replicated: every rank needs its slice of coefficients_for_H
    if (rc == VSP_OK) rc = witness_map_dev(ctx, s, pk, w, stride, 1, abc.as<uint32_t>());
    if (rc == VSP_OK) rc = msm_run(ctx, s, pk->H, pk->H_n, abc.as<uint32_t>() + 8 * pk->H_lo, o + 48, true);
    // join on every path, so that scratch owned by this call is not released under a side stream still reading it
    cudaStreamWaitEvent(s, ctx->ev_join[2], 0);
    cudaStreamWaitEvent(s, ctx->ev_join[3], 0);
    if (rc != VSP_OK) return rc;
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// sums[j] = sum over `world` ranks of gathered[rank][j-th point]   (thread j: 0..3 G1, 4 G2)
__global__ void k_sum_partials(const uint64_t* __restrict__ gathered, unsigned world, uint32_t* __restrict__ ev) {
    const unsigned j = blockIdx.x;
    if (threadIdx.x != 0) return;
    if (j < 4) {
        G1XYZZ acc = G1XYZZ::inf();
        for (unsigned g = 0; g < world; ++g) ec_add<FqOps>(acc, G1XYZZ::load(gathered + uint64_t(g) * PARTIAL_LIMBS + 24 * j));
        acc.store(ev + 48 * j);
    } else {
        G2XYZZ acc = G2XYZZ::inf();
        for (unsigned g = 0; g < world; ++g) ec_add<Fq2Ops>(acc, G2XYZZ::load(gathered + uint64_t(g) * PARTIAL_LIMBS + 96));
        acc.store(ev + 48 * 4);
    }
}
// XYZZ shares -> canonical affine (the host-facing form of vsp_prove_partial)
__global__ void k_partials_to_affine(const uint64_t* __restrict__ xyzz, uint64_t* __restrict__ out) {
    const unsigned j = blockIdx.x;
    if (threadIdx.x != 0) return;
    if (j < 4)
        write_canonical<FqOps>(out + 12 * j, G1XYZZ::load(xyzz + 24 * j), true);
    else
        write_canonical<Fq2Ops>(out + 48, G2XYZZ::load(xyzz + 96), true);
}
struct AsmScalars {
    uint64_t fsc[16];  // r | s | -rs | r_saver
    uint64_t vsc[8];   // s | r
};
__global__ void k_store_scalars(AsmScalars v, uint64_t* fsc, uint64_t* vsc, FixedJob* jobs) {
    if (threadIdx.x < 16) fsc[threadIdx.x] = v.fsc[threadIdx.x];
    if (threadIdx.x < 8) vsc[threadIdx.x] = v.vsc[threadIdx.x];
    if (threadIdx.x == 0) {
        jobs[0] = FixedJob{0, 0};
        jobs[1] = FixedJob{0, 1};
        jobs[2] = FixedJob{0, 2};
        jobs[3] = FixedJob{1, 3};
        jobs[4] = FixedJob{0, 1};
    }
}

// ev: evA | evB1 | evH | evL | evB2 summed over all ranks (device, XYZZ Montgomery); d_proof: 48 limbs on the device.
// Stream-ordered: the host scalars travel as kernel arguments.
static int assemble_impl(vsp_ctx* ctx, vsp_pk* pk, const uint32_t* ev, const uint64_t r[4], const uint64_t sv[4],
                         const uint64_t* r_saver, const uint64_t* gamma_inverse_sum_s_g1, uint64_t* d_proof, DevBuf* keep) {
    using namespace hostfr;
    const bool enc = r_saver && gamma_inverse_sum_s_g1;
    if (!fr_reduced(r) || !fr_reduced(sv) || (enc && !fr_reduced(r_saver)))
        return set_error(ctx, VSP_E_INVALID, "proving randomness not reduced mod r");
    if (enc) VSP_TRY(ensure_gamma_table(ctx, pk, gamma_inverse_sum_s_g1));
    cudaStream_t s = ctx->stream;
    constexpr size_t X1 = G1XYZZ::BYTES, X2 = G2XYZZ::BYTES;
    Fr4 fr, fs;
    memcpy(fr.v, r, 32);
    memcpy(fs.v, sv, 32);
    const Fr4 frs = sub(from_u64(0), mul(fr, fs));
    AsmScalars sc;
    memset(&sc, 0, sizeof(sc));
    memcpy(sc.fsc, r, 32);
    memcpy(sc.fsc + 4, sv, 32);
    memcpy(sc.fsc + 8, frs.v, 32);
    if (enc) memcpy(sc.fsc + 12, r_saver, 32);
    memcpy(sc.vsc, sv, 32);
    memcpy(sc.vsc + 4, r, 32);
    // one scratch block: fsc 128 | vsc 64 | jobs 64 | fx1 4 X1 | fx2 X2 | gab 192 | sg 2 X1
    DevBuf& buf = *keep;
    VSP_TRY(buf.alloc(ctx, 256 + 4 * X1 + X2 + 192 + 2 * X1, s));
    char* bp = buf.as<char>();
    uint64_t* d_fsc = reinterpret_cast<uint64_t*>(bp);
    uint64_t* d_vsc = reinterpret_cast<uint64_t*>(bp + 128);
    FixedJob* jobs1 = reinterpret_cast<FixedJob*>(bp + 192);
    uint32_t* fx1 = reinterpret_cast<uint32_t*>(bp + 256);
    uint32_t* fx2 = reinterpret_cast<uint32_t*>(bp + 256 + 4 * X1);
    uint64_t* gab = reinterpret_cast<uint64_t*>(bp + 256 + 4 * X1 + X2);
    uint32_t* sg = reinterpret_cast<uint32_t*>(bp + 256 + 4 * X1 + X2 + 192);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_store_scalars, 1, 32, 0, s, sc, d_fsc, d_vsc, jobs1);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<Fq2Ops>, 1, 32 * FIXED_WARPS, FIXED_WARPS * 32 * X2, s, pk->fx2tab.d_pts,
                 pk->fx2tab.n, jobs1 + 4, 1u, 1u, reinterpret_cast<uint32_t*>(d_fsc), 4u, (const uint32_t*)nullptr, uint64_t(0),
                 fx2);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<FqOps>, (4 + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS,
                 FIXED_WARPS * 32 * X1, s, pk->fx1tab.d_pts, pk->fx1tab.n, jobs1, 4u, 4u, reinterpret_cast<uint32_t*>(d_fsc), 4u,
                 (const uint32_t*)nullptr, uint64_t(0), fx1);
    AsmPtrs ap;
    ap.evA = ev;
    ap.evB1 = ap.evA + X1 / 4;
    ap.evH = ap.evB1 + X1 / 4;
    ap.evL = ap.evH + X1 / 4;
    ap.evB2 = ap.evL + X1 / 4;
    ap.fx1 = fx1;
    ap.fx2 = fx2;
    ap.fixed1 = pk->fixed1.d_pts;
    ap.fixed2 = pk->fixed2.d_pts;
    ap.gab = gab;
    ap.sg = sg;
    ap.proof = d_proof;
    ap.B = 1;
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase1, dim3(1, 3), 32, 0, s, ap);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, 2, 32, 0, s, gab, reinterpret_cast<uint32_t*>(d_vsc), 2u, 0, sg);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase3, 1, 32, 0, s, ap);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}
}  // namespace vsp

// vsp_prove_partial_dev in two calls with room for an exchange between them (sharded.ShardedProver.prove_dev):
//   begin   the A / L / B multiexps start on the side streams; on the main stream the constraint evaluation and the coset
//           extension of the vectors in `vector_mask` (bit 0 aA, 1 aB, 2 aC), left in d_abc (3 m Fr, caller-owned)
//   ...     the ranks exchange the extended vectors (three broadcasts of m Fr over NCCL) while the multiexps run
//   finish  H on the coset, its coefficients, the H multiexp; joins the side streams; d_partials complete
extern "C" int vsp_prove_partial_begin_dev(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* d_assignment, unsigned vector_mask,
                                           uint64_t* d_abc, uint64_t* d_partials) {
    if (!ctx || !pk || !d_assignment || !d_abc || !d_partials) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStream_t s = ctx->stream;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(d_assignment);
    const uint64_t stride = pk->nv + 1;
    uint64_t* o = d_partials;
    // the gathered B scalars must outlive this call: they live behind the three vectors' scratch in a context-owned buffer
    if (ctx->partial_bsc_cap < (pk->B_count ? pk->B_count : 1) * 32) {
        if (ctx->partial_bsc) cudaFree(ctx->partial_bsc);
        ctx->partial_bsc = nullptr;
        ctx->partial_bsc_cap = 0;
        VSP_CUDA_CHECK(ctx, cudaMalloc(&ctx->partial_bsc, (pk->B_count ? pk->B_count : 1) * 32));
        ctx->partial_bsc_cap = (pk->B_count ? pk->B_count : 1) * 32;
    }
    uint32_t* bsc = reinterpret_cast<uint32_t*>(ctx->partial_bsc);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_fork, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 2), ctx->ev_fork, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 3), ctx->ev_fork, 0));
    int rc = msm_run(ctx, side_stream(ctx, 2), pk->A, pk->A_n, w + 8 * pk->A_lo, o, true);
    if (rc == VSP_OK) rc = msm_run(ctx, side_stream(ctx, 2), pk->L, pk->L_n, w + 8 * (pk->ni + 1 + pk->L_lo), o + 72, true);
    cudaEventRecord(ctx->ev_join[2], side_stream(ctx, 2));
    cudaStream_t s3 = side_stream(ctx, 3);
    if (rc == VSP_OK && pk->B_count)
        VSP_LAUNCH_K(ctx, KC_QAP, k_gather_fr, dim3((unsigned)((pk->B_count + 127) / 128), 1), 128, 0, s3, w, stride, pk->d_B_idx,
                     pk->B_count, bsc);
    if (rc == VSP_OK) rc = msm_run(ctx, s3, pk->Bg2, pk->B_count, bsc, o + 96, true);
    if (rc == VSP_OK) rc = msm_run(ctx, s3, pk->Bg1, pk->B_count, bsc, o + 24, true);
    cudaEventRecord(ctx->ev_join[3], s3);
    if (rc == VSP_OK)
        rc = witness_map_evaluate(ctx, s, pk, w, stride, 1, reinterpret_cast<uint32_t*>(d_abc), vector_mask & 7u);
    if (rc != VSP_OK) {  // nothing of this call may still be running when the caller releases its buffers
        cudaStreamWaitEvent(s, ctx->ev_join[2], 0);
        cudaStreamWaitEvent(s, ctx->ev_join[3], 0);
    }
    return rc;
}
extern "C" int vsp_prove_partial_finish_dev(vsp_ctx* ctx, const vsp_pk* pk, uint64_t* d_abc, uint64_t* d_partials) {
    if (!ctx || !pk || !d_abc || !d_partials) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStream_t s = ctx->stream;
    uint32_t* abc = reinterpret_cast<uint32_t*>(d_abc);
    int rc = witness_map_finish(ctx, s, pk, 1, abc);
    if (rc == VSP_OK) rc = msm_run(ctx, s, pk->H, pk->H_n, abc + 8 * pk->H_lo, d_partials + 48, true);
    cudaStreamWaitEvent(s, ctx->ev_join[2], 0);
    cudaStreamWaitEvent(s, ctx->ev_join[3], 0);
    if (rc != VSP_OK) return rc;
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// out_partials: evA | evB1 | evH | evL (G1 affine) | evB2 (G2 affine) = 4 * 12 + 24 = 72 limbs, canonical
extern "C" int vsp_prove_partial(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux,
                                 uint64_t* out_partials) {
    if (!ctx || !pk || !out_partials || (!primary && pk->ni) || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStream_t s = ctx->stream;
    DevBuf d_w, raw, out;
    VSP_TRY(upload_witness(ctx, pk, 1, primary, aux, d_w));
    VSP_TRY(raw.alloc(ctx, PARTIAL_LIMBS * 8, s));
    VSP_TRY(out.alloc(ctx, 72 * 8, s));
    VSP_TRY(prove_partial_impl(ctx, pk, d_w.as<uint32_t>(), raw.as<uint64_t>()));
    VSP_LAUNCH_K(ctx, KC_FIXED, k_partials_to_affine, 5, 32, 0, s, raw.as<uint64_t>(), out.as<uint64_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_partials, out.p, 72 * 8, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

// device-resident variant: d_assignment = (1 || primary || auxiliary) canonical Fr; d_partials = 144 limbs of raw
// Montgomery XYZZ shares (evA | evB1 | evH | evL | evB2) for vsp_prove_assemble_dev.  Stream-ordered, no host wait.
extern "C" int vsp_prove_partial_dev(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* d_assignment, uint64_t* d_partials) {
    if (!ctx || !pk || !d_assignment || !d_partials) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return prove_partial_impl(ctx, pk, reinterpret_cast<const uint32_t*>(d_assignment), d_partials);
}

// sums: the five multiexp results summed over all ranks, same layout as out_partials
extern "C" int vsp_prove_assemble(vsp_ctx* ctx, const vsp_pk* cpk, const uint64_t* sums, const uint64_t r[4],
                                  const uint64_t sv[4], const uint64_t* r_saver, const uint64_t* gamma_inverse_sum_s_g1,
                                  uint64_t* out_proof) {
    if (!ctx || !cpk || !sums || !r || !sv || !out_proof) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    vsp_pk* pk = const_cast<vsp_pk*>(cpk);
    cudaStream_t s = ctx->stream;
    constexpr size_t X1 = G1XYZZ::BYTES, X2 = G2XYZZ::BYTES;
    DevBuf d_sums, ev, proof, scratch;
    VSP_TRY(d_sums.alloc(ctx, 72 * 8, s));
    VSP_TRY(ev.alloc(ctx, 4 * X1 + X2, s));
    VSP_TRY(proof.alloc(ctx, 384, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_sums.p, sums, 72 * 8, cudaMemcpyHostToDevice, s));
    // sums (evA | evB1 | evH | evL | evB2, affine canonical) -> the XYZZ Montgomery layout of the assembly kernels
    VSP_LAUNCH_K(ctx, KC_FIXED, k_affine_to_xyzz<FqOps>, 1, 32, 0, s, d_sums.as<uint64_t>(), 4u, ev.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_affine_to_xyzz<Fq2Ops>, 1, 32, 0, s, d_sums.as<uint64_t>() + 48, 1u,
                 ev.as<uint32_t>() + 4 * X1 / 4);
    VSP_TRY(assemble_impl(ctx, pk, ev.as<uint32_t>(), r, sv, r_saver, gamma_inverse_sum_s_g1, proof.as<uint64_t>(), &scratch));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_proof, proof.p, 384, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

// device-resident variant: d_gathered = world x 144 limbs (every rank's vsp_prove_partial_dev output, e.g. straight out
// of an all_gather); the shares are summed on the device and the proof (48 limbs) is left at d_proof.  Stream-ordered.
extern "C" int vsp_prove_assemble_dev(vsp_ctx* ctx, const vsp_pk* cpk, uint64_t world, const uint64_t* d_gathered,
                                      const uint64_t r[4], const uint64_t sv[4], const uint64_t* r_saver,
                                      const uint64_t* gamma_inverse_sum_s_g1, uint64_t* d_proof) {
    if (!ctx || !cpk || !d_gathered || !r || !sv || !d_proof || world == 0 || world > 4096) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    vsp_pk* pk = const_cast<vsp_pk*>(cpk);
    cudaStream_t s = ctx->stream;
    DevBuf ev, scratch;
    VSP_TRY(ev.alloc(ctx, 4 * G1XYZZ::BYTES + G2XYZZ::BYTES, s));
    VSP_LAUNCH_K(ctx, KC_FIXED, k_sum_partials, 5, 32, 0, s, d_gathered, (unsigned)world, ev.as<uint32_t>());
    return assemble_impl(ctx, pk, ev.as<uint32_t>(), r, sv, r_saver, gamma_inverse_sum_s_g1, d_proof, &scratch);
}

// ---------------------------------------------------------------------------------------------
// SAVER
// ---------------------------------------------------------------------------------------------
namespace vsp {

// jobs: out group g gets sum of scalars[i] * points[i]; all host inputs canonical
static int run_mul_jobs_g1(vsp_ctx* ctx, const std::vector<const uint64_t*>& pts, const std::vector<const uint64_t*>& scs,
                           const std::vector<uint32_t>& grp, unsigned n_groups, const uint64_t* addend_host,
                           uint64_t* out_host) {
    cudaStream_t s = ctx->stream;
    const unsigned n = (unsigned)pts.size();
    std::vector<uint64_t> hp(12 * (size_t)n), hs(4 * (size_t)n);
    for (unsigned i = 0; i < n; ++i) {
        memcpy(&hp[12 * i], pts[i], 96);
        memcpy(&hs[4 * i], scs[i], 32);
        if (!fr_reduced(scs[i])) return set_error(ctx, VSP_E_INVALID, "scalar not reduced mod r");
    }
    DevBuf dp, ds, dg, dx, da, dout;
    VSP_TRY(dp.alloc(ctx, hp.size() * 8, s));
    VSP_TRY(ds.alloc(ctx, hs.size() * 8, s));
    VSP_TRY(dg.alloc(ctx, (size_t)n * 4, s));
    VSP_TRY(dx.alloc(ctx, (size_t)n * G1XYZZ::BYTES, s));
    VSP_TRY(dout.alloc(ctx, (size_t)n_groups * 96, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(dp.p, hp.data(), hp.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(ds.p, hs.data(), hs.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(dg.p, grp.data(), (size_t)n * 4, cudaMemcpyHostToDevice, s));
    if (addend_host) {
        VSP_TRY(da.alloc(ctx, (size_t)n_groups * 96, s));
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(da.p, addend_host, (size_t)n_groups * 96, cudaMemcpyHostToDevice, s));
    }
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, n, 32, 0, s, dp.as<uint64_t>(), ds.as<uint32_t>(), n, 0,
               dx.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_group_sum<FqOps>, (n_groups + 31) / 32, 32, 0, s, dx.as<uint32_t>(), dg.as<uint32_t>(), n,
               addend_host ? da.as<uint64_t>() : nullptr, n_groups, dout.as<uint64_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_host, dout.p, (size_t)n_groups * 96, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

}  // namespace vsp

extern "C" int vsp_g1_mul_batch(vsp_ctx* ctx, uint64_t n, const uint64_t* points, const uint64_t* scalars, uint64_t* out) {
    if (!ctx) return VSP_E_INVALID;
    if (n == 0) return VSP_OK;
    if (!points || !scalars || !out || n > (1u << 20)) return set_error(ctx, VSP_E_INVALID, "vsp_g1_mul_batch: bad argument");
    cudaSetDevice(ctx->device);
    std::vector<const uint64_t*> pts(n), scs(n);
    std::vector<uint32_t> grp(n);
    for (uint64_t i = 0; i < n; ++i) {
        pts[i] = points + 12 * i;
        scs[i] = scalars + 4 * i;
        grp[i] = (uint32_t)i;
    }
    return run_mul_jobs_g1(ctx, pts, scs, grp, (unsigned)n, nullptr, out);
}

extern "C" int vsp_saver_ciphertext(vsp_ctx* ctx, const vsp_saver_pubkey_t* pub, const uint64_t* gamma_abc_rest,
                                    const uint64_t* msg, const uint64_t r_saver[4], uint64_t* out_ct) {
    if (!ctx || !pub || !gamma_abc_rest || (!msg && pub->msg_size) || !r_saver || !out_ct) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    const unsigned n = (unsigned)pub->msg_size;
    // ct[0] = r delta; ct[1+i] = r delta_s[i] + m_i gamma_abc[i]; ct[n+1] = r delta_sum_s + sum m_i t[i]
    // (elgamal_verifiable.hpp:341-352)
    std::vector<const uint64_t*> pts, scs;
    std::vector<uint32_t> grp;
    pts.push_back(pub->delta_g1); scs.push_back(r_saver); grp.push_back(0);
    for (unsigned i = 0; i < n; ++i) {
        pts.push_back(pub->delta_s_g1 + 12 * i); scs.push_back(r_saver); grp.push_back(1 + i);
        pts.push_back(gamma_abc_rest + 12 * i); scs.push_back(msg + 4 * i); grp.push_back(1 + i);
        pts.push_back(pub->t_g1 + 12 * i); scs.push_back(msg + 4 * i); grp.push_back(n + 1);
    }
    pts.push_back(pub->delta_sum_s_g1); scs.push_back(r_saver); grp.push_back(n + 1);
    return run_mul_jobs_g1(ctx, pts, scs, grp, n + 2, nullptr, out_ct);
}

extern "C" int vsp_saver_encrypt(vsp_ctx* ctx, const vsp_pk* pk, const vsp_saver_pubkey_t* pub,
                                 const uint64_t* gamma_abc_rest, const uint64_t* msg, const uint64_t* primary,
                                 const uint64_t* aux, const uint64_t r_saver[4], const uint64_t r[4], const uint64_t s[4],
                                 uint64_t* out_ct, uint64_t* out_proof) {
    if (!ctx || !pk || !pub || !msg || !primary || !out_ct || !out_proof) return VSP_E_INVALID;
    // "Plain text should be a prefix of primary input" (elgamal_verifiable.hpp:336-339)
    // and "Primary input should be longer than plain text." (:325-326)
    if (pub->msg_size >= pk->ni || memcmp(msg, primary, pub->msg_size * 32) != 0)
        return set_error(ctx, VSP_E_INVALID, "plain text should be a proper prefix of the primary input");
    VSP_TRY(vsp_saver_ciphertext(ctx, pub, gamma_abc_rest, msg, r_saver, out_ct));
    return vsp_prove(ctx, pk, primary, aux, r, s, r_saver, pub->gamma_inverse_sum_s_g1, out_proof);
}

extern "C" int vsp_saver_rerandomize(vsp_ctx* ctx, const vsp_saver_pubkey_t* pub, const uint64_t* delta_g2,
                                     const uint64_t* rnd, const uint64_t* ct_in, const uint64_t* proof_in,
                                     uint64_t* ct_out, uint64_t* proof_out) {
    using namespace hostfr;
    if (!ctx || !pub || !delta_g2 || !rnd || !ct_in || !proof_in || !ct_out || !proof_out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    const unsigned n = (unsigned)pub->msg_size;
    const uint64_t *r = rnd, *z1 = rnd + 4, *z2 = rnd + 8;
    if (!fr_reduced(r) || !fr_reduced(z1) || !fr_reduced(z2)) return set_error(ctx, VSP_E_INVALID, "rnd not reduced");
    Fr4 fz1, fz2;
    memcpy(fz1.v, z1, 32);
    memcpy(fz2.v, z2, 32);
    if (is_zero(fz1)) return set_error(ctx, VSP_E_INVALID, "z1 must be invertible");
    const Fr4 z1inv = inv(fz1), z1z2 = mul(fz1, fz2);
    // G1 part: ct'[i] = ct[i] + r base_i ; A' = z1 A ; C' = C + (z1 z2) A + r gamma_inv     (:653-663)
    std::vector<const uint64_t*> pts, scs;
    std::vector<uint32_t> grp;
    pts.push_back(pub->delta_g1); scs.push_back(r); grp.push_back(0);
    for (unsigned i = 0; i < n; ++i) { pts.push_back(pub->delta_s_g1 + 12 * i); scs.push_back(r); grp.push_back(1 + i); }
    pts.push_back(pub->delta_sum_s_g1); scs.push_back(r); grp.push_back(n + 1);
    pts.push_back(proof_in); scs.push_back(z1); grp.push_back(n + 2);                               // A'
    pts.push_back(proof_in); scs.push_back(z1z2.v); grp.push_back(n + 3);                           // z2 A'
    pts.push_back(pub->gamma_inverse_sum_s_g1); scs.push_back(r); grp.push_back(n + 3);
    std::vector<uint64_t> addend(12 * (size_t)(n + 4), 0), out(12 * (size_t)(n + 4));
    memcpy(addend.data(), ct_in, (size_t)(n + 2) * 96);
    memcpy(&addend[12 * (size_t)(n + 3)], proof_in + 36, 96);  // C
    VSP_TRY(run_mul_jobs_g1(ctx, pts, scs, grp, n + 4, addend.data(), out.data()));
    memcpy(ct_out, out.data(), (size_t)(n + 2) * 96);
    memcpy(proof_out, &out[12 * (size_t)(n + 2)], 96);
    memcpy(proof_out + 36, &out[12 * (size_t)(n + 3)], 96);
    // G2 part: B' = z1^-1 B + z2 delta_g2
    cudaStream_t s = ctx->stream;
    uint64_t hp[48], hs[8];
    uint32_t hg[2] = {0, 0};
    memcpy(hp, proof_in + 12, 192);
    memcpy(hp + 24, delta_g2, 192);
    memcpy(hs, z1inv.v, 32);
    memcpy(hs + 4, z2, 32);
    DevBuf dp, ds, dg, dx, dout;
    VSP_TRY(dp.alloc(ctx, sizeof(hp), s));
    VSP_TRY(ds.alloc(ctx, sizeof(hs), s));
    VSP_TRY(dg.alloc(ctx, sizeof(hg), s));
    VSP_TRY(dx.alloc(ctx, 2 * G2XYZZ::BYTES, s));
    VSP_TRY(dout.alloc(ctx, 192, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(dp.p, hp, sizeof(hp), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(ds.p, hs, sizeof(hs), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(dg.p, hg, sizeof(hg), cudaMemcpyHostToDevice, s));
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<Fq2Ops>, 2, 32, 0, s, dp.as<uint64_t>(), ds.as<uint32_t>(), 2u, 0, dx.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_group_sum<Fq2Ops>, 1, 32, 0, s, dx.as<uint32_t>(), dg.as<uint32_t>(), 2u, nullptr, 1u,
               dout.as<uint64_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(proof_out + 12, dout.p, 192, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// SAVER, batched: resident public key + ciphertexts and proofs of k ballots in one call
// ---------------------------------------------------------------------------------------------
extern "C" void vsp_saver_pub_destroy(vsp_saver_pub* sp) {
    if (!sp) return;
    cudaSetDevice(sp->ctx->device);
    cudaDeviceSynchronize();
    bases_free(&sp->tab);
    delete sp;
}

extern "C" int vsp_saver_pub_load(vsp_ctx* ctx, const vsp_saver_pubkey_t* pub, const uint64_t* gamma_abc_rest,
                                  vsp_saver_pub** out) {
    if (!ctx || !pub || !out || !pub->delta_g1 || !pub->delta_sum_s_g1 || !pub->gamma_inverse_sum_s_g1 ||
        (pub->msg_size && (!pub->delta_s_g1 || !pub->t_g1 || !gamma_abc_rest)))
        return VSP_E_INVALID;
    *out = nullptr;
    if (pub->msg_size > 4096) return set_error(ctx, VSP_E_INVALID, "msg_size too large");
    cudaSetDevice(ctx->device);
    vsp_saver_pub* sp = new (std::nothrow) vsp_saver_pub();
    if (!sp) return VSP_E_NOMEM;
    sp->ctx = ctx;
    sp->n = (unsigned)pub->msg_size;
    const unsigned n = sp->n;
    std::vector<uint64_t> pts(12 * size_t(3 * n + 2));
    memcpy(&pts[0], pub->delta_g1, 96);
    if (n) {
        memcpy(&pts[12], pub->delta_s_g1, size_t(n) * 96);
        memcpy(&pts[12 * size_t(1 + n)], gamma_abc_rest, size_t(n) * 96);
        memcpy(&pts[12 * size_t(1 + 2 * n)], pub->t_g1, size_t(n) * 96);
    }
    memcpy(&pts[12 * size_t(1 + 3 * n)], pub->delta_sum_s_g1, 96);
    memcpy(sp->gamma_inv, pub->gamma_inverse_sum_s_g1, 96);
    sp->tab.force_c = 8;
    int rc = bases_upload(ctx, false, 3 * n + 2, pts.data(), &sp->tab, true);
    if (rc == VSP_OK && !sp->tab.table) rc = set_error(ctx, VSP_E_NOMEM, "fixed-base table not built");
    if (rc) {
        bases_free(&sp->tab);
        delete sp;
        return rc;
    }
    *out = sp;
    return VSP_OK;
}

namespace vsp {
// ciphertexts of B ballots: messages are the first msg_size primary inputs of each assignment (the reference asserts
// the plain text is a prefix of the primary input, elgamal_verifiable.hpp:336-339); r_saver = rs[b][2]
static int ciphertext_chunk(vsp_ctx* ctx, const vsp_saver_pub* sp, unsigned B, const uint32_t* d_w, uint64_t w_stride,
                            const uint64_t* rs, uint64_t* d_cts) {
    cudaStream_t s = side_stream(ctx, 0);
    const unsigned n = sp->n, J = 3 * n + 2;
    std::vector<FixedJob> spec(J);
    spec[0] = FixedJob{0, 0};
    for (unsigned i = 0; i < n; ++i) {
        spec[1 + i] = FixedJob{int(1 + i), 0};                      // r delta_s[i]
        spec[1 + n + i] = FixedJob{int(1 + n + i), -int(1 + i)};    // m_i gamma_abc[i]   (w[1 + i] = primary[i])
        spec[1 + 2 * n + i] = FixedJob{int(1 + 2 * n + i), -int(1 + i)};  // m_i t[i]
    }
    spec[3 * n + 1] = FixedJob{int(3 * n + 1), 0};
    std::vector<uint64_t> sc(size_t(B) * 4);
    for (unsigned b = 0; b < B; ++b) {
        if (!fr_reduced(rs + 12 * b + 8)) return set_error(ctx, VSP_E_INVALID, "r_saver not reduced mod r");
        memcpy(&sc[size_t(b) * 4], rs + 12 * b + 8, 32);
    }
    DevBuf d_spec, d_sc, mul;
    VSP_TRY(d_spec.alloc(ctx, J * sizeof(FixedJob), s));
    VSP_TRY(d_sc.alloc(ctx, sc.size() * 8, s));
    VSP_TRY(mul.alloc(ctx, size_t(B) * J * G1XYZZ::BYTES, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_spec.p, spec.data(), J * sizeof(FixedJob), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_sc.p, sc.data(), sc.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<FqOps>, (B * J + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS,
                 FIXED_WARPS * 32 * G1XYZZ::BYTES, s, sp->tab.d_pts, sp->tab.n, d_spec.as<FixedJob>(), J, B * J,
                 d_sc.as<uint32_t>(), 1u, d_w, w_stride, mul.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_ct_sum, (B * (n + 2) + 63) / 64, 64, 0, s, mul.as<uint32_t>(), n, B, d_cts);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

static int saver_encrypt_batch_impl(vsp_ctx* ctx, const vsp_pk* cpk, const vsp_saver_pub* sp, uint64_t k,
                                    const uint64_t* primary, const uint64_t* aux, const uint64_t* d_assignments,
                                    const uint64_t* rs, uint64_t* d_cts, uint64_t* d_proofs) {
    vsp_pk* pk = const_cast<vsp_pk*>(cpk);
    if (sp->n >= pk->ni)  // "Primary input should be longer than plain text." (elgamal_verifiable.hpp:325-326)
        return set_error(ctx, VSP_E_INVALID, "plain text should be a proper prefix of the primary input");
    const uint64_t stride = pk->nv + 1;
    const unsigned chunk = batch_chunk(ctx, pk);
    const uint64_t ct_limbs = uint64_t(sp->n + 2) * 12;
    WitnessPrefetch pre(ctx, pk, primary, aux, k, chunk);
    if (!d_assignments) VSP_TRY(pre.init());
    for (uint64_t i = 0, c = 0; i < k; i += chunk, ++c) {
        const unsigned B = (unsigned)((k - i < chunk) ? (k - i) : chunk);
        const uint32_t* w;
        if (d_assignments)
            w = reinterpret_cast<const uint32_t*>(d_assignments + 4 * stride * i);
        else
            VSP_TRY(pre.acquire(c, &w));
        // the ciphertext runs on side stream 0 beside the proof
        VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[0], ctx->stream));
        VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(side_stream(ctx, 0), ctx->ev_join[0], 0));
        VSP_TRY(ciphertext_chunk(ctx, sp, B, w, stride, rs + 12 * i, d_cts + ct_limbs * i));
        VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[0], side_stream(ctx, 0)));
        VSP_TRY(prove_chunk(ctx, pk, B, w, stride, rs + 12 * i, sp->gamma_inv, d_proofs + 48 * i));
        VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join[0], 0));
        if (!d_assignments) VSP_TRY(pre.release(c));
    }
    return join_tails(ctx);
}
}  // namespace vsp

extern "C" int vsp_saver_encrypt_batch(vsp_ctx* ctx, const vsp_pk* pk, const vsp_saver_pub* sp, uint64_t k,
                                       const uint64_t* primary, const uint64_t* aux, const uint64_t* rs, uint64_t* out_cts,
                                       uint64_t* out_proofs) {
    if (!ctx || !pk || !sp || !rs || !out_cts || !out_proofs || !primary || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    if (k == 0) return VSP_OK;
    const size_t ct_bytes = size_t(sp->n + 2) * 96;
    DevBuf d_cts, d_proofs;
    VSP_TRY(d_cts.alloc(ctx, k * ct_bytes, ctx->stream));
    VSP_TRY(d_proofs.alloc(ctx, k * 384, ctx->stream));
    VSP_TRY(saver_encrypt_batch_impl(ctx, pk, sp, k, primary, aux, nullptr, rs, d_cts.as<uint64_t>(), d_proofs.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_cts, d_cts.p, k * ct_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_proofs, d_proofs.p, k * 384, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

extern "C" int vsp_saver_encrypt_batch_dev(vsp_ctx* ctx, const vsp_pk* pk, const vsp_saver_pub* sp, uint64_t k,
                                           const uint64_t* d_assignments, const uint64_t* rs, uint64_t* d_cts,
                                           uint64_t* d_proofs) {
    if (!ctx || !pk || !sp || !rs || !d_cts || !d_proofs || !d_assignments) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return saver_encrypt_batch_impl(ctx, pk, sp, k, nullptr, nullptr, d_assignments, rs, d_cts, d_proofs);
}

// ---------------------------------------------------------------------------------------------
// rerandomize_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:632-667) for k ballots at once
//   ct'[0] = ct[0] + r delta;  ct'[i] = ct[i] + r delta_s[i-1];  ct'[n+1] = ct[n+1] + r delta_sum_s
//   A' = z1 A;   B' = z1^-1 B + z2 delta_g2;   C' = C + z2 A' + r gamma_inverse_sum_s_g1
// The n + 3 multiples of key material (r delta, r delta_s, r delta_sum_s from the resident SAVER table; r gamma_inv and
// z2 delta_g2 from the proving key's tables) are fixed-base: one warp per job, 8 doublings + 5 additions deep.  z1 A and
// (z1 z2) A share one launch of the endomorphism-split variable-base kernel, z1^-1 B is the only 255-bit G2 walk.
// z1^-1 comes from the host (one batched inversion for the call).
// ---------------------------------------------------------------------------------------------
namespace vsp {
// pts[2 b] = pts[2 b + 1] = proofs[b].A (the two G1 variable-base jobs of ballot b);  g2[b] = proofs[b].B
__global__ void k_rerand_gather(const uint64_t* __restrict__ proofs, unsigned k, uint64_t* __restrict__ pts, uint64_t* __restrict__ g2) {
    const unsigned q = blockIdx.x * blockDim.x + threadIdx.x;  // (ballot, limb)
    if (q >= k * 24) return;
    const unsigned b = q / 24, l = q % 24;
    const uint64_t* pr = proofs + uint64_t(b) * 48;
    g2[uint64_t(b) * 24 + l] = pr[12 + l];
    if (l < 12) {
        pts[uint64_t(2 * b) * 12 + l] = pr[l];
        pts[uint64_t(2 * b + 1) * 12 + l] = pr[l];
    }
}
// thread = (ballot, slot): slots 0 .. n+1 the cipher text, n+2: A', n+3: C', n+4: B'
__global__ void __launch_bounds__(64) k_rerand_sum(const uint64_t* __restrict__ cts_in, const uint64_t* __restrict__ proofs_in,
                                                   const uint32_t* __restrict__ fx, const uint32_t* __restrict__ fg,
                                                   const uint32_t* __restrict__ fd2, const uint32_t* __restrict__ v1,
                                                   const uint32_t* __restrict__ v2, unsigned n, unsigned k,
                                                   uint64_t* __restrict__ cts_out, uint64_t* __restrict__ proofs_out) {
    const unsigned q = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned slots = n + 5;
    if (q >= k * slots) return;
    const unsigned b = q / slots, slot = q - b * slots;
    constexpr int L1 = G1XYZZ::BYTES / 4, L2 = G2XYZZ::BYTES / 4;
    auto mont = [](G1Affine a) {
        if (!a.is_inf()) {
            a.x = fp_to_mont(a.x);
            a.y = fp_to_mont(a.y);
        }
        return a;
    };
    if (slot <= n + 1) {
        G1XYZZ acc = G1XYZZ::load(fx + (uint64_t(b) * (n + 2) + slot) * L1);
        ec_madd<FqOps>(acc, mont(G1Affine::load(cts_in + (uint64_t(b) * (n + 2) + slot) * 12)));
        write_canonical<FqOps>(cts_out + (uint64_t(b) * (n + 2) + slot) * 12, acc);
    } else if (slot == n + 2) {
        write_canonical<FqOps>(proofs_out + uint64_t(b) * 48, G1XYZZ::load(v1 + uint64_t(2 * b) * L1));
    } else if (slot == n + 3) {
        G1XYZZ acc = G1XYZZ::load(v1 + uint64_t(2 * b + 1) * L1);
        ec_add<FqOps>(acc, G1XYZZ::load(fg + uint64_t(b) * L1));
        ec_madd<FqOps>(acc, mont(G1Affine::load(proofs_in + uint64_t(b) * 48 + 36)));
        write_canonical<FqOps>(proofs_out + uint64_t(b) * 48 + 36, acc);
    } else {
        G2XYZZ acc = G2XYZZ::load(v2 + uint64_t(b) * L2);
        ec_add<Fq2Ops>(acc, G2XYZZ::load(fd2 + uint64_t(b) * L2));
        write_canonical<Fq2Ops>(proofs_out + uint64_t(b) * 48 + 12, acc);
    }
}

// d_cts_in / d_cts_out: k x (n + 2) G1;  d_proofs_in / d_proofs_out: k x 48 limbs;  rnd (host): k x (r | z1 | z2).
// Stream-ordered on ctx->stream except for the upload of the per-ballot scalars (pageable host memory: synchronous).
static int rerandomize_batch_impl(vsp_ctx* ctx, vsp_pk* pk, const vsp_saver_pub* sp, uint64_t k, const uint64_t* rnd,
                                  const uint64_t* d_cts_in, const uint64_t* d_proofs_in, uint64_t* d_cts_out,
                                  uint64_t* d_proofs_out) {
    using namespace hostfr;
    if (k > (1u << 20)) return set_error(ctx, VSP_E_INVALID, "rerandomize: batch too large");
    const unsigned n = sp->n, K = (unsigned)k;
    VSP_TRY(ensure_gamma_table(ctx, pk, sp->gamma_inv));
    cudaStream_t s = ctx->stream;
    constexpr size_t X1 = G1XYZZ::BYTES, X2 = G2XYZZ::BYTES;
    // per ballot: fixed-base slots r | z2, then the variable-base scalars z1 | z1 z2 (G1) and z1^-1 (G2)
    std::vector<uint64_t> fsc(size_t(K) * 8), v1sc(size_t(K) * 8), v2sc(size_t(K) * 4);
    std::vector<Fr4> z1s(K), pre(K);
    for (unsigned b = 0; b < K; ++b) {
        const uint64_t *r = rnd + 12 * b, *z1 = r + 4, *z2 = r + 8;
        if (!fr_reduced(r) || !fr_reduced(z1) || !fr_reduced(z2)) return set_error(ctx, VSP_E_INVALID, "rnd not reduced");
        memcpy(z1s[b].v, z1, 32);
        if (is_zero(z1s[b])) return set_error(ctx, VSP_E_INVALID, "z1 must be invertible");
        Fr4 fz2;
        memcpy(fz2.v, z2, 32);
        const Fr4 z1z2 = mul(z1s[b], fz2);
        memcpy(&fsc[size_t(b) * 8], r, 32);
        memcpy(&fsc[size_t(b) * 8 + 4], z2, 32);
        memcpy(&v1sc[size_t(b) * 8], z1, 32);
        memcpy(&v1sc[size_t(b) * 8 + 4], z1z2.v, 32);
        pre[b] = b ? mul(pre[b - 1], z1s[b]) : z1s[b];
    }
    Fr4 run = inv(pre[K - 1]);  // Montgomery's trick: one field inversion for the call
    for (unsigned b = K; b-- > 0;) {
        const Fr4 zi = b ? mul(run, pre[b - 1]) : run;
        run = mul(run, z1s[b]);
        memcpy(&v2sc[size_t(b) * 4], zi.v, 32);
    }
    std::vector<FixedJob> spec(n + 4);
    spec[0] = FixedJob{0, 0};
    for (unsigned i = 0; i < n; ++i) spec[1 + i] = FixedJob{int(1 + i), 0};
    spec[n + 1] = FixedJob{int(3 * n + 1), 0};
    spec[n + 2] = FixedJob{1, 0};  // fx1tab: r gamma_inv
    spec[n + 3] = FixedJob{0, 1};  // fx2tab: z2 delta_g2
    DevBuf d_spec, d_fsc, d_v1sc, d_v2sc, fx, fg, fd2, pts, g2, v1, v2;
    VSP_TRY(d_spec.alloc(ctx, spec.size() * sizeof(FixedJob), s));
    VSP_TRY(d_fsc.alloc(ctx, fsc.size() * 8, s));
    VSP_TRY(d_v1sc.alloc(ctx, v1sc.size() * 8, s));
    VSP_TRY(d_v2sc.alloc(ctx, v2sc.size() * 8, s));
    VSP_TRY(fx.alloc(ctx, size_t(K) * (n + 2) * X1, s));
    VSP_TRY(fg.alloc(ctx, size_t(K) * X1, s));
    VSP_TRY(fd2.alloc(ctx, size_t(K) * X2, s));
    VSP_TRY(pts.alloc(ctx, size_t(K) * 2 * 96, s));
    VSP_TRY(g2.alloc(ctx, size_t(K) * 192, s));
    VSP_TRY(v1.alloc(ctx, size_t(K) * 2 * X1, s));
    VSP_TRY(v2.alloc(ctx, size_t(K) * X2, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_spec.p, spec.data(), spec.size() * sizeof(FixedJob), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_fsc.p, fsc.data(), fsc.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_v1sc.p, v1sc.data(), v1sc.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_v2sc.p, v2sc.data(), v2sc.size() * 8, cudaMemcpyHostToDevice, s));
    const FixedJob* js = d_spec.as<FixedJob>();
    VSP_LAUNCH_K(ctx, KC_FIXED, k_rerand_gather, (K * 24 + 127) / 128, 128, 0, s, d_proofs_in, K, pts.as<uint64_t>(), g2.as<uint64_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<FqOps>, (K * (n + 2) + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS,
                 FIXED_WARPS * 32 * X1, s, sp->tab.d_pts, sp->tab.n, js, n + 2, K * (n + 2), d_fsc.as<uint32_t>(), 2u,
                 (const uint32_t*)nullptr, uint64_t(0), fx.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<FqOps>, (K + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS, FIXED_WARPS * 32 * X1, s,
                 pk->fx1tab.d_pts, pk->fx1tab.n, js + n + 2, 1u, K, d_fsc.as<uint32_t>(), 2u, (const uint32_t*)nullptr, uint64_t(0),
                 fg.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_mul<Fq2Ops>, (K + FIXED_WARPS - 1) / FIXED_WARPS, 32 * FIXED_WARPS, FIXED_WARPS * 32 * X2, s,
                 pk->fx2tab.d_pts, pk->fx2tab.n, js + n + 3, 1u, K, d_fsc.as<uint32_t>(), 2u, (const uint32_t*)nullptr, uint64_t(0),
                 fd2.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, 2 * K, 32, 0, s, pts.as<uint64_t>(), d_v1sc.as<uint32_t>(), 2 * K, 0,
                 v1.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<Fq2Ops>, K, 32, 0, s, g2.as<uint64_t>(), d_v2sc.as<uint32_t>(), K, 0,
                 v2.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_rerand_sum, (K * (n + 5) + 63) / 64, 64, 0, s, d_cts_in, d_proofs_in, fx.as<uint32_t>(),
                 fg.as<uint32_t>(), fd2.as<uint32_t>(), v1.as<uint32_t>(), v2.as<uint32_t>(), n, K, d_cts_out, d_proofs_out);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));  // the scalar uploads read host vectors of this frame
    return VSP_OK;
}
}  // namespace vsp

extern "C" int vsp_saver_rerandomize_batch(vsp_ctx* ctx, const vsp_pk* pk, const vsp_saver_pub* sp, uint64_t k,
                                           const uint64_t* rnd, const uint64_t* cts_in, const uint64_t* proofs_in,
                                           uint64_t* cts_out, uint64_t* proofs_out) {
    if (!ctx || !pk || !sp || !rnd || !cts_in || !proofs_in || !cts_out || !proofs_out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    if (k == 0) return VSP_OK;
    cudaStream_t s = ctx->stream;
    const size_t ctb = size_t(sp->n + 2) * 96 * k, prb = size_t(384) * k;
    DevBuf ci, pi, co, po;
    VSP_TRY(ci.alloc(ctx, ctb, s));
    VSP_TRY(pi.alloc(ctx, prb, s));
    VSP_TRY(co.alloc(ctx, ctb, s));
    VSP_TRY(po.alloc(ctx, prb, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(ci.p, cts_in, ctb, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(pi.p, proofs_in, prb, cudaMemcpyHostToDevice, s));
    VSP_TRY(rerandomize_batch_impl(ctx, const_cast<vsp_pk*>(pk), sp, k, rnd, ci.as<uint64_t>(), pi.as<uint64_t>(),
                                   co.as<uint64_t>(), po.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(cts_out, co.p, ctb, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(proofs_out, po.p, prb, cudaMemcpyDeviceToHost, s));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

// device-resident variant: cipher texts and proofs in and out stay on the device (e.g. straight from
// vsp_saver_encrypt_batch_dev); in and out may be the same buffers
extern "C" int vsp_saver_rerandomize_batch_dev(vsp_ctx* ctx, const vsp_pk* pk, const vsp_saver_pub* sp, uint64_t k,
                                               const uint64_t* rnd, const uint64_t* d_cts_in, const uint64_t* d_proofs_in,
                                               uint64_t* d_cts_out, uint64_t* d_proofs_out) {
    if (!ctx || !pk || !sp || !rnd || !d_cts_in || !d_proofs_in || !d_cts_out || !d_proofs_out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    if (k == 0) return VSP_OK;
    return rerandomize_batch_impl(ctx, const_cast<vsp_pk*>(pk), sp, k, rnd, d_cts_in, d_proofs_in, d_cts_out, d_proofs_out);
}
