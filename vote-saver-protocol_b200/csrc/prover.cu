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
// Stream plan for one proof (fork after the witness upload, join before assembly):
//   main   : A w / B w / C w  ->  aA: iNTT+coset, NTT  ->  pointwise H  ->  coset iNTT  ->  multiexp H  ->  assembly
//   side 0 : aB: iNTT+coset, NTT                         side 1 : aC: iNTT+coset, NTT
//   side 2 : multiexp A, multiexp L                      side 3 : multiexp B (G2), multiexp B (G1), fixed-base r/s multiples
#include "common.cuh"
#include "ec.cuh"
#include "host_fr.h"
#include "msm.cuh"

#include <cstring>
#include <new>
#include <vector>

namespace vsp {
int domain_plan(uint64_t min_size, int* kind, uint64_t* m);
int fill_power_table(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_out, uint64_t n, const hostfr::Fr4& g,
                     const hostfr::Fr4& k);
}  // namespace vsp

using namespace vsp;

struct vsp_pk {
    vsp_ctx* ctx = nullptr;
    uint64_t ni = 0, na = 0, nc = 0, nv = 0, m = 0;
    int kind = 0;
    // CSR (device): rows u64, cols u32, values Montgomery (A, C: v R;  B: v R^2 so that B w comes out in Montgomery form)
    uint64_t* d_row[3] = {nullptr, nullptr, nullptr};
    uint32_t* d_col[3] = {nullptr, nullptr, nullptr};
    uint32_t* d_val[3] = {nullptr, nullptr, nullptr};
    BasesDev A, Bg1, Bg2, H, L;
    BasesDev fixed1;  // alpha_g1, beta_g1, delta_g1 (Montgomery affine)
    BasesDev fixed2;  // beta_g2, delta_g2
    uint32_t* d_B_idx = nullptr;
    uint64_t B_count = 0;
    uint32_t* d_gpow = nullptr;     // g^i, i < m (Montgomery)
    uint32_t* d_ginvpow = nullptr;  // basic domain: Z(g)^-1 g^-i ; step domain: g^-i
    uint32_t* d_zinv = nullptr;     // step domain only: 1 / Z on the coset, per element (Montgomery)
    uint64_t h_delta_g1[12];        // canonical copies for the fixed-base scalar multiplications
    uint64_t h_delta_g2[24];
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

// one warp per domain row: aA / aB / aC[row] = <matrix row, (1 || primary || aux)>; rows n_c .. n_c + n_in of aA hold
// the input-consistency values, everything above is zero.  aB comes out in Montgomery form (values carry R^2).
__global__ void __launch_bounds__(256) k_r1cs_eval(CsrDev csr, const uint32_t* __restrict__ w, uint64_t nc, uint64_t ni,
                                                   uint64_t m, uint32_t* aA, uint32_t* aB, uint32_t* aC) {
    const uint64_t row = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (row >= m) return;
    uint32_t* outs[3] = {aA, aB, aC};
    if (row >= nc) {
        if (lane == 0) {
            Fr z = Fr::zero();
            Fr a = (row - nc <= ni) ? fp_load<FrCfg>(w + 8 * (row - nc)) : z;
            fp_store(aA + 8 * row, a);
            fp_store(aB + 8 * row, z);
            fp_store(aC + 8 * row, z);
        }
        return;
    }
#pragma unroll
    for (int mtx = 0; mtx < 3; ++mtx) {
        const uint64_t lo = csr.row[mtx][row], hi = csr.row[mtx][row + 1];
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

// h[i] = (a[i] * b_mont[i] - c[i]) [* zinv[i]]
__global__ void k_h_pointwise(uint32_t* a, const uint32_t* b, const uint32_t* c, const uint32_t* zinv, uint64_t m) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    Fr x = fp_sub(fp_mul(fp_load<FrCfg>(a + 8 * i), fp_load<FrCfg>(b + 8 * i)), fp_load<FrCfg>(c + 8 * i));
    if (zinv) x = fp_mul(x, fp_load<FrCfg>(zinv + 8 * i));
    fp_store(a + 8 * i, x);
}
__global__ void k_mul_table2(uint32_t* a, uint64_t n, const uint32_t* tbl) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, fp_mul(fp_load<FrCfg>(a + 8 * i), fp_load<FrCfg>(tbl + 8 * i)));
}
__global__ void k_fill_const(uint32_t* a, uint64_t n, Fr v) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, v);
}
__global__ void k_gather_fr(const uint32_t* __restrict__ w, const uint32_t* __restrict__ idx, uint64_t n, uint32_t* out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(out + 8 * i, fp_load<FrCfg>(w + 8 * uint64_t(idx[i])));
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

// out[i] (XYZZ, Montgomery) = scalars[i] * points[i];  points canonical affine (mont == 0) or Montgomery affine
template <class F>
__global__ void k_scalar_mul_batch(const uint64_t* __restrict__ points, const uint32_t* __restrict__ scalars, unsigned n,
                                   int mont, uint32_t* __restrict__ out) {
    // one job per block (lane 0 of one warp): independent scalars diverge at every add step, so jobs sharing a warp
    // would serialise; separate warps run concurrently on separate schedulers / SMs
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
    XYZZ<F> r = ec_scalar_mul<F>(p, k);
    r.store(out + uint64_t(i) * (XYZZ<F>::BYTES / 4));
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

// proof assembly, phase 1: three threads (separate warps)
//   t0: g_A  = alpha + evA  + r delta1      -> out_proof[0..12)  and scratch for phase 2
//   t1: g1_B = beta1 + evB1 + s delta1      -> scratch
//   t2: g_B  = beta2 + evB2 + s delta2      -> out_proof[12..36)
struct AsmPtrs {
    const uint32_t* evA;    // XYZZ G1
    const uint32_t* evB1;   // XYZZ G1
    const uint32_t* evB2;   // XYZZ G2
    const uint32_t* evH;    // XYZZ G1
    const uint32_t* evL;    // XYZZ G1
    const uint32_t* fx1;    // XYZZ G1 x4: r d1, s d1, (-rs) d1, r_saver gamma_inv (or infinity)
    const uint32_t* fx2;    // XYZZ G2 x1: s d2
    const uint32_t* fixed1; // affine Montgomery: alpha, beta1, delta1
    const uint32_t* fixed2; // affine Montgomery: beta2, delta2
    uint64_t* gA_gB1;       // scratch: canonical affine g_A, g1_B (2 x 12 limbs)
    uint64_t* proof;        // 48 limbs
};

template <class F>
VSP_DEV void write_canonical(uint64_t* out, const XYZZ<F>& p) {
    Affine<F> a = ec_to_affine<F>(p);
    a.x = F::from_mont(a.x);
    a.y = F::from_mont(a.y);
    a.store(out);
}

__global__ void __launch_bounds__(96) k_prove_phase1(AsmPtrs p) {
    const unsigned warp = threadIdx.x >> 5;
    if (threadIdx.x & 31) return;
    constexpr int L1 = G1XYZZ::BYTES / 4, A1 = G1Affine::BYTES / 4;
    if (warp == 0) {
        G1XYZZ acc = G1XYZZ::load(p.evA);
        ec_madd<FqOps>(acc, G1Affine::load(p.fixed1));
        ec_add<FqOps>(acc, G1XYZZ::load(p.fx1));
        write_canonical<FqOps>(p.gA_gB1, acc);
        write_canonical<FqOps>(p.proof, acc);
    } else if (warp == 1) {
        G1XYZZ acc = G1XYZZ::load(p.evB1);
        ec_madd<FqOps>(acc, G1Affine::load(p.fixed1 + A1));
        ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + L1));
        write_canonical<FqOps>(p.gA_gB1 + 12, acc);
    } else {
        G2XYZZ acc = G2XYZZ::load(p.evB2);
        ec_madd<Fq2Ops>(acc, G2Affine::load(p.fixed2));
        ec_add<Fq2Ops>(acc, G2XYZZ::load(p.fx2));
        write_canonical<Fq2Ops>(p.proof + 12, acc);
    }
}

// phase 3: g_C = evH + evL + s g_A + r g1_B + (-rs) delta1 + r_saver gamma_inv
__global__ void k_prove_phase3(AsmPtrs p, const uint32_t* sg) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    constexpr int L1 = G1XYZZ::BYTES / 4;
    G1XYZZ acc = G1XYZZ::load(p.evH);
    ec_add<FqOps>(acc, G1XYZZ::load(p.evL));
    ec_add<FqOps>(acc, G1XYZZ::load(sg));
    ec_add<FqOps>(acc, G1XYZZ::load(sg + L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + 2 * L1));
    ec_add<FqOps>(acc, G1XYZZ::load(p.fx1 + 3 * L1));
    write_canonical<FqOps>(p.proof + 36, acc);
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
    free_dev(pk->d_B_idx);
    free_dev(pk->d_gpow);
    free_dev(pk->d_ginvpow);
    free_dev(pk->d_zinv);
    delete pk;
}

extern "C" uint64_t vsp_pk_domain_size(const vsp_pk* pk) { return pk ? pk->m : 0; }

static int pk_load_impl(vsp_ctx* ctx, const vsp_proving_key_t* d, vsp_pk* pk) {
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
    if (d->H_count + 1 < pk->m) return set_error(ctx, VSP_E_INVALID, "H_query shorter than m - 1");
    cudaStream_t s = ctx->stream;
    const uint64_t* rows[3] = {cs.a_row, cs.b_row, cs.c_row};
    const uint32_t* cols[3] = {cs.a_col, cs.b_col, cs.c_col};
    const uint64_t* vals[3] = {cs.a_val, cs.b_val, cs.c_val};
    for (int i = 0; i < 3; ++i) {
        if (!rows[i]) return set_error(ctx, VSP_E_INVALID, "null CSR row pointer");
        const uint64_t nnz = rows[i][pk->nc];
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
    VSP_TRY(bases_upload(ctx, false, pk->nv + 1, d->A_query, &pk->A, true));
    VSP_TRY(bases_upload(ctx, false, d->B_count, d->B_g1, &pk->Bg1, true));
    VSP_TRY(bases_upload(ctx, true, d->B_count, d->B_g2, &pk->Bg2, true));
    VSP_TRY(bases_upload(ctx, false, pk->m - 1, d->H_query, &pk->H, true));
    VSP_TRY(bases_upload(ctx, false, pk->na, d->L_query, &pk->L, true));
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
    int rc = pk_load_impl(ctx, desc, pk);
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

static int coset_lde(vsp_ctx* ctx, cudaStream_t s, const vsp_pk* pk, uint32_t* d_a) {
    const uint64_t m = pk->m;
    if (pk->kind == VSP_DOMAIN_BASIC_RADIX2) {
        const unsigned L = ceil_log2_u64(m);
        VSP_TRY(ntt_pow2(ctx, s, d_a, d_a, L, true, nullptr, pk->d_gpow));  // inverse_fft then multiply_by_coset(g)
        VSP_TRY(ntt_pow2(ctx, s, d_a, d_a, L, false, nullptr, nullptr));    // fft
    } else {
        VSP_TRY(domain_fft_dev(ctx, s, m, true, d_a));
        VSP_LAUNCH_K(ctx, KC_QAP, k_mul_table2, (unsigned)((m + 127) / 128), 128, 0, s, d_a, m, pk->d_gpow);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        VSP_TRY(domain_fft_dev(ctx, s, m, false, d_a));
    }
    return VSP_OK;
}

// d_w: (1 || primary || aux) canonical.  d_h: m Fr (coefficients_for_H[0..m)), canonical.  Uses side streams 0 and 1.
static int witness_map_dev(vsp_ctx* ctx, const vsp_pk* pk, const uint32_t* d_w, uint32_t* d_h, DevBuf& aB, DevBuf& aC) {
    cudaStream_t s = ctx->stream;
    const uint64_t m = pk->m;
    VSP_TRY(aB.alloc(ctx, m * 32, s));
    VSP_TRY(aC.alloc(ctx, m * 32, s));
    CsrDev csr;
    for (int i = 0; i < 3; ++i) {
        csr.row[i] = pk->d_row[i];
        csr.col[i] = pk->d_col[i];
        csr.val[i] = pk->d_val[i];
    }
    uint32_t* aA = d_h;
    VSP_LAUNCH_K(ctx, KC_QAP, k_r1cs_eval, (unsigned)((m * 32 + 255) / 256), 256, 0, s, csr, d_w, pk->nc, pk->ni, m, aA,
               aB.as<uint32_t>(), aC.as<uint32_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_fork, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->side[0], ctx->ev_fork, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->side[1], ctx->ev_fork, 0));
    VSP_TRY(coset_lde(ctx, ctx->side[0], pk, aB.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[0], ctx->side[0]));
    VSP_TRY(coset_lde(ctx, ctx->side[1], pk, aC.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[1], ctx->side[1]));
    VSP_TRY(coset_lde(ctx, s, pk, aA));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[0], 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[1], 0));
    VSP_LAUNCH_K(ctx, KC_QAP, k_h_pointwise, (unsigned)((m + 127) / 128), 128, 0, s, aA, aB.as<uint32_t>(), aC.as<uint32_t>(),
               pk->d_zinv, m);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    if (pk->kind == VSP_DOMAIN_BASIC_RADIX2) {
        // inverse_fft, multiply_by_coset(g^-1) and divide_by_z_on_coset in one pass chain (table = Z^-1 g^-i)
        VSP_TRY(ntt_pow2(ctx, s, aA, aA, ceil_log2_u64(m), true, nullptr, pk->d_ginvpow));
    } else {
        VSP_TRY(domain_fft_dev(ctx, s, m, true, aA));
        VSP_LAUNCH_K(ctx, KC_QAP, k_mul_table2, (unsigned)((m + 127) / 128), 128, 0, s, aA, m, pk->d_ginvpow);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
    }
    return VSP_OK;
}

static int upload_witness(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux, DevBuf& d_w) {
    cudaStream_t s = ctx->stream;
    VSP_TRY(d_w.alloc(ctx, (pk->nv + 1) * 32, s));
    static const uint64_t one[4] = {1, 0, 0, 0};
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_w.p, one, 32, cudaMemcpyHostToDevice, s));
    if (pk->ni)
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_w.as<char>() + 32, primary, pk->ni * 32, cudaMemcpyHostToDevice, s));
    if (pk->na)
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_w.as<char>() + 32 + pk->ni * 32, aux, pk->na * 32, cudaMemcpyHostToDevice, s));
    return VSP_OK;
}

static bool fr_reduced(const uint64_t* x) { return !hostfr::geq(x, hostfr::MOD); }

// One proof, fully asynchronous: on return everything is enqueued and d_proof (48 limbs) will hold the result once
// ctx->stream completes.  `keep` owns the scratch so it outlives this function until the caller has synchronised.
struct ProofScratch {
    DevBuf w, h, aB, aC, bsc, ev, fxpts, fxsc, fx1, fx2, gab, sgpts, sgsc, sg;
};

static int prove_async(vsp_ctx* ctx, const vsp_pk* pk, const uint32_t* d_assignment, const uint64_t* primary,
                       const uint64_t* aux, const uint64_t* r, const uint64_t* sv, const uint64_t* r_saver,
                       const uint64_t* gamma_inv, uint64_t* d_proof, ProofScratch& k) {
    using namespace hostfr;
    if (!fr_reduced(r) || !fr_reduced(sv) || (r_saver && !fr_reduced(r_saver)))
        return set_error(ctx, VSP_E_INVALID, "proving randomness not reduced mod r");
    cudaStream_t s = ctx->stream;
    const uint64_t m = pk->m;
    constexpr size_t X1 = G1XYZZ::BYTES, X2 = G2XYZZ::BYTES;
    // the full assignment (1 || primary || auxiliary): already on the device, or uploaded from the host buffers
    const uint32_t* d_w = d_assignment;
    if (!d_w) {
        VSP_TRY(upload_witness(ctx, pk, primary, aux, k.w));
        d_w = k.w.as<uint32_t>();
    }

    // fork: multiexps that only need the witness start now
    VSP_TRY(k.ev.alloc(ctx, 4 * X1 + X2, s));  // evA, evB1, evH, evL (G1) | evB2 (G2)
    uint64_t* evA = k.ev.as<uint64_t>();
    uint64_t* evB1 = evA + X1 / 8;
    uint64_t* evH = evB1 + X1 / 8;
    uint64_t* evL = evH + X1 / 8;
    uint64_t* evB2 = evL + X1 / 8;
    VSP_TRY(k.bsc.alloc(ctx, (pk->B_count ? pk->B_count : 1) * 32, s));
    // fixed-base multiples: r d1, s d1, (-rs) d1, r_saver gamma_inv | s d2
    Fr4 fr, fs, frs;
    memcpy(fr.v, r, 32);
    memcpy(fs.v, sv, 32);
    frs = sub(from_u64(0), mul(fr, fs));
    uint64_t pts1[4 * 12], sc1[4 * 4];
    for (int i = 0; i < 3; ++i) memcpy(pts1 + 12 * i, pk->h_delta_g1, 96);
    memcpy(sc1, r, 32);
    memcpy(sc1 + 4, sv, 32);
    memcpy(sc1 + 8, frs.v, 32);
    if (r_saver && gamma_inv) {
        memcpy(pts1 + 36, gamma_inv, 96);
        memcpy(sc1 + 12, r_saver, 32);
    } else {
        memset(pts1 + 36, 0, 96);
        memset(sc1 + 12, 0, 32);
    }
    VSP_TRY(k.fxpts.alloc(ctx, sizeof(pts1) + 192, s));
    VSP_TRY(k.fxsc.alloc(ctx, sizeof(sc1) + 32, s));
    VSP_TRY(k.fx1.alloc(ctx, 4 * X1, s));
    VSP_TRY(k.fx2.alloc(ctx, X2, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(k.fxpts.p, pts1, sizeof(pts1), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(k.fxpts.as<char>() + sizeof(pts1), pk->h_delta_g2, 192, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(k.fxsc.p, sc1, sizeof(sc1), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(k.fxsc.as<char>() + sizeof(sc1), sv, 32, cudaMemcpyHostToDevice, s));
    VSP_TRY(k.h.alloc(ctx, m * 32, s));
    VSP_TRY(k.gab.alloc(ctx, 2 * 96, s));
    VSP_TRY(k.sgsc.alloc(ctx, 64, s));
    VSP_TRY(k.sg.alloc(ctx, 2 * X1, s));
    uint64_t sgs[8];
    memcpy(sgs, sv, 32);     // s * g_A
    memcpy(sgs + 4, r, 32);  // r * g1_B
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(k.sgsc.p, sgs, 64, cudaMemcpyHostToDevice, s));

    cudaEvent_t ev_w = ctx->ev_join[3];  // reuse as "witness ready" before side 2/3 start (re-recorded below)
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ev_w, s));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->side[2], ev_w, 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->side[3], ev_w, 0));
    // side 2: A, L
    VSP_TRY(msm_run(ctx, ctx->side[2], pk->A, pk->nv + 1, d_w, evA, true));
    VSP_TRY(msm_run(ctx, ctx->side[2], pk->L, pk->na, d_w + 8 * (pk->ni + 1), evL, true));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[2], ctx->side[2]));
    // side 3: fixed-base multiples, B (G2), B (G1)
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<Fq2Ops>, 1, 32, 0, ctx->side[3],
               reinterpret_cast<const uint64_t*>(k.fxpts.as<char>() + sizeof(pts1)),
               reinterpret_cast<const uint32_t*>(k.fxsc.as<char>() + sizeof(sc1)), 1u, 0, k.fx2.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, 4, 32, 0, ctx->side[3], k.fxpts.as<uint64_t>(), k.fxsc.as<uint32_t>(), 4u,
               0, k.fx1.as<uint32_t>());
    if (pk->B_count)
        VSP_LAUNCH_K(ctx, KC_QAP, k_gather_fr, (unsigned)((pk->B_count + 127) / 128), 128, 0, ctx->side[3], d_w, pk->d_B_idx,
                   pk->B_count, k.bsc.as<uint32_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_TRY(msm_run(ctx, ctx->side[3], pk->Bg2, pk->B_count, k.bsc.as<uint32_t>(), evB2, true));
    VSP_TRY(msm_run(ctx, ctx->side[3], pk->Bg1, pk->B_count, k.bsc.as<uint32_t>(), evB1, true));
    VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[3], ctx->side[3]));

    // main: witness map, H multiexp
    VSP_TRY(witness_map_dev(ctx, pk, d_w, k.h.as<uint32_t>(), k.aB, k.aC));
    VSP_TRY(msm_run(ctx, s, pk->H, m - 1, k.h.as<uint32_t>(), evH, true));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[2], 0));
    VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join[3], 0));

    AsmPtrs ap;
    ap.evA = reinterpret_cast<uint32_t*>(evA);
    ap.evB1 = reinterpret_cast<uint32_t*>(evB1);
    ap.evB2 = reinterpret_cast<uint32_t*>(evB2);
    ap.evH = reinterpret_cast<uint32_t*>(evH);
    ap.evL = reinterpret_cast<uint32_t*>(evL);
    ap.fx1 = k.fx1.as<uint32_t>();
    ap.fx2 = k.fx2.as<uint32_t>();
    ap.fixed1 = pk->fixed1.d_pts;
    ap.fixed2 = pk->fixed2.d_pts;
    ap.gA_gB1 = k.gab.as<uint64_t>();
    ap.proof = d_proof;
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase1, 1, 96, 0, s, ap);
    VSP_LAUNCH_K(ctx, KC_FIXED, k_scalar_mul_batch<FqOps>, 2, 32, 0, s, k.gab.as<uint64_t>(), k.sgsc.as<uint32_t>(), 2u, 0,
               k.sg.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_FIXED, k_prove_phase3, 1, 32, 0, s, ap, k.sg.as<uint32_t>());
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

}  // namespace vsp

extern "C" int vsp_witness_map(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux,
                               uint64_t* h_out) {
    if (!ctx || !pk || !h_out || (!primary && pk->ni) || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf d_w, d_h, aB, aC;
    VSP_TRY(upload_witness(ctx, pk, primary, aux, d_w));
    VSP_TRY(d_h.alloc(ctx, pk->m * 32, ctx->stream));
    VSP_TRY(witness_map_dev(ctx, pk, d_w.as<uint32_t>(), d_h.as<uint32_t>(), aB, aC));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(h_out, d_h.p, pk->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    memset(h_out + 4 * pk->m, 0, 32);  // coefficients_for_H[m] = 0 for d1 = d2 = d3 = 0
    return VSP_OK;
}

extern "C" int vsp_prove(vsp_ctx* ctx, const vsp_pk* pk, const uint64_t* primary, const uint64_t* aux,
                         const uint64_t r[4], const uint64_t s[4], const uint64_t* r_saver,
                         const uint64_t* gamma_inverse_sum_s_g1, uint64_t* out_proof) {
    if (!ctx || !pk || !r || !s || !out_proof || (!primary && pk->ni) || (!aux && pk->na)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    ProofScratch k;
    DevBuf d_proof;
    VSP_TRY(d_proof.alloc(ctx, 384, ctx->stream));
    VSP_TRY(prove_async(ctx, pk, nullptr, primary, aux, r, s, r_saver, gamma_inverse_sum_s_g1, d_proof.as<uint64_t>(), k));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out_proof, d_proof.p, 384, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 4; ++i) VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->side[i]));
    return VSP_OK;
}

static int prove_batch_impl(vsp_ctx* ctx, const vsp_pk* pk, uint64_t k, const uint64_t* primary, const uint64_t* aux,
                            const uint64_t* d_assignments, const uint64_t* rs, const uint64_t* gamma_inverse_sum_s_g1,
                            uint64_t* d_proofs) {
    // ballots are enqueued back to back without host synchronisation; scratch of ballot i is released (stream-ordered)
    // once ballot i + 1 has been enqueued, so the allocator pool recycles it.
    for (uint64_t i = 0; i < k; ++i) {
        ProofScratch sc;
        const uint64_t* rr = rs + 12 * i;
        const uint32_t* d_w =
            d_assignments ? reinterpret_cast<const uint32_t*>(d_assignments + 4 * (pk->nv + 1) * i) : nullptr;
        VSP_TRY(prove_async(ctx, pk, d_w, primary ? primary + 4 * pk->ni * i : nullptr, aux ? aux + 4 * pk->na * i : nullptr,
                            rr, rr + 4, gamma_inverse_sum_s_g1 ? rr + 8 : nullptr, gamma_inverse_sum_s_g1,
                            d_proofs + 48 * i, sc));
        // side streams must finish with this ballot's scratch before it is recycled
        for (int j = 0; j < 4; ++j) {
            VSP_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev_join[j], ctx->side[j]));
            VSP_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join[j], 0));
        }
    }
    return VSP_OK;
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

// device-resident variant: d_assignments = k full assignments (1 || primary || auxiliary), (num_inputs + num_aux + 1)
// canonical Fr each; d_proofs = k * 48 limbs on the device.  Stream-ordered: returns once everything is enqueued.
extern "C" int vsp_prove_batch_dev(vsp_ctx* ctx, const vsp_pk* pk, uint64_t k, const uint64_t* d_assignments,
                                   const uint64_t* rs, const uint64_t* gamma_inverse_sum_s_g1, uint64_t* d_proofs) {
    if (!ctx || !pk || !rs || !d_proofs || !d_assignments) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return prove_batch_impl(ctx, pk, k, nullptr, nullptr, d_assignments, rs, gamma_inverse_sum_s_g1, d_proofs);
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
    if (pub->msg_size >= pk->ni + 1 || memcmp(msg, primary, pub->msg_size * 32) != 0)
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
