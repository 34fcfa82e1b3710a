// keygen.cu — key generation either side of the ballot-proof path: the Groth16
// (r1cs_gg_ppzksnark) generator and the SAVER (elgamal_verifiable) key pair, as batched fixed-base multiples on the GPU.
//
// Replaces (paths relative to /root/reference):
//   zk::generate<r1cs_gg_ppzksnark<bls12<381>, ..., proving_mode::encrypted_input>>(constraint_system)
//       call site bin/cli/include/nil/vote_saver/common.hpp:916-917; the generator itself lives in the un-vendored
//       submodule crypto3-zk @ d335bb8b (.SUBMODULES.json:199) — restated:
//         u   = domain.evaluate_all_lagrange_polynomials(t)     (basic_radix2_domain_aux.hpp:85-140,
//                                                                 step_radix2_domain.hpp:168-206)
//         At[j] = sum_i u[i] A[i][j] (+ u[n_c + j] for j <= n_in), Bt, Ct likewise
//         A_query[j] = At[j] G1;  B_query[j] = (Bt[j] G2, Bt[j] G1), zero entries dropped (sparse vector)
//         gamma_ABC[j] = (beta At[j] + alpha Bt[j] + Ct[j]) / gamma G1   (j <= n_in)
//         L_query[j]   = (beta At[j] + alpha Bt[j] + Ct[j]) / delta G1   (j >  n_in)
//         H_query[i]   = t^i Z(t) / delta G1                             (i <  m - 1)
//       key layouts: libs/marshalling/zk/.../r1cs_gg_ppzksnark/fast_proving_key.hpp:66-103, verification_key.hpp
//   generate_keypair_op<elgamal_verifiable<Curve>>::process      libs/pubkey/.../elgamal_verifiable.hpp:218-271
//
// The reference draws t, alpha, beta, gamma, delta inside zk::generate; here they are explicit inputs so results can be
// compared bit for bit (as r, s are for the prover).
//
// Device plan: the Lagrange vector comes from one batch-inverted table (no per-element inversion), the three
// transposed sparse products run one warp per variable over a CSC copy of the constraint system, and every group
// element is a fixed-base multiple of a generator: one thread per scalar walks an 8-bit table T[w][d] = d 2^(8w) G
// (32 mixed additions, no doublings), then points go back to canonical affine 8 at a time on one shared inversion.
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
                     const hostfr::Fr4& k, bool canonical_out = false);
int inv_affine_table_dev(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_out, uint64_t n, const hostfr::Fr4& elt,
                         const hostfr::Fr4& A, const hostfr::Fr4& B);

namespace {

const uint64_t G1_GEN[12] = {0xfb3af00adb22c6bbULL, 0x6c55e83ff97a1aefULL, 0xa14e3a3f171bac58ULL, 0xc3688c4f9774b905ULL,
                             0x2695638c4fa9ac0fULL, 0x17f1d3a73197d794ULL, 0x0caa232946c5e7e1ULL, 0xd03cc744a2888ae4ULL,
                             0x00db18cb2c04b3edULL, 0xfcf5e095d5d00af6ULL, 0xa09e30ed741d8ae4ULL, 0x08b3f481e3aaa0f1ULL};
// x.c0 | x.c1 | y.c0 | y.c1
const uint64_t G2_GEN[24] = {
    0xd48056c8c121bdb8ULL, 0x0bac0326a805bbefULL, 0xb4510b647ae3d177ULL, 0xc6e47ad4fa403b02ULL, 0x260805272dc51051ULL, 0x024aa2b2f08f0a91ULL,
    0xe5ac7d055d042b7eULL, 0x334cf11213945d57ULL, 0xb5da61bbdc7f5049ULL, 0x596bd0d09920b61aULL, 0x7dacd3a088274f65ULL, 0x13e02b6052719f60ULL,
    0xe193548608b82801ULL, 0x923ac9cc3baca289ULL, 0x6d429a695160d12cULL, 0xadfd9baa8cbdd3a7ULL, 0x8cc9cdc6da2e351aULL, 0x0ce5d527727d6e11ULL,
    0xaaa9075ff05f79beULL, 0x3f370d275cec1da1ULL, 0x267492ab572e99abULL, 0xcb3e287e85a763afULL, 0x32acd2b02bc28b99ULL, 0x0606c4a02ea734ccULL};

// ---------------------------------------------------------------------------------------------
// Fr kernels
// ---------------------------------------------------------------------------------------------
__global__ void k_kg_to_mont(uint32_t* v, uint64_t n) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(v + 8 * i, fp_mul(fp_load<FrCfg>(v + 8 * i), Fr::r2()));
}
// a[i] *= tbl[i]
__global__ void k_kg_mul_table(uint32_t* a, uint64_t n, const uint32_t* tbl) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, fp_mul(fp_load<FrCfg>(a + 8 * i), fp_load<FrCfg>(tbl + 8 * i)));
}
// t lies in the domain: u[i] = (omega^i == t) ? scale : 0     (basic_radix2_domain_aux.hpp:97-110)
__global__ void k_kg_unit_where(uint32_t* u, uint64_t n, const uint32_t* pow_mont, Fr t_mont, Fr scale) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(u + 8 * i, fp_load<FrCfg>(pow_mont + 8 * i) == t_mont ? scale : Fr::zero());
}
__global__ void k_kg_fill(uint32_t* a, uint64_t n, Fr v) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, v);
}

struct CscDev {
    const uint64_t* ptr[3];  // n_v + 2
    const uint32_t* row[3];
    const uint32_t* val[3];  // Montgomery
};

VSP_DEV Fr kg_warp_sum(Fr a) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        Fr b;
#pragma unroll
        for (int k = 0; k < 8; ++k) b.l[k] = __shfl_down_sync(0xffffffffu, a.l[k], off);
        a = fp_add(a, b);
    }
    return a;
}

// one warp per variable j: abct[mtx][j] = sum over the column's entries of val * u[row]; At[j] += u[n_c + j], j <= n_in
__global__ void __launch_bounds__(256) k_kg_columns(CscDev csc, const uint32_t* __restrict__ u, uint64_t nc, uint64_t ni,
                                                    uint64_t nvp1, uint32_t* abct) {
    const uint64_t j = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (j >= nvp1) return;
#pragma unroll 1
    for (int mtx = 0; mtx < 3; ++mtx) {
        const uint64_t lo = csc.ptr[mtx][j], hi = csc.ptr[mtx][j + 1];
        Fr acc = Fr::zero();
        for (uint64_t k = lo + lane; k < hi; k += 32)
            acc = fp_add(acc, fp_mul(fp_load<FrCfg>(csc.val[mtx] + 8 * k), fp_load<FrCfg>(u + 8 * uint64_t(csc.row[mtx][k]))));
        acc = kg_warp_sum(acc);
        if (lane == 0) {
            if (mtx == 0 && j <= ni) acc = fp_add(acc, fp_load<FrCfg>(u + 8 * (nc + j)));
            fp_store(abct + 8 * (uint64_t(mtx) * nvp1 + j), acc);
        }
    }
}

// K[j] = (beta At[j] + alpha Bt[j] + Ct[j]) * (j <= n_in ? 1/gamma : 1/delta)     (constants in Montgomery form)
__global__ void k_kg_combine(const uint32_t* __restrict__ abct, uint64_t nvp1, uint64_t ni, Fr alpha, Fr beta, Fr ginv,
                             Fr dinv, uint32_t* out) {
    const uint64_t j = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= nvp1) return;
    Fr x = fp_mul(beta, fp_load<FrCfg>(abct + 8 * j));
    x = fp_add(x, fp_mul(alpha, fp_load<FrCfg>(abct + 8 * (nvp1 + j))));
    x = fp_add(x, fp_load<FrCfg>(abct + 8 * (2 * nvp1 + j)));
    fp_store(out + 8 * j, fp_mul(x, j <= ni ? ginv : dinv));
}

// ---------------------------------------------------------------------------------------------
// fixed-base multiples of a generator
// ---------------------------------------------------------------------------------------------
// tab[w][d - 1] = d * T1[w], d = 1 .. 255, from T1[w] = 2^(8 w) G (Montgomery affine)
template <class F>
__global__ void __launch_bounds__(64) k_gen_table(const uint32_t* __restrict__ t1, uint32_t* __restrict__ tab) {
    constexpr int AL = Affine<F>::BYTES / 4;
    const unsigned q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= 32 * 255) return;
    const unsigned w = q / 255, d = q - w * 255 + 1;
    const Affine<F> p = Affine<F>::load(t1 + uint64_t(w) * AL);
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int bit = 31 - __clz(d); bit >= 0; --bit) {
        acc = ec_dbl<F>(acc);
        if ((d >> bit) & 1u) ec_madd<F>(acc, p);
    }
    ec_to_affine<F>(acc).store(tab + uint64_t(q) * AL);
}

// out[i] (XYZZ, Montgomery) = scalars[i] * G: one thread per scalar, 32 table lookups + mixed additions
template <class F>
__global__ void __launch_bounds__(128) k_fixed_base_batch(const uint32_t* __restrict__ tab, const uint32_t* __restrict__ scalars,
                                                          uint64_t n, uint32_t* __restrict__ out) {
    constexpr int AL = Affine<F>::BYTES / 4, L = XYZZ<F>::BYTES / 4;
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t k[8];
    {
        const uint4 a = *reinterpret_cast<const uint4*>(scalars + 8 * i), b = *reinterpret_cast<const uint4*>(scalars + 8 * i + 4);
        k[0] = a.x; k[1] = a.y; k[2] = a.z; k[3] = a.w; k[4] = b.x; k[5] = b.y; k[6] = b.z; k[7] = b.w;
    }
    XYZZ<F> acc = XYZZ<F>::inf();
#pragma unroll 1
    for (unsigned w = 0; w < 32; ++w) {
        const uint32_t d = (k[w >> 2] >> ((w & 3) * 8)) & 0xffu;
        if (d) ec_madd<F>(acc, Affine<F>::load(tab + (uint64_t(w) * 255 + d - 1) * AL));
    }
    acc.store(out + i * L);
}

// canonical affine of n XYZZ points, AFF_CH per thread on one shared inversion (Montgomery's trick over ZZ * ZZZ)
constexpr int AFF_CH = 8;
template <class F>
__global__ void __launch_bounds__(64) k_batch_to_affine(const uint32_t* __restrict__ xyzz, uint64_t n, uint64_t* __restrict__ out) {
    using E = typename F::T;
    constexpr int L = XYZZ<F>::BYTES / 4, A8 = Affine<F>::BYTES / 8;
    const uint64_t base = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) * AFF_CH;
    if (base >= n) return;
    const int cnt = (int)(n - base < (uint64_t)AFF_CH ? n - base : AFF_CH);
    E den[AFF_CH], pre[AFF_CH];
    for (int q = 0; q < cnt; ++q) {
        const char* p = reinterpret_cast<const char*>(xyzz + (base + q) * L);
        const E zz = F::load(p + 2 * F::BYTES), zzz = F::load(p + 3 * F::BYTES);
        den[q] = F::is_zero(zz) ? F::one() : F::mul(zz, zzz);
        pre[q] = q ? F::mul(pre[q - 1], den[q]) : den[q];
    }
    E inv = F::inv(pre[cnt - 1]);
    for (int q = cnt - 1; q >= 0; --q) {
        const E dinv = q ? F::mul(inv, pre[q - 1]) : inv;
        inv = F::mul(inv, den[q]);
        const XYZZ<F> r = XYZZ<F>::load(xyzz + (base + q) * L);
        Affine<F> a = Affine<F>::inf();
        if (!r.is_inf()) {
            a.x = F::from_mont(F::mul(r.X, F::mul(dinv, r.ZZZ)));
            a.y = F::from_mont(F::mul(r.Y, F::mul(dinv, r.ZZ)));
        }
        a.store(out + (base + q) * A8);
    }
}

// the 8-bit generator table of one group, built on the device (G1 783 KB, G2 1.5 MB: L2-resident)
struct GenTable {
    uint32_t* d_tab = nullptr;
    ~GenTable() {
        if (d_tab) cudaFree(d_tab);
    }
};

template <class F>
int gen_table_build(vsp_ctx* ctx, cudaStream_t s, bool g2, GenTable* t) {
    BasesDev t1;
    t1.force_c = 8;
    int rc = bases_upload(ctx, g2, 1, g2 ? G2_GEN : G1_GEN, &t1, true);
    if (rc == VSP_OK && (!t1.table || t1.windows != 32)) rc = set_error(ctx, VSP_E_NOMEM, "generator table not built");
    if (rc == VSP_OK && cudaMalloc(&t->d_tab, size_t(32) * 255 * Affine<F>::BYTES) != cudaSuccess) {
        t->d_tab = nullptr;
        rc = set_error(ctx, VSP_E_NOMEM, "cudaMalloc(generator table) failed");
    }
    if (rc == VSP_OK) {
        VSP_LAUNCH_K(ctx, KC_FIXED, k_gen_table<F>, (32 * 255 + 63) / 64, 64, 0, s, t1.d_pts, t->d_tab);
        if (cudaStreamSynchronize(s) != cudaSuccess || cudaGetLastError() != cudaSuccess)
            rc = set_error(ctx, VSP_E_CUDA, "generator table kernel failed");
    }
    bases_free(&t1);
    return rc;
}

// d_out[i] (canonical affine) = d_scalars[i] * G, i < n
template <class F>
int fixed_base_batch(vsp_ctx* ctx, cudaStream_t s, const GenTable& t, const uint32_t* d_scalars, uint64_t n, uint64_t* d_out) {
    if (n == 0) return VSP_OK;
    DevBuf x;
    VSP_TRY(x.alloc(ctx, n * XYZZ<F>::BYTES, s));
    VSP_LAUNCH_K(ctx, KC_FIXED, k_fixed_base_batch<F>, (unsigned)((n + 127) / 128), 128, 0, s, t.d_tab, d_scalars, n,
                 x.as<uint32_t>());
    const uint64_t threads = (n + AFF_CH - 1) / AFF_CH;
    VSP_LAUNCH_K(ctx, KC_FIXED, k_batch_to_affine<F>, (unsigned)((threads + 63) / 64), 64, 0, s, x.as<uint32_t>(), n, d_out);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

inline Fr fr_mont_const(const hostfr::Fr4& x) {
    Fr r;
    memcpy(r.l, hostfr::to_mont(x).v, 32);
    return r;
}
inline Fr fr_canon_const(const hostfr::Fr4& x) {
    Fr r;
    memcpy(r.l, x.v, 32);
    return r;
}
inline bool fr4_reduced(const uint64_t* x) { return !hostfr::geq(x, hostfr::MOD); }
inline hostfr::Fr4 fr4_load(const uint64_t* x) {
    hostfr::Fr4 r;
    memcpy(r.v, x, 32);
    return r;
}
inline bool fr4_eq(const hostfr::Fr4& a, const hostfr::Fr4& b) { return memcmp(a.v, b.v, 32) == 0; }

// d_u[i] = scale * L_i(t) over the basic radix-2 domain of size n (canonical values); d_tmp: n Fr of scratch
//   L_i(t) = (t^n - 1)/n * omega^i / (t - omega^i)          basic_radix2_domain_aux.hpp:85-140
int lagrange_basic(vsp_ctx* ctx, cudaStream_t s, uint64_t n, const hostfr::Fr4& t, const hostfr::Fr4& scale, uint32_t* d_u,
                   uint32_t* d_tmp) {
    using namespace hostfr;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (n == 1) {
        VSP_LAUNCH_K(ctx, KC_QAP, k_kg_fill, 1, 128, 0, s, d_u, n, fr_canon_const(scale));
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        return VSP_OK;
    }
    const Fr4 omega = unity_root_pow2(ceil_log2_u64(n));
    const Fr4 tn = pow_u64(t, n);
    if (fr4_eq(tn, from_u64(1))) {
        VSP_TRY(fill_power_table(ctx, s, d_tmp, n, omega, from_u64(1)));
        VSP_LAUNCH_K(ctx, KC_QAP, k_kg_unit_where, grid, 128, 0, s, d_u, n, d_tmp, fr_mont_const(t), fr_canon_const(scale));
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        return VSP_OK;
    }
    const Fr4 l0 = mul(scale, mul(sub(tn, from_u64(1)), inv(from_u64(n))));
    const Fr4 zero = from_u64(0);
    VSP_TRY(inv_affine_table_dev(ctx, s, d_tmp, n, omega, sub(zero, from_u64(1)), sub(zero, t)));  // 1 / (t - omega^i)
    VSP_TRY(fill_power_table(ctx, s, d_u, n, omega, l0, true));                                    // l0 omega^i
    VSP_LAUNCH_K(ctx, KC_QAP, k_kg_mul_table, grid, 128, 0, s, d_u, n, d_tmp);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// d_u = evaluate_all_lagrange_polynomials(t) of the size-m domain (canonical); also Z(t)
int lagrange_domain(vsp_ctx* ctx, cudaStream_t s, int kind, uint64_t m, const hostfr::Fr4& t, uint32_t* d_u, uint32_t* d_tmp,
                    hostfr::Fr4* Zt) {
    using namespace hostfr;
    const Fr4 one = from_u64(1);
    if (kind == VSP_DOMAIN_BASIC_RADIX2) {
        *Zt = sub(pow_u64(t, m), one);
        return lagrange_basic(ctx, s, m, t, one, d_u, d_tmp);
    }
    // step_radix2_domain.hpp:168-206, 208-211
    const unsigned log_2big = ceil_log2_u64(m);
    const uint64_t big_m = uint64_t(1) << (log_2big - 1), small_m = m - big_m;
    const Fr4 omega = unity_root_pow2(log_2big), big_omega = mul(omega, omega);
    const Fr4 w_small = pow_u64(omega, small_m);
    const Fr4 L0 = sub(pow_u64(t, small_m), w_small);
    *Zt = mul(sub(pow_u64(t, big_m), one), L0);
    // big half: inner_big[i] * L0 / (big_omega^(small_m i) - w_small)
    VSP_TRY(lagrange_basic(ctx, s, big_m, t, L0, d_u, d_tmp));
    VSP_TRY(inv_affine_table_dev(ctx, s, d_tmp, big_m, pow_u64(big_omega, small_m), one, w_small));
    VSP_LAUNCH_K(ctx, KC_QAP, k_kg_mul_table, (unsigned)((big_m + 127) / 128), 128, 0, s, d_u, big_m, d_tmp);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    // small half: L1 * inner_small(t / omega),  L1 = (t^big_m - 1) / (omega^big_m - 1)
    const Fr4 L1 = mul(sub(pow_u64(t, big_m), one), inv(sub(pow_u64(omega, big_m), one)));
    return lagrange_basic(ctx, s, small_m, mul(t, inv(omega)), L1, d_u + 8 * big_m, d_tmp);
}

// CSR (host) -> CSC (device) of one constraint matrix
int upload_csc(vsp_ctx* ctx, cudaStream_t s, uint64_t nc, uint64_t nvp1, const uint64_t* row, const uint32_t* col,
               const uint64_t* val, DevBuf* d_ptr, DevBuf* d_row, DevBuf* d_val) {
    const uint64_t nnz = row[nc];
    std::vector<uint64_t> ptr(nvp1 + 1, 0), v(4 * (nnz ? nnz : 1));
    std::vector<uint32_t> r(nnz ? nnz : 1);
    for (uint64_t k = 0; k < nnz; ++k) {
        if (col[k] >= nvp1) return set_error(ctx, VSP_E_INVALID, "constraint refers to variable %u", col[k]);
        if (!fr4_reduced(val + 4 * k)) return set_error(ctx, VSP_E_INVALID, "constraint coefficient not reduced mod r");
        ++ptr[col[k] + 1];
    }
    for (uint64_t j = 0; j < nvp1; ++j) ptr[j + 1] += ptr[j];
    std::vector<uint64_t> fill(ptr.begin(), ptr.end() - 1);
    for (uint64_t i = 0; i < nc; ++i) {
        if (row[i + 1] < row[i] || row[i + 1] > nnz) return set_error(ctx, VSP_E_INVALID, "CSR row pointers not monotone");
        for (uint64_t k = row[i]; k < row[i + 1]; ++k) {
            const uint64_t dst = fill[col[k]]++;
            r[dst] = (uint32_t)i;
            memcpy(&v[4 * dst], val + 4 * k, 32);
        }
    }
    VSP_TRY(d_ptr->alloc(ctx, ptr.size() * 8, s));
    VSP_TRY(d_row->alloc(ctx, r.size() * 4, s));
    VSP_TRY(d_val->alloc(ctx, v.size() * 8, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_ptr->p, ptr.data(), ptr.size() * 8, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_row->p, r.data(), r.size() * 4, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_val->p, v.data(), v.size() * 8, cudaMemcpyHostToDevice, s));
    if (nnz) VSP_LAUNCH_K(ctx, KC_QAP, k_kg_to_mont, (unsigned)((nnz + 127) / 128), 128, 0, s, d_val->as<uint32_t>(), nnz);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));  // the host staging vectors die with this frame
    return VSP_OK;
}

int copy_out(vsp_ctx* ctx, cudaStream_t s, uint64_t* dst, const uint64_t* d_src, uint64_t limbs) {
    if (!limbs) return VSP_OK;
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(dst, d_src, limbs * 8, cudaMemcpyDeviceToHost, s));
    return VSP_OK;
}

int generate_impl(vsp_ctx* ctx, const vsp_r1cs_t* cs, const uint64_t* toxic, vsp_keypair_t* out) {
    using namespace hostfr;
    const uint64_t ni = cs->num_inputs, na = cs->num_aux, nc = cs->num_constraints, nv = ni + na, nvp1 = nv + 1;
    if (nvp1 >= (uint64_t(1) << 31)) return set_error(ctx, VSP_E_INVALID, "too many variables");
    int kind = 0;
    uint64_t m = 0;
    if (domain_plan(nc + ni + 1, &kind, &m) != VSP_OK || kind == VSP_DOMAIN_EXTENDED_RADIX2)
        return set_error(ctx, VSP_E_DOMAIN, "no executable evaluation domain for %llu constraints", (unsigned long long)nc);
    for (int i = 0; i < 5; ++i)
        if (!fr4_reduced(toxic + 4 * i)) return set_error(ctx, VSP_E_INVALID, "generator randomness not reduced mod r");
    const Fr4 t = fr4_load(toxic), alpha = fr4_load(toxic + 4), beta = fr4_load(toxic + 8), gamma = fr4_load(toxic + 12),
              delta = fr4_load(toxic + 16);
    if (is_zero(gamma) || is_zero(delta)) return set_error(ctx, VSP_E_INVALID, "gamma and delta must be invertible");
    if (!cs->a_row || !cs->b_row || !cs->c_row) return set_error(ctx, VSP_E_INVALID, "null CSR row pointer");
    cudaStream_t s = ctx->stream;

    // 1. u = L(t), Z(t)
    DevBuf d_u, d_tmp;
    VSP_TRY(d_u.alloc(ctx, m * 32, s));
    VSP_TRY(d_tmp.alloc(ctx, m * 32, s));
    Fr4 Zt;
    VSP_TRY(lagrange_domain(ctx, s, kind, m, t, d_u.as<uint32_t>(), d_tmp.as<uint32_t>(), &Zt));

    // 2. At | Bt | Ct over the CSC copy
    DevBuf ptr[3], row[3], val[3];
    const uint64_t* rows[3] = {cs->a_row, cs->b_row, cs->c_row};
    const uint32_t* cols[3] = {cs->a_col, cs->b_col, cs->c_col};
    const uint64_t* vals[3] = {cs->a_val, cs->b_val, cs->c_val};
    CscDev csc;
    for (int i = 0; i < 3; ++i) {
        VSP_TRY(upload_csc(ctx, s, nc, nvp1, rows[i], cols[i], vals[i], &ptr[i], &row[i], &val[i]));
        csc.ptr[i] = ptr[i].as<uint64_t>();
        csc.row[i] = row[i].as<uint32_t>();
        csc.val[i] = val[i].as<uint32_t>();
    }
    // G1 scalar list: At | Bt | K (gamma_ABC then L) | H | alpha beta delta gamma;  G2 list: Bt | beta delta gamma
    const uint64_t n1 = 3 * nvp1 + (m - 1) + 4, n2 = nvp1 + 3;
    DevBuf d_s1, d_s2;
    VSP_TRY(d_s1.alloc(ctx, n1 * 32, s));
    VSP_TRY(d_s2.alloc(ctx, n2 * 32, s));
    uint32_t* s1 = d_s1.as<uint32_t>();
    uint32_t* s2 = d_s2.as<uint32_t>();
    DevBuf d_ct;  // Ct sits outside the scalar list
    VSP_TRY(d_ct.alloc(ctx, 3 * nvp1 * 32, s));
    uint32_t* abct = d_ct.as<uint32_t>();
    VSP_LAUNCH_K(ctx, KC_QAP, k_kg_columns, (unsigned)((nvp1 * 32 + 255) / 256), 256, 0, s, csc, d_u.as<uint32_t>(), nc, ni, nvp1,
                 abct);
    const Fr4 ginv = inv(gamma), dinv = inv(delta);
    VSP_LAUNCH_K(ctx, KC_QAP, k_kg_combine, (unsigned)((nvp1 + 127) / 128), 128, 0, s, abct, nvp1, ni, fr_mont_const(alpha),
                 fr_mont_const(beta), fr_mont_const(ginv), fr_mont_const(dinv), s1 + 8 * (2 * nvp1));
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(s1, abct, 2 * nvp1 * 32, cudaMemcpyDeviceToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(s2, abct + 8 * nvp1, nvp1 * 32, cudaMemcpyDeviceToDevice, s));
    // H scalars: t^i Z(t) / delta
    VSP_TRY(fill_power_table(ctx, s, s1 + 8 * (3 * nvp1), m - 1, t, mul(Zt, dinv), true));
    uint64_t tail1[16], tail2[12];
    memcpy(tail1, alpha.v, 32);
    memcpy(tail1 + 4, beta.v, 32);
    memcpy(tail1 + 8, delta.v, 32);
    memcpy(tail1 + 12, gamma.v, 32);
    memcpy(tail2, beta.v, 32);
    memcpy(tail2 + 4, delta.v, 32);
    memcpy(tail2 + 8, gamma.v, 32);
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(s1 + 8 * (n1 - 4), tail1, sizeof(tail1), cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(s2 + 8 * nvp1, tail2, sizeof(tail2), cudaMemcpyHostToDevice, s));

    // 3. every group element is a fixed-base multiple of a generator
    GenTable tab1, tab2;
    VSP_TRY(gen_table_build<FqOps>(ctx, s, false, &tab1));
    VSP_TRY(gen_table_build<Fq2Ops>(ctx, s, true, &tab2));
    DevBuf d_p1, d_p2;
    VSP_TRY(d_p1.alloc(ctx, n1 * 96, s));
    VSP_TRY(d_p2.alloc(ctx, n2 * 192, s));
    VSP_TRY(fixed_base_batch<FqOps>(ctx, s, tab1, s1, n1, d_p1.as<uint64_t>()));
    VSP_TRY(fixed_base_batch<Fq2Ops>(ctx, s, tab2, s2, n2, d_p2.as<uint64_t>()));

    // 4. results to the caller's arrays; the sparse B_query is compacted on the host (zero entries are infinity)
    const uint64_t* p1 = d_p1.as<uint64_t>();
    const uint64_t* p2 = d_p2.as<uint64_t>();
    std::vector<uint64_t> b1(12 * nvp1), b2(24 * nvp1);
    VSP_TRY(copy_out(ctx, s, out->A_query, p1, 12 * nvp1));
    VSP_TRY(copy_out(ctx, s, b1.data(), p1 + 12 * nvp1, 12 * nvp1));
    VSP_TRY(copy_out(ctx, s, out->gamma_ABC_g1, p1 + 12 * (2 * nvp1), 12 * (ni + 1)));
    VSP_TRY(copy_out(ctx, s, out->L_query, p1 + 12 * (2 * nvp1 + ni + 1), 12 * na));
    VSP_TRY(copy_out(ctx, s, out->H_query, p1 + 12 * (3 * nvp1), 12 * (m - 1)));
    VSP_TRY(copy_out(ctx, s, out->alpha_g1, p1 + 12 * (n1 - 4), 12));
    VSP_TRY(copy_out(ctx, s, out->beta_g1, p1 + 12 * (n1 - 3), 12));
    VSP_TRY(copy_out(ctx, s, out->delta_g1, p1 + 12 * (n1 - 2), 12));
    VSP_TRY(copy_out(ctx, s, out->gamma_g1, p1 + 12 * (n1 - 1), 12));
    VSP_TRY(copy_out(ctx, s, b2.data(), p2, 24 * nvp1));
    VSP_TRY(copy_out(ctx, s, out->beta_g2, p2 + 24 * nvp1, 24));
    VSP_TRY(copy_out(ctx, s, out->delta_g2, p2 + 24 * (nvp1 + 1), 24));
    VSP_TRY(copy_out(ctx, s, out->gamma_g2, p2 + 24 * (nvp1 + 2), 24));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    uint64_t cnt = 0;
    for (uint64_t j = 0; j < nvp1; ++j) {
        bool inf = true;
        for (int q = 0; q < 12 && inf; ++q) inf = b1[12 * j + q] == 0;
        if (inf) continue;
        out->B_idx[cnt] = (uint32_t)j;
        memcpy(out->B_g1 + 12 * cnt, &b1[12 * j], 96);
        memcpy(out->B_g2 + 24 * cnt, &b2[24 * j], 192);
        ++cnt;
    }
    out->B_count = cnt;
    out->H_count = m - 1;
    return VSP_OK;
}

}  // namespace
}  // namespace vsp

using namespace vsp;

extern "C" int vsp_groth16_generate(vsp_ctx* ctx, const vsp_r1cs_t* cs, const uint64_t* toxic, vsp_keypair_t* out) {
    if (!ctx) return VSP_E_INVALID;
    if (!cs || !toxic || !out || !out->alpha_g1 || !out->beta_g1 || !out->beta_g2 || !out->delta_g1 || !out->delta_g2 ||
        !out->A_query || !out->B_idx || !out->B_g2 || !out->B_g1 || !out->H_query || (!out->L_query && cs->num_aux) ||
        !out->gamma_g2 || !out->gamma_g1 || !out->gamma_ABC_g1)
        return set_error(ctx, VSP_E_INVALID, "vsp_groth16_generate: null argument");
    cudaSetDevice(ctx->device);
    try {
        return generate_impl(ctx, cs, toxic, out);
    } catch (const std::bad_alloc&) {
        return set_error(ctx, VSP_E_NOMEM, "host allocation failed");
    }
}

// generate_keypair_op::process (elgamal_verifiable.hpp:218-271).  rnd = rho | t_0 | (s_i, v_i, t_i) for i < msg_size.
//   delta_s[i] = s_i delta;  t_g1[i] = t_i gamma_ABC.rest[i];  t_g2 = (t_0 G2, t_1 G2, ...)
//   delta_sum_s = t_0 delta + sum_i (s_i t_i) delta;  gamma_inverse_sum_s = -(gamma_g1 + sum_i s_i gamma_g1)
//   rho_g2 = rho G2;  rho_sv[i] = (s_i v_i) G2;  rho_rhov[i] = v_i rho_g2
// The sums over one base are taken on the scalars (group law: same point), the G2 side is fixed-base on the generator.
extern "C" int vsp_saver_keygen(vsp_ctx* ctx, uint64_t msg_size, const uint64_t* delta_g1, const uint64_t* gamma_g1,
                                const uint64_t* gamma_abc_rest, const uint64_t* rnd, vsp_saver_keypair_t* out) {
    using namespace hostfr;
    if (!ctx) return VSP_E_INVALID;
    if (!delta_g1 || !gamma_g1 || !rnd || !out || !out->t_g2 || !out->delta_sum_s_g1 || !out->gamma_inverse_sum_s_g1 ||
        !out->rho_g2 || !out->rho ||
        (msg_size && (!gamma_abc_rest || !out->delta_s_g1 || !out->t_g1 || !out->rho_sv_g2 || !out->rho_rhov_g2)))
        return set_error(ctx, VSP_E_INVALID, "vsp_saver_keygen: null argument");
    if (msg_size > 4096) return set_error(ctx, VSP_E_INVALID, "msg_size too large");
    cudaSetDevice(ctx->device);
    const unsigned n = (unsigned)msg_size;
    for (unsigned i = 0; i < 3 * n + 2; ++i)
        if (!fr4_reduced(rnd + 4 * i)) return set_error(ctx, VSP_E_INVALID, "randomness not reduced mod r");
    cudaStream_t s = ctx->stream;
    const Fr4 rho = fr4_load(rnd), t0 = fr4_load(rnd + 4);
    // G1 scalars: n x s_i (delta) | n x t_i (gamma_ABC.rest[i]) | t_0 + sum s_i t_i (delta) | -(1 + sum s_i) (gamma_g1)
    // G2 scalars: t_0, t_1 .. t_n | rho | n x s_i v_i | n x v_i rho
    std::vector<uint64_t> sc1(4 * size_t(2 * n + 2)), pt1(12 * size_t(2 * n + 2)), sc2(4 * size_t(3 * n + 2));
    Fr4 sum_st = t0, sum_s = from_u64(1);
    memcpy(&sc2[0], t0.v, 32);
    memcpy(&sc2[4 * size_t(n + 1)], rho.v, 32);
    for (unsigned i = 0; i < n; ++i) {
        const Fr4 si = fr4_load(rnd + 4 * (2 + 3 * i)), vi = fr4_load(rnd + 4 * (3 + 3 * i)), ti = fr4_load(rnd + 4 * (4 + 3 * i));
        memcpy(&sc1[4 * size_t(i)], si.v, 32);
        memcpy(&pt1[12 * size_t(i)], delta_g1, 96);
        memcpy(&sc1[4 * size_t(n + i)], ti.v, 32);
        memcpy(&pt1[12 * size_t(n + i)], gamma_abc_rest + 12 * i, 96);
        sum_st = add(sum_st, mul(si, ti));
        sum_s = add(sum_s, si);
        memcpy(&sc2[4 * size_t(1 + i)], ti.v, 32);
        memcpy(&sc2[4 * size_t(n + 2 + i)], mul(si, vi).v, 32);
        memcpy(&sc2[4 * size_t(2 * n + 2 + i)], mul(vi, rho).v, 32);
    }
    const Fr4 neg_sum_s = sub(from_u64(0), sum_s);
    memcpy(&sc1[4 * size_t(2 * n)], sum_st.v, 32);
    memcpy(&pt1[12 * size_t(2 * n)], delta_g1, 96);
    memcpy(&sc1[4 * size_t(2 * n + 1)], neg_sum_s.v, 32);
    memcpy(&pt1[12 * size_t(2 * n + 1)], gamma_g1, 96);

    // G2: fixed-base on the generator
    GenTable tab2;
    VSP_TRY(gen_table_build<Fq2Ops>(ctx, s, true, &tab2));
    DevBuf d_s2, d_p2;
    const uint64_t n2 = 3 * uint64_t(n) + 2;
    VSP_TRY(d_s2.alloc(ctx, n2 * 32, s));
    VSP_TRY(d_p2.alloc(ctx, n2 * 192, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_s2.p, sc2.data(), n2 * 32, cudaMemcpyHostToDevice, s));
    VSP_TRY(fixed_base_batch<Fq2Ops>(ctx, s, tab2, d_s2.as<uint32_t>(), n2, d_p2.as<uint64_t>()));
    const uint64_t* p2 = d_p2.as<uint64_t>();
    VSP_TRY(copy_out(ctx, s, out->t_g2, p2, 24 * uint64_t(n + 1)));
    VSP_TRY(copy_out(ctx, s, out->rho_g2, p2 + 24 * uint64_t(n + 1), 24));
    VSP_TRY(copy_out(ctx, s, out->rho_sv_g2, p2 + 24 * uint64_t(n + 2), 24 * uint64_t(n)));
    VSP_TRY(copy_out(ctx, s, out->rho_rhov_g2, p2 + 24 * uint64_t(2 * n + 2), 24 * uint64_t(n)));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));

    // G1: variable bases (delta_g1, gamma_g1, gamma_ABC.rest[i]) through the batched scalar-multiple entry point
    std::vector<uint64_t> r1(12 * size_t(2 * n + 2));
    VSP_TRY(vsp_g1_mul_batch(ctx, 2 * uint64_t(n) + 2, pt1.data(), sc1.data(), r1.data()));
    if (n) {
        memcpy(out->delta_s_g1, &r1[0], size_t(n) * 96);
        memcpy(out->t_g1, &r1[12 * size_t(n)], size_t(n) * 96);
    }
    memcpy(out->delta_sum_s_g1, &r1[12 * size_t(2 * n)], 96);
    memcpy(out->gamma_inverse_sum_s_g1, &r1[12 * size_t(2 * n + 1)], 96);
    memcpy(out->rho, rho.v, 32);
    return VSP_OK;
}

// out[i] = scalars[i] * G (canonical affine), fixed-base on the group generator: the generator's building block, exposed
// for tests and for callers that need commitments to many scalars (e.g. [t^i] G)
extern "C" int vsp_g1_generator_mul_batch(vsp_ctx* ctx, uint64_t n, const uint64_t* scalars, uint64_t* out) {
    if (!ctx) return VSP_E_INVALID;
    if ((!scalars || !out) && n) return set_error(ctx, VSP_E_INVALID, "null argument");
    cudaSetDevice(ctx->device);
    for (uint64_t i = 0; i < n; ++i)
        if (!fr4_reduced(scalars + 4 * i)) return set_error(ctx, VSP_E_INVALID, "scalar not reduced mod r");
    cudaStream_t s = ctx->stream;
    GenTable tab;
    VSP_TRY(gen_table_build<FqOps>(ctx, s, false, &tab));
    DevBuf d_s, d_p;
    VSP_TRY(d_s.alloc(ctx, n * 32, s));
    VSP_TRY(d_p.alloc(ctx, n * 96, s));
    if (n) VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_s.p, scalars, n * 32, cudaMemcpyHostToDevice, s));
    VSP_TRY(fixed_base_batch<FqOps>(ctx, s, tab, d_s.as<uint32_t>(), n, d_p.as<uint64_t>()));
    VSP_TRY(copy_out(ctx, s, out, d_p.as<uint64_t>(), 12 * n));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}

extern "C" int vsp_g2_generator_mul_batch(vsp_ctx* ctx, uint64_t n, const uint64_t* scalars, uint64_t* out) {
    if (!ctx) return VSP_E_INVALID;
    if ((!scalars || !out) && n) return set_error(ctx, VSP_E_INVALID, "null argument");
    cudaSetDevice(ctx->device);
    for (uint64_t i = 0; i < n; ++i)
        if (!fr4_reduced(scalars + 4 * i)) return set_error(ctx, VSP_E_INVALID, "scalar not reduced mod r");
    cudaStream_t s = ctx->stream;
    GenTable tab;
    VSP_TRY(gen_table_build<Fq2Ops>(ctx, s, true, &tab));
    DevBuf d_s, d_p;
    VSP_TRY(d_s.alloc(ctx, n * 32, s));
    VSP_TRY(d_p.alloc(ctx, n * 192, s));
    if (n) VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_s.p, scalars, n * 32, cudaMemcpyHostToDevice, s));
    VSP_TRY(fixed_base_batch<Fq2Ops>(ctx, s, tab, d_s.as<uint32_t>(), n, d_p.as<uint64_t>()));
    VSP_TRY(copy_out(ctx, s, out, d_p.as<uint64_t>(), 24 * n));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    return VSP_OK;
}
