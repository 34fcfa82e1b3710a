// ntt.cu — Fr radix-2 NTT / iNTT, coset shift and the evaluation-domain layer on sm_100a.
//
// Replaces (reference paths relative to /root/reference/libs/math/include/nil/crypto3/math):
//   detail::basic_radix2_fft                 domains/detail/basic_radix2_domain_aux.hpp:46-83
//   basic_radix2_domain::fft / inverse_fft   domains/basic_radix2_domain.hpp:69-96
//   step_radix2_domain::fft / inverse_fft    domains/step_radix2_domain.hpp:73-166
//   multiply_by_coset                        coset.hpp:37-43
//   divide_by_z_on_coset                     basic_radix2_domain.hpp:118-124, step_radix2_domain.hpp:222-246
//   make_evaluation_domain (size selection)  algorithms/make_evaluation_domain.hpp:51-106
//
// The reference bit-reverses then runs log n in-place radix-2 stages, recomputing each twiddle with a serial
// `w *= w_m`.  Here one transform is 1-3 Stockham passes: a pass loads 2^k strided values per sub-transform into
// shared memory (bit-reversed on the way in), runs k butterfly stages there, and writes the auto-sorted result,
// so input and output are both in natural order and no global bit-reversal pass exists.  Twiddles come from one
// table of powers of the 2^K-th root (K = largest size seen); every smaller root, its inverse and the inter-pass
// twiddles are strided / negated reads of that table.  Twiddles are kept in Montgomery form and data in whatever
// form the caller uses: montmul(w_mont, x) = w * x, so a transform maps canonical input to canonical output with
// no conversion passes.
#define VSP_DEFINE_CONSTANT_TABLES
#include "common.cuh"
#include "fp.cuh"
#include "host_fr.h"

#include <map>
#include <mutex>

namespace vsp {

struct StepDivTable {
    DevBuf inv_big;  // big_m Montgomery-form inverses of Z on the coset's first half
    hostfr::Fr4 z1_inv_mont;
};

struct NttTables {
    unsigned log_w = 0;       // table holds omega_{2^log_w}^e, e in [0, 2^(log_w-1))
    uint32_t* d_w = nullptr;  // Montgomery form, 8 limbs each
    std::map<uint64_t, StepDivTable*> step_div;
};

// ---------------------------------------------------------------------------------------------
// twiddle table
// ---------------------------------------------------------------------------------------------
__global__ void k_build_twiddles(uint32_t* w, unsigned log_w) {
    const uint64_t half = uint64_t(1) << (log_w - 1);
    const uint64_t e = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= half) return;
    // omega^e = prod over set bits b of e of omega^(2^b); omega^(2^b) is the primitive 2^(log_w-b)-th root
    Fr acc = Fr::one();
    for (unsigned b = 0; b + 1 < log_w; ++b) {
        if ((e >> b) & 1) {
            Fr f;
#pragma unroll
            for (int i = 0; i < 8; ++i) f.l[i] = FR_ROOT_OF_UNITY_POW2[log_w - b][i];
            acc = fp_mul(acc, f);
        }
    }
    fp_store(w + 8 * e, acc);
}

int ntt_tables_ensure(vsp_ctx* ctx, unsigned log_n) {
    if (!ctx->ntt) ctx->ntt = new NttTables();
    NttTables* t = ctx->ntt;
    if (log_n < 1) log_n = 1;
    if (t->d_w && t->log_w >= log_n) return VSP_OK;
    if (log_n > 32) return set_error(ctx, VSP_E_DOMAIN, "NTT size 2^%u exceeds the field's 2-adicity", log_n);
    // grow generously so mixed sizes do not rebuild repeatedly
    unsigned want = log_n < 16 ? 16 : log_n;
    if (t->d_w) {
        VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        for (int i = 0; i < 4; ++i)
            if (ctx->side[i]) VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->side[i]));
        cudaFree(t->d_w);
        t->d_w = nullptr;
    }
    const uint64_t half = uint64_t(1) << (want - 1);
    VSP_CUDA_CHECK(ctx, cudaMalloc(&t->d_w, half * 32));
    t->log_w = want;
    VSP_LAUNCH_K(ctx, KC_NTT, k_build_twiddles, (unsigned)((half + 255) / 256), 256, 0, ctx->stream, t->d_w, want);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    // side streams may use the table immediately: order them after the build
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

void ntt_tables_free(vsp_ctx* ctx) {
    if (!ctx->ntt) return;
    if (ctx->ntt->d_w) cudaFree(ctx->ntt->d_w);
    for (auto& kv : ctx->ntt->step_div) delete kv.second;
    delete ctx->ntt;
    ctx->ntt = nullptr;
}

// omega_N^e (N = 2^log_n, e in [0, N)), or its inverse, from the master table
VSP_DEV Fr twiddle(const uint32_t* __restrict__ w, unsigned log_w, unsigned log_n, uint32_t e, bool inverse) {
    const uint32_t n_mask = (log_n >= 32) ? 0xffffffffu : ((1u << log_n) - 1u);
    if (inverse) e = (0u - e) & n_mask;
    const uint32_t half = 1u << (log_n - 1);
    const bool neg = (e & half) != 0;
    e &= half - 1u;
    Fr t = fp_load<FrCfg>(w + 8 * (uint64_t(e) << (log_w - log_n)));
    return neg ? fp_neg(t) : t;
}

// ---------------------------------------------------------------------------------------------
// one Stockham pass of radix R = 2^k
// ---------------------------------------------------------------------------------------------
struct PassArgs {
    const uint32_t* in;
    uint32_t* out;
    const uint32_t* w;
    const uint32_t* pre;   // optional per-input-index multiplier (first pass only)
    const uint32_t* post;  // optional per-output-index multiplier (last pass only)
    unsigned log_w, log_n, k, ns_log, g_log;
    uint64_t in_stride, out_stride;  // elements between consecutive vectors of the batch (blockIdx.y)
    int inverse;
    int scale_log;  // multiply outputs by 2^-scale_log (last pass of an inverse transform), 0 = none
};

// shared layout: element e of sub-transform jj lives at lo[jj*R + e] / hi[jj*R + e] (two 16-byte halves kept in
// separate arrays so 128-bit accesses of consecutive elements are bank-conflict free)
__global__ void __launch_bounds__(512) k_ntt_pass(PassArgs a) {
    extern __shared__ uint4 smem[];
    const unsigned R = 1u << a.k, G = 1u << a.g_log;
    uint4* lo = smem;
    uint4* hi = smem + (size_t)G * R;
    uint4* tw_lo = smem + (size_t)2 * G * R;  // omega_R^e, e < R/2: every stage's butterfly twiddles, staged once
    uint4* tw_hi = tw_lo + (R >> 1);
    const unsigned T = blockDim.x;  // G * R / 2 (or 1 when R == 1... R >= 2 here)
    const unsigned t = threadIdx.x;
    const uint64_t nj = uint64_t(1) << (a.log_n - a.k);
    const uint64_t j0 = uint64_t(blockIdx.x) << a.g_log;
    const uint64_t ns_mask = (uint64_t(1) << a.ns_log) - 1;
    const uint32_t* vin = a.in + 8 * (uint64_t(blockIdx.y) * a.in_stride);
    uint32_t* vout = a.out + 8 * (uint64_t(blockIdx.y) * a.out_stride);

    // ---- load (+ coset pre-scale, + inter-pass twiddle), bit-reversed into shared memory
#pragma unroll
    for (unsigned h = 0; h < 2; ++h) {
        const unsigned q = t + h * T;
        const unsigned jj = q & (G - 1), r = q >> a.g_log;
        const uint64_t j = j0 + jj;
        const uint64_t gidx = j + uint64_t(r) * nj;
        Fr x = fp_load<FrCfg>(vin + 8 * gidx);
        if (a.pre) x = fp_mul(x, fp_load<FrCfg>(a.pre + 8 * gidx));
        if (a.ns_log) {
            // omega_{Ns*R}^{(j mod Ns) * r} = omega_N^{(j mod Ns) * r * N / (Ns * R)}
            const uint32_t e = (uint32_t)(((j & ns_mask) * r) << (a.log_n - a.ns_log - a.k));
            if (e) x = fp_mul(x, twiddle(a.w, a.log_w, a.log_n, e, a.inverse));
        }
        const unsigned rr = __brev(r) >> (32 - a.k);
        lo[jj * R + rr] = make_uint4(x.l[0], x.l[1], x.l[2], x.l[3]);
        hi[jj * R + rr] = make_uint4(x.l[4], x.l[5], x.l[6], x.l[7]);
    }
    for (unsigned e = t; e < (R >> 1); e += T) {
        const Fr w = twiddle(a.w, a.log_w, a.log_n, e << (a.log_n - a.k), a.inverse);
        tw_lo[e] = make_uint4(w.l[0], w.l[1], w.l[2], w.l[3]);
        tw_hi[e] = make_uint4(w.l[4], w.l[5], w.l[6], w.l[7]);
    }
    __syncthreads();

    // ---- k radix-2 DIT stages in shared memory
    for (unsigned s = 1; s <= a.k; ++s) {
        const unsigned m = 1u << (s - 1);
        const unsigned jj = t >> (a.k - 1), i = t & ((R >> 1) - 1);
        const unsigned jpos = i & (m - 1);
        const unsigned i0 = jj * R + ((i >> (s - 1)) << s) + jpos, i1 = i0 + m;
        Fr u, v;
        {
            uint4 p = lo[i0], q = hi[i0];
            u.l[0] = p.x; u.l[1] = p.y; u.l[2] = p.z; u.l[3] = p.w;
            u.l[4] = q.x; u.l[5] = q.y; u.l[6] = q.z; u.l[7] = q.w;
            p = lo[i1]; q = hi[i1];
            v.l[0] = p.x; v.l[1] = p.y; v.l[2] = p.z; v.l[3] = p.w;
            v.l[4] = q.x; v.l[5] = q.y; v.l[6] = q.z; v.l[7] = q.w;
        }
        if (jpos) {  // omega_{2^s}^jpos = omega_R^(jpos << (k - s))
            const uint4 p = tw_lo[jpos << (a.k - s)], q = tw_hi[jpos << (a.k - s)];
            Fr w;
            w.l[0] = p.x; w.l[1] = p.y; w.l[2] = p.z; w.l[3] = p.w;
            w.l[4] = q.x; w.l[5] = q.y; w.l[6] = q.z; w.l[7] = q.w;
            v = fp_mul(v, w);
        }
        Fr x = fp_add(u, v), y = fp_sub(u, v);
        lo[i0] = make_uint4(x.l[0], x.l[1], x.l[2], x.l[3]);
        hi[i0] = make_uint4(x.l[4], x.l[5], x.l[6], x.l[7]);
        lo[i1] = make_uint4(y.l[0], y.l[1], y.l[2], y.l[3]);
        hi[i1] = make_uint4(y.l[4], y.l[5], y.l[6], y.l[7]);
        __syncthreads();
    }

    // ---- auto-sorted store (+ inverse scaling, + coset post-scale)
    Fr scale;
    if (a.scale_log) {
#pragma unroll
        for (int i = 0; i < 8; ++i) scale.l[i] = FR_INV_POW2[a.scale_log][i];
    }
#pragma unroll
    for (unsigned h = 0; h < 2; ++h) {
        const unsigned q = t + h * T;
        unsigned jj, r;
        if (a.ns_log == 0) {
            r = q & (R - 1);
            jj = q >> a.k;
        } else {
            jj = q & (G - 1);
            r = q >> a.g_log;
        }
        const uint64_t j = j0 + jj;
        const uint64_t didx = ((j >> a.ns_log) << (a.ns_log + a.k)) + (j & ns_mask) + (uint64_t(r) << a.ns_log);
        uint4 p = lo[jj * R + r], qq = hi[jj * R + r];
        Fr y;
        y.l[0] = p.x; y.l[1] = p.y; y.l[2] = p.z; y.l[3] = p.w;
        y.l[4] = qq.x; y.l[5] = qq.y; y.l[6] = qq.z; y.l[7] = qq.w;
        if (a.scale_log) y = fp_mul(y, scale);
        if (a.post) y = fp_mul(y, fp_load<FrCfg>(a.post + 8 * didx));
        fp_store(vout + 8 * didx, y);
    }
}

// The same pass with radix-4 butterflies: a thread owns four elements and runs TWO stages on them in registers between
// shared-memory exchanges (half the shared traffic and barriers of the radix-2 kernel); an odd k ends with one radix-2
// stage.  Launched with G * R / 4 threads; k >= 2.
// 16-byte slot i of a shared vector lives at swz(i): the radix-4 stages read slots 4 apart (stage pair 1) and runs of 4
// slots 16 apart (stage pair 2); XOR-ing bits 2..4 into bits 0..2 spreads both patterns (and contiguous runs) over all
// eight 16-byte bank groups of a quarter-warp
VSP_DEV unsigned swz(unsigned i) { return i ^ ((i >> 2) & 7u); }
VSP_DEV Fr sm_load(const uint4* lo, const uint4* hi, unsigned i) {
    i = swz(i);
    const uint4 p = lo[i], q = hi[i];
    Fr x;
    x.l[0] = p.x; x.l[1] = p.y; x.l[2] = p.z; x.l[3] = p.w;
    x.l[4] = q.x; x.l[5] = q.y; x.l[6] = q.z; x.l[7] = q.w;
    return x;
}
VSP_DEV void sm_store(uint4* lo, uint4* hi, unsigned i, const Fr& x) {
    i = swz(i);
    lo[i] = make_uint4(x.l[0], x.l[1], x.l[2], x.l[3]);
    hi[i] = make_uint4(x.l[4], x.l[5], x.l[6], x.l[7]);
}
__global__ void __launch_bounds__(512) k_ntt_pass4(PassArgs a) {
    extern __shared__ uint4 smem[];
    const unsigned R = 1u << a.k, G = 1u << a.g_log;
    uint4* lo = smem;
    uint4* hi = smem + (size_t)G * R;
    uint4* tw_lo = smem + (size_t)2 * G * R;
    uint4* tw_hi = tw_lo + (R >> 1);
    const unsigned T = blockDim.x;  // G * R / 4
    const unsigned t = threadIdx.x;
    const uint64_t nj = uint64_t(1) << (a.log_n - a.k);
    const uint64_t j0 = uint64_t(blockIdx.x) << a.g_log;
    const uint64_t ns_mask = (uint64_t(1) << a.ns_log) - 1;
    const uint32_t* vin = a.in + 8 * (uint64_t(blockIdx.y) * a.in_stride);
    uint32_t* vout = a.out + 8 * (uint64_t(blockIdx.y) * a.out_stride);

#pragma unroll
    for (unsigned h = 0; h < 4; ++h) {
        const unsigned q = t + h * T;
        const unsigned jj = q & (G - 1), r = q >> a.g_log;
        const uint64_t j = j0 + jj;
        const uint64_t gidx = j + uint64_t(r) * nj;
        Fr x = fp_load<FrCfg>(vin + 8 * gidx);
        if (a.pre) x = fp_mul(x, fp_load<FrCfg>(a.pre + 8 * gidx));
        if (a.ns_log) {
            const uint32_t e = (uint32_t)(((j & ns_mask) * r) << (a.log_n - a.ns_log - a.k));
            if (e) x = fp_mul(x, twiddle(a.w, a.log_w, a.log_n, e, a.inverse));
        }
        sm_store(lo, hi, jj * R + (__brev(r) >> (32 - a.k)), x);
    }
    for (unsigned e = t; e < (R >> 1); e += T) {
        const Fr w = twiddle(a.w, a.log_w, a.log_n, e << (a.log_n - a.k), a.inverse);
        sm_store(tw_lo, tw_hi, e, w);
    }
    __syncthreads();

    const unsigned jj4 = t >> (a.k - 2), g = t & ((R >> 2) - 1);
    unsigned s = 1;
    for (; s + 1 <= a.k; s += 2) {
        const unsigned m = 1u << (s - 1);
        const unsigned jpos = g & (m - 1);
        const unsigned base = jj4 * R + ((g >> (s - 1)) << (s + 1)) + jpos;
        Fr e0 = sm_load(lo, hi, base), e1 = sm_load(lo, hi, base + m), e2 = sm_load(lo, hi, base + 2 * m),
           e3 = sm_load(lo, hi, base + 3 * m);
        if (jpos) {  // stage s: omega_{2^s}^jpos on both pairs
            const Fr w1 = sm_load(tw_lo, tw_hi, jpos << (a.k - s));
            e1 = fp_mul(e1, w1);
            e3 = fp_mul(e3, w1);
        }
        Fr x = fp_add(e0, e1);
        e1 = fp_sub(e0, e1);
        e0 = x;
        x = fp_add(e2, e3);
        e3 = fp_sub(e2, e3);
        e2 = x;
        // stage s + 1: pairs (e0, e2) with omega_{2^(s+1)}^jpos and (e1, e3) with omega_{2^(s+1)}^(jpos + m)
        if (jpos) e2 = fp_mul(e2, sm_load(tw_lo, tw_hi, jpos << (a.k - s - 1)));
        e3 = fp_mul(e3, sm_load(tw_lo, tw_hi, (jpos + m) << (a.k - s - 1)));
        x = fp_add(e0, e2);
        e2 = fp_sub(e0, e2);
        e0 = x;
        x = fp_add(e1, e3);
        e3 = fp_sub(e1, e3);
        e1 = x;
        sm_store(lo, hi, base, e0);
        sm_store(lo, hi, base + m, e1);
        sm_store(lo, hi, base + 2 * m, e2);
        sm_store(lo, hi, base + 3 * m, e3);
        __syncthreads();
    }
    if (s == a.k) {  // odd k: the last stage as two radix-2 butterflies per thread
        const unsigned m = 1u << (s - 1);
#pragma unroll
        for (unsigned h = 0; h < 2; ++h) {
            const unsigned b = t + h * T;  // butterfly index over G * R / 2
            const unsigned jj = b >> (a.k - 1), i = b & ((R >> 1) - 1);
            const unsigned i0 = jj * R + i, i1 = i0 + m;  // s == k: one group per sub-transform, jpos = i
            Fr u = sm_load(lo, hi, i0), v = sm_load(lo, hi, i1);
            if (i) v = fp_mul(v, sm_load(tw_lo, tw_hi, i));
            sm_store(lo, hi, i0, fp_add(u, v));
            sm_store(lo, hi, i1, fp_sub(u, v));
        }
        __syncthreads();
    }

    Fr scale;
    if (a.scale_log) {
#pragma unroll
        for (int i = 0; i < 8; ++i) scale.l[i] = FR_INV_POW2[a.scale_log][i];
    }
#pragma unroll
    for (unsigned h = 0; h < 4; ++h) {
        const unsigned q = t + h * T;
        unsigned jj, r;
        if (a.ns_log == 0) {
            r = q & (R - 1);
            jj = q >> a.k;
        } else {
            jj = q & (G - 1);
            r = q >> a.g_log;
        }
        const uint64_t j = j0 + jj;
        const uint64_t didx = ((j >> a.ns_log) << (a.ns_log + a.k)) + (j & ns_mask) + (uint64_t(r) << a.ns_log);
        Fr y = sm_load(lo, hi, jj * R + r);
        if (a.scale_log) y = fp_mul(y, scale);
        if (a.post) y = fp_mul(y, fp_load<FrCfg>(a.post + 8 * didx));
        fp_store(vout + 8 * didx, y);
    }
}

__global__ void k_scale_copy(const uint32_t* in, uint32_t* out, uint64_t n, const uint32_t* pre, const uint32_t* post,
                             uint64_t in_stride, uint64_t out_stride) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    in += 8 * (uint64_t(blockIdx.y) * in_stride);
    out += 8 * (uint64_t(blockIdx.y) * out_stride);
    Fr x = fp_load<FrCfg>(in + 8 * i);
    if (pre) x = fp_mul(x, fp_load<FrCfg>(pre + 8 * i));
    if (post) x = fp_mul(x, fp_load<FrCfg>(post + 8 * i));
    fp_store(out + 8 * i, x);
}

int ntt_pow2(vsp_ctx* ctx, cudaStream_t s, const uint32_t* d_in, uint32_t* d_out, unsigned log_n, bool inverse,
             const uint32_t* d_pre, const uint32_t* d_post, unsigned batch, uint64_t in_stride, uint64_t out_stride) {
    const uint64_t n = uint64_t(1) << log_n;
    if (batch == 0) return VSP_OK;
    if (batch > 65535) return set_error(ctx, VSP_E_INVALID, "NTT batch too large");
    if (in_stride == 0) in_stride = n;
    if (out_stride == 0) out_stride = n;
    if (log_n == 0) {
        if (d_in != d_out || d_pre || d_post) {
            VSP_LAUNCH_K(ctx, KC_NTT, k_scale_copy, dim3(1, batch), 32, 0, s, d_in, d_out, n, d_pre, d_post, in_stride,
                         out_stride);
            VSP_CUDA_CHECK(ctx, cudaGetLastError());
        }
        return VSP_OK;
    }
    VSP_TRY(ntt_tables_ensure(ctx, log_n));
    NttTables* tb = ctx->ntt;
    unsigned ks[3], np;
    if (log_n <= 10) {
        np = 1;
        ks[0] = log_n;
    } else if (log_n <= 20) {
        np = 2;
        ks[0] = (log_n + 1) / 2;
        ks[1] = log_n - ks[0];
    } else {
        np = 3;
        ks[0] = (log_n + 2) / 3;
        ks[1] = (log_n - ks[0] + 1) / 2;
        ks[2] = log_n - ks[0] - ks[1];
    }
    DevBuf tmp_a, tmp_b;
    const uint32_t* src = d_in;
    uint64_t src_stride = in_stride;
    unsigned ns_log = 0;
    for (unsigned p = 0; p < np; ++p) {
        const unsigned k = ks[p];
        // G sub-transforms per block: 2^block_log values (128 threads on four values each); 9 measured best of 9/10/11
        static const unsigned block_log = getenv("VSP_NTT_BLOCK_LOG") ? (unsigned)atoi(getenv("VSP_NTT_BLOCK_LOG")) : 9u;
        unsigned g_log = (k >= block_log) ? 0 : (block_log - k);
        if (g_log > log_n - k) g_log = log_n - k;
        const bool last = (p + 1 == np);
        uint32_t* dst;
        uint64_t dst_stride = n;
        const bool single_block = (log_n - k - g_log) == 0;
        if (last) {
            dst = d_out;
            dst_stride = out_stride;
            if (np == 1 && d_in == d_out && !single_block) {
                VSP_TRY(tmp_a.alloc(ctx, n * batch * 32, s));
                dst = tmp_a.as<uint32_t>();
                dst_stride = n;
            }
        } else if (p == 0) {
            VSP_TRY(tmp_a.alloc(ctx, n * batch * 32, s));
            dst = tmp_a.as<uint32_t>();
        } else {
            VSP_TRY(tmp_b.alloc(ctx, n * batch * 32, s));
            dst = tmp_b.as<uint32_t>();
        }
        PassArgs a;
        a.in = src;
        a.out = dst;
        a.w = tb->d_w;
        a.pre = (p == 0) ? d_pre : nullptr;
        a.post = last ? d_post : nullptr;
        a.log_w = tb->log_w;
        a.log_n = log_n;
        a.k = k;
        a.ns_log = ns_log;
        a.g_log = g_log;
        a.in_stride = src_stride;
        a.out_stride = dst_stride;
        a.inverse = inverse ? 1 : 0;
        a.scale_log = (inverse && last) ? (int)log_n : 0;
        const unsigned blocks = 1u << (log_n - k - g_log);
        const size_t smem = ((size_t)32 << (k + g_log)) + ((size_t)16 << k);
        if (k >= 2 && !getenv("VSP_NTT_RADIX2")) {
            if (smem > 48 * 1024) cudaFuncSetAttribute(k_ntt_pass4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            VSP_LAUNCH_K(ctx, KC_NTT, k_ntt_pass4, dim3(blocks, batch), 1u << (k - 2 + g_log), smem, s, a);
        } else {
            VSP_LAUNCH_K(ctx, KC_NTT, k_ntt_pass, dim3(blocks, batch), 1u << (k - 1 + g_log), smem, s, a);
        }
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        if (last && dst != d_out) {
            if (out_stride == n)
                VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_out, dst, n * batch * 32, cudaMemcpyDeviceToDevice, s));
            else
                VSP_CUDA_CHECK(ctx, cudaMemcpy2DAsync(d_out, out_stride * 32, dst, n * 32, n * 32, batch,
                                                      cudaMemcpyDeviceToDevice, s));
        }
        src = dst;
        src_stride = dst_stride;
        ns_log += k;
    }
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// domain selection (make_evaluation_domain.hpp:51-106, type_traits.hpp:56-85; s = 32 for Fr)
// ---------------------------------------------------------------------------------------------
static bool is_basic(uint64_t m) { return m > 1 && ceil_log2_u64(m) <= 32 && m == (uint64_t(1) << ceil_log2_u64(m)); }
static bool is_extended(uint64_t m) {
    if (m <= 1) return false;
    const uint64_t small_m = m / 2;
    return ceil_log2_u64(m) == 33 && small_m == (uint64_t(1) << ceil_log2_u64(small_m));
}
static bool is_step(uint64_t m) {
    if (m <= 1) return false;
    const unsigned log_m = ceil_log2_u64(m);
    const uint64_t small_m = m - (uint64_t(1) << (log_m - 1));
    if (small_m == 0) return false;
    return small_m == (uint64_t(1) << ceil_log2_u64(small_m)) && log_m <= 32;
}

int domain_plan(uint64_t min_size, int* kind, uint64_t* m) {
    if (min_size <= 1) return VSP_E_DOMAIN;
    const uint64_t big = uint64_t(1) << (ceil_log2_u64(min_size) - 1);
    const uint64_t rounded_small = uint64_t(1) << ceil_log2_u64(min_size - big);
    const uint64_t cand[2] = {min_size, big + rounded_small};
    for (int i = 0; i < 2; ++i) {
        if (is_basic(cand[i])) { *kind = VSP_DOMAIN_BASIC_RADIX2; *m = cand[i]; return VSP_OK; }
        if (is_extended(cand[i])) { *kind = VSP_DOMAIN_EXTENDED_RADIX2; *m = cand[i]; return VSP_OK; }
        if (is_step(cand[i])) { *kind = VSP_DOMAIN_STEP_RADIX2; *m = cand[i]; return VSP_OK; }
    }
    return VSP_E_DOMAIN;
}

// ---------------------------------------------------------------------------------------------
// step_radix2_domain kernels (step_radix2_domain.hpp:73-166)
// ---------------------------------------------------------------------------------------------
// every kernel below takes the batch in blockIdx.y: vector v of `a` starts at element v * a_stride, the scratch
// vectors c / u0 (big_m each) and e / u1 (small_m each) are packed
// c[i] = i < small ? a[i] + a[i + big] : a[i]
__global__ void k_step_fft_c(const uint32_t* a, uint32_t* c, uint64_t big_m, uint64_t small_m, uint64_t a_stride) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= big_m) return;
    a += 8 * (uint64_t(blockIdx.y) * a_stride);
    c += 8 * (uint64_t(blockIdx.y) * big_m);
    Fr x = fp_load<FrCfg>(a + 8 * i);
    if (i < small_m) x = fp_add(x, fp_load<FrCfg>(a + 8 * (i + big_m)));
    fp_store(c + 8 * i, x);
}
// e[i] = sum_j omega^(i + j small) * (idx < small ? a[idx] - a[idx + big] : a[idx]),  idx = i + j small
__global__ void k_step_fft_e(const uint32_t* a, uint32_t* e, uint64_t big_m, uint64_t small_m, const uint32_t* w,
                             unsigned log_w, unsigned log_2big, uint64_t a_stride) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= small_m) return;
    a += 8 * (uint64_t(blockIdx.y) * a_stride);
    e += 8 * (uint64_t(blockIdx.y) * small_m);
    Fr acc = Fr::zero();
    for (uint64_t idx = i; idx < big_m; idx += small_m) {
        Fr x = fp_load<FrCfg>(a + 8 * idx);
        if (idx < small_m) x = fp_sub(x, fp_load<FrCfg>(a + 8 * (idx + big_m)));
        if (idx) x = fp_mul(x, twiddle(w, log_w, log_2big, (uint32_t)idx, false));
        acc = fp_add(acc, x);
    }
    fp_store(e + 8 * i, acc);
}
// inverse: combine U0 (big_m, already scaled) and U1 (small_m, already scaled) into a
__global__ void k_step_ifft_combine(const uint32_t* u0, const uint32_t* u1, uint32_t* a, uint64_t big_m,
                                    uint64_t small_m, const uint32_t* w, unsigned log_w, unsigned log_2big,
                                    uint64_t a_stride) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= big_m) return;
    a += 8 * (uint64_t(blockIdx.y) * a_stride);
    u0 += 8 * (uint64_t(blockIdx.y) * big_m);
    u1 += 8 * (uint64_t(blockIdx.y) * small_m);
    Fr x0 = fp_load<FrCfg>(u0 + 8 * i);
    if (i >= small_m) {
        fp_store(a + 8 * i, x0);
        return;
    }
    Fr x1 = fp_load<FrCfg>(u1 + 8 * i);
    for (uint64_t idx = i + small_m; idx < big_m; idx += small_m) {
        Fr t = fp_mul(fp_load<FrCfg>(u0 + 8 * idx), twiddle(w, log_w, log_2big, (uint32_t)idx, false));
        x1 = fp_sub(x1, t);
    }
    if (i) x1 = fp_mul(x1, twiddle(w, log_w, log_2big, (uint32_t)i, true));
    Fr half;
#pragma unroll
    for (int k = 0; k < 8; ++k) half.l[k] = FR_INV_POW2[1][k];
    fp_store(a + 8 * i, fp_mul(fp_add(x0, x1), half));
    fp_store(a + 8 * (i + big_m), fp_mul(fp_sub(x0, x1), half));
}

// fft / inverse_fft of `batch` vectors of the size-m domain, in place; vector v starts at element v * stride
int domain_fft_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, bool inverse, uint32_t* d, unsigned batch, uint64_t stride) {
    if (stride == 0) stride = m;
    if (batch == 0) return VSP_OK;
    if (batch > 65535) return set_error(ctx, VSP_E_INVALID, "NTT batch too large");
    if (is_basic(m)) return ntt_pow2(ctx, s, d, d, ceil_log2_u64(m), inverse, nullptr, nullptr, batch, stride, stride);
    if (!is_step(m)) return set_error(ctx, VSP_E_DOMAIN, "no executable radix-2 domain of size %llu", (unsigned long long)m);
    const unsigned log_2big = ceil_log2_u64(m);
    const uint64_t big_m = uint64_t(1) << (log_2big - 1), small_m = m - big_m;
    const unsigned log_big = log_2big - 1, log_small = ceil_log2_u64(small_m);
    VSP_TRY(ntt_tables_ensure(ctx, log_2big));
    NttTables* tb = ctx->ntt;
    DevBuf c, e;
    VSP_TRY(c.alloc(ctx, big_m * batch * 32, s));
    VSP_TRY(e.alloc(ctx, small_m * batch * 32, s));
    const dim3 gb((unsigned)((big_m + 127) / 128), batch), gs((unsigned)((small_m + 127) / 128), batch);
    if (!inverse) {
        VSP_LAUNCH_K(ctx, KC_NTT, k_step_fft_c, gb, 128, 0, s, d, c.as<uint32_t>(), big_m, small_m, stride);
        VSP_LAUNCH_K(ctx, KC_NTT, k_step_fft_e, gs, 128, 0, s, d, e.as<uint32_t>(), big_m, small_m, tb->d_w, tb->log_w,
                     log_2big, stride);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        VSP_TRY(ntt_pow2(ctx, s, c.as<uint32_t>(), d, log_big, false, nullptr, nullptr, batch, big_m, stride));
        VSP_TRY(ntt_pow2(ctx, s, e.as<uint32_t>(), d + 8 * big_m, log_small, false, nullptr, nullptr, batch, small_m, stride));
    } else {
        VSP_TRY(ntt_pow2(ctx, s, d, c.as<uint32_t>(), log_big, true, nullptr, nullptr, batch, stride, big_m));
        VSP_TRY(ntt_pow2(ctx, s, d + 8 * big_m, e.as<uint32_t>(), log_small, true, nullptr, nullptr, batch, stride, small_m));
        VSP_LAUNCH_K(ctx, KC_NTT, k_step_ifft_combine, gb, 128, 0, s, c.as<uint32_t>(), e.as<uint32_t>(), d, big_m, small_m,
                     tb->d_w, tb->log_w, log_2big, stride);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
    }
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// multiply_by_coset (coset.hpp:37-43): a[i] *= g^i
// ---------------------------------------------------------------------------------------------
struct PowTable {
    uint32_t p[32][8];  // g^(2^b), Montgomery form
};
__global__ void k_mul_powers(uint32_t* a, uint64_t n, PowTable tbl) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || i == 0) return;
    Fr acc = Fr::one();
    for (unsigned b = 0; b < 32 && (i >> b); ++b) {
        if ((i >> b) & 1) {
            Fr f;
#pragma unroll
            for (int k = 0; k < 8; ++k) f.l[k] = tbl.p[b][k];
            acc = fp_mul(acc, f);
        }
    }
    fp_store(a + 8 * i, fp_mul(fp_load<FrCfg>(a + 8 * i), acc));
}
// out[i] = k * g^i  (Montgomery form): coset tables for the prover
__global__ void k_fill_powers(uint32_t* out, uint64_t n, PowTable tbl, Fr k) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr acc = k;
    for (unsigned b = 0; b < 32 && (i >> b); ++b) {
        if ((i >> b) & 1) {
            Fr f;
#pragma unroll
            for (int q = 0; q < 8; ++q) f.l[q] = tbl.p[b][q];
            acc = fp_mul(acc, f);
        }
    }
    fp_store(out + 8 * i, acc);
}

static void fr4_to_limbs(uint32_t* dst, const hostfr::Fr4& x) { memcpy(dst, x.v, 32); }

static PowTable make_pow_table(const hostfr::Fr4& g) {
    PowTable t;
    hostfr::Fr4 p = g;
    for (int b = 0; b < 32; ++b) {
        fr4_to_limbs(t.p[b], hostfr::to_mont(p));
        p = hostfr::mul(p, p);
    }
    return t;
}

int multiply_by_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t n, const uint64_t g[4], uint32_t* d) {
    if (n <= 1) return VSP_OK;
    if (n > (uint64_t(1) << 32)) return set_error(ctx, VSP_E_INVALID, "multiply_by_coset: n too large");
    hostfr::Fr4 gg;
    memcpy(gg.v, g, 32);
    if (hostfr::geq(gg.v, hostfr::MOD)) return set_error(ctx, VSP_E_INVALID, "coset generator not reduced");
    PowTable t = make_pow_table(gg);
    VSP_LAUNCH_K(ctx, KC_NTT, k_mul_powers, (unsigned)((n + 127) / 128), 128, 0, s, d, n, t);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// table[i] = k * g^i, i < n (k, g canonical): Montgomery form, or canonical values with canonical_out
int fill_power_table(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_out, uint64_t n, const hostfr::Fr4& g,
                     const hostfr::Fr4& k, bool canonical_out) {
    PowTable t = make_pow_table(g);
    Fr kk;
    fr4_to_limbs(kk.l, canonical_out ? k : hostfr::to_mont(k));
    VSP_LAUNCH_K(ctx, KC_NTT, k_fill_powers, (unsigned)((n + 127) / 128), 128, 0, s, d_out, n, t, kk);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// divide_by_z_on_coset
// ---------------------------------------------------------------------------------------------
__global__ void k_scale_const(uint32_t* a, uint64_t n, Fr k) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, fp_mul(fp_load<FrCfg>(a + 8 * i), k));
}
__global__ void k_mul_table(uint32_t* a, uint64_t n, const uint32_t* tbl) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(a + 8 * i, fp_mul(fp_load<FrCfg>(a + 8 * i), fp_load<FrCfg>(tbl + 8 * i)));
}
// den[i] = A * elt^i - B for i < n (Montgomery), then batch-inverted in chunks of CH per thread (Montgomery's trick,
// one Fermat inversion per chunk) — replaces the per-element `.inversed()` of step_radix2_domain.hpp:232-235
constexpr int STEP_INV_CHUNK = 32;
__global__ void k_step_div_table(uint32_t* out, uint64_t n, PowTable elt_pows, Fr A, Fr B) {
    const uint64_t c = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t base = c * STEP_INV_CHUNK;
    if (base >= n) return;
    const int cnt = (int)((n - base < (uint64_t)STEP_INV_CHUNK) ? (n - base) : STEP_INV_CHUNK);
    Fr e = Fr::one();  // elt^base
    for (unsigned b = 0; b < 32 && (base >> b); ++b) {
        if ((base >> b) & 1) {
            Fr f;
#pragma unroll
            for (int q = 0; q < 8; ++q) f.l[q] = elt_pows.p[b][q];
            e = fp_mul(e, f);
        }
    }
    Fr step;
#pragma unroll
    for (int q = 0; q < 8; ++q) step.l[q] = elt_pows.p[0][q];
    Fr den[STEP_INV_CHUNK];
    Fr run = Fr::one();
    for (int i = 0; i < cnt; ++i) {
        den[i] = fp_sub(fp_mul(A, e), B);
        run = fp_mul(run, den[i]);
        fp_store(out + 8 * (base + i), run);  // prefix products, overwritten below
        e = fp_mul(e, step);
    }
    Fr inv = fp_inv_safegcd(run);
    for (int i = cnt - 1; i >= 0; --i) {
        Fr prev = (i > 0) ? fp_load<FrCfg>(out + 8 * (base + i - 1)) : Fr::one();
        fp_store(out + 8 * (base + i), fp_mul(inv, prev));
        inv = fp_mul(inv, den[i]);
    }
}

// out[i] = 1 / (A * elt^i - B), i < n, Montgomery form (A, B, elt canonical); a zero denominator poisons its chunk,
// so callers rule that case out first (the generator does: t outside the domain)
int inv_affine_table_dev(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_out, uint64_t n, const hostfr::Fr4& elt,
                         const hostfr::Fr4& A, const hostfr::Fr4& B) {
    if (n == 0) return VSP_OK;
    PowTable pt = make_pow_table(elt);
    Fr a, b;
    fr4_to_limbs(a.l, hostfr::to_mont(A));
    fr4_to_limbs(b.l, hostfr::to_mont(B));
    const uint64_t chunks = (n + STEP_INV_CHUNK - 1) / STEP_INV_CHUNK;
    VSP_LAUNCH_K(ctx, KC_NTT, k_step_div_table, (unsigned)((chunks + 63) / 64), 64, 0, s, d_out, n, pt, a, b);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

int divide_by_z_on_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, uint32_t* d) {
    using namespace hostfr;
    const Fr4 coset = from_u64(7);
    if (is_basic(m)) {
        Fr4 z = sub(pow_u64(coset, m), from_u64(1));
        Fr k;
        fr4_to_limbs(k.l, to_mont(inv(z)));
        VSP_LAUNCH_K(ctx, KC_NTT, k_scale_const, (unsigned)((m + 127) / 128), 128, 0, s, d, m, k);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        return VSP_OK;
    }
    if (!is_step(m)) return set_error(ctx, VSP_E_DOMAIN, "no executable radix-2 domain of size %llu", (unsigned long long)m);
    const unsigned log_2big = ceil_log2_u64(m);
    const uint64_t big_m = uint64_t(1) << (log_2big - 1), small_m = m - big_m;
    if (!ctx->ntt) ctx->ntt = new NttTables();
    StepDivTable*& tbl = ctx->ntt->step_div[m];
    if (!tbl) {
        tbl = new StepDivTable();
        const Fr4 omega = unity_root_pow2(log_2big);
        const Fr4 Z0 = sub(pow_u64(coset, big_m), from_u64(1));
        const Fr4 c_small_Z0 = mul(pow_u64(coset, small_m), Z0);
        const Fr4 w_small = pow_u64(omega, small_m);
        const Fr4 w_small_Z0 = mul(w_small, Z0);
        const Fr4 w_2small = pow_u64(omega, 2 * small_m);
        const Fr4 cw = mul(coset, omega);
        const Fr4 Z1 = mul(sub(pow_u64(cw, big_m), from_u64(1)), sub(pow_u64(cw, small_m), w_small));
        tbl->z1_inv_mont = to_mont(inv(Z1));
        int rc = tbl->inv_big.alloc(ctx, big_m * 32, s);
        if (rc) {
            delete tbl;
            tbl = nullptr;
            return rc;
        }
        PowTable pt = make_pow_table(w_2small);
        Fr A, B;
        fr4_to_limbs(A.l, to_mont(c_small_Z0));
        fr4_to_limbs(B.l, to_mont(w_small_Z0));
        const uint64_t chunks = (big_m + STEP_INV_CHUNK - 1) / STEP_INV_CHUNK;
        VSP_LAUNCH_K(ctx, KC_NTT, k_step_div_table, (unsigned)((chunks + 63) / 64), 64, 0, s, tbl->inv_big.as<uint32_t>(), big_m,
                   pt, A, B);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        // the table outlives this stream-ordered call: make it visible to every stream of the context
        VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    }
    VSP_LAUNCH_K(ctx, KC_NTT, k_mul_table, (unsigned)((big_m + 127) / 128), 128, 0, s, d, big_m, tbl->inv_big.as<uint32_t>());
    Fr k;
    fr4_to_limbs(k.l, tbl->z1_inv_mont);
    VSP_LAUNCH_K(ctx, KC_NTT, k_scale_const, (unsigned)((small_m + 127) / 128), 128, 0, s, d + 8 * big_m, small_m, k);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

}  // namespace vsp

// ---------------------------------------------------------------------------------------------
// vsp_fr_fft_noalloc_dev: the basic radix-2 transform for callers that may not allocate.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and the
// same pass plan as ntt_pow2 above, but the Stockham ping-pong buffer that ntt_pow2 takes from the stream-ordered
// pool, and the twiddle table that ntt_tables_ensure cudaMallocs, live in two file-scope __device__ arrays of fixed
// size: VSP_NOALLOC_FFT_MAX_LOG = 16, i.e. 2^16 x 32 B = 2 MiB of scratch + 2^15 x 32 B = 1 MiB of twiddles, held by
// every GPU that loads the library until the process exits.
//
// In host code the arrays' names are host-side stand-ins, not device addresses: the addresses the kernels get come
// from cudaGetSymbolAddress, called once per device with that device current (noalloc_setup).  Nothing in the scratch
// has to start at zero: pass 0 of a two-pass transform writes all m elements before pass 1 reads any, and a one-pass
// in-place transform writes all m before they are copied back, so the previous call's contents are never read.
// All calls on one GPU share the one scratch array, whatever context or stream they come from: calls that overlap on
// the GPU race.  Calls on the same stream are ordered and safe.
// ---------------------------------------------------------------------------------------------
#define VSP_NOALLOC_FFT_MAX_LOG 16
static __device__ __align__(256) uint32_t g_noalloc_fft_scratch[8u << VSP_NOALLOC_FFT_MAX_LOG];
static __device__ __align__(256) uint32_t g_noalloc_fft_twiddles[8u << (VSP_NOALLOC_FFT_MAX_LOG - 1)];

namespace vsp {

struct NoAllocFft {
    uint32_t* scratch = nullptr;   // device address of g_noalloc_fft_scratch on this device
    uint32_t* twiddles = nullptr;  // device address of g_noalloc_fft_twiddles on this device, filled
    bool ready = false;
};
static const int NOALLOC_MAX_DEVICES = 64;
static NoAllocFft g_noalloc[NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mutex;

// once per device, that device current: resolve the two symbols and fill the twiddle table
static int noalloc_setup(vsp_ctx* ctx, NoAllocFft** out) {
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return set_error(ctx, VSP_E_INVALID, "vsp_fr_fft_noalloc_dev: device ordinal %d out of range", ctx->device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    NoAllocFft* st = &g_noalloc[ctx->device];
    if (!st->ready) {
        uint32_t *scratch = nullptr, *tw = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&scratch, g_noalloc_fft_scratch);
        if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&tw, g_noalloc_fft_twiddles);
        if (e != cudaSuccess || !scratch || !tw)
            return set_error(ctx, VSP_E_CUDA, "vsp_fr_fft_noalloc_dev: cudaGetSymbolAddress failed on device %d: %s", ctx->device,
                             cudaGetErrorString(e));
        const uint64_t half = uint64_t(1) << (VSP_NOALLOC_FFT_MAX_LOG - 1);
        k_build_twiddles<<<(unsigned)((half + 255) / 256), 256, 0, ctx->stream>>>(tw, VSP_NOALLOC_FFT_MAX_LOG);
        ++ctx->launches;
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        // other contexts and streams of this device use the table from now on: the build must be complete
        VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        st->scratch = scratch;
        st->twiddles = tw;
        st->ready = true;
    }
    *out = st;
    return VSP_OK;
}

}  // namespace vsp

extern "C" int vsp_fr_fft_noalloc_dev(vsp_ctx* ctx, uint64_t m, int inverse, uint64_t* d_inout) {
    using namespace vsp;
    if (!ctx || !d_inout) return VSP_E_INVALID;
    if (m < 2 || (m & (m - 1)) != 0)
        return set_error(ctx, VSP_E_DOMAIN, "vsp_fr_fft_noalloc_dev: m = %llu is not a basic radix-2 domain size (a power of two >= 2)",
                         (unsigned long long)m);
    if (m > (uint64_t(1) << VSP_NOALLOC_FFT_MAX_LOG))
        return set_error(ctx, VSP_E_INVALID,
                         "vsp_fr_fft_noalloc_dev: m = %llu exceeds the static scratch, maximum 2^%d = %llu elements "
                         "(use vsp_fr_fft_dev for larger domains)",
                         (unsigned long long)m, VSP_NOALLOC_FFT_MAX_LOG, (unsigned long long)(uint64_t(1) << VSP_NOALLOC_FFT_MAX_LOG));
    VSP_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    NoAllocFft* st = nullptr;
    VSP_TRY(noalloc_setup(ctx, &st));
    const unsigned log_n = ceil_log2_u64(m);
    cudaStream_t s = ctx->stream;
    uint32_t* d = reinterpret_cast<uint32_t*>(d_inout);
    // the pass plan of ntt_pow2 for log_n <= 16: one pass up to 2^10, two above; 2^9 values per block
    unsigned ks[2], np;
    if (log_n <= 10) {
        np = 1;
        ks[0] = log_n;
    } else {
        np = 2;
        ks[0] = (log_n + 1) / 2;
        ks[1] = log_n - ks[0];
    }
    const unsigned block_log = 9;
    const uint32_t* src = d;
    unsigned ns_log = 0;
    for (unsigned p = 0; p < np; ++p) {
        const unsigned k = ks[p];
        unsigned g_log = (k >= block_log) ? 0 : (block_log - k);
        if (g_log > log_n - k) g_log = log_n - k;
        const bool last = (p + 1 == np);
        const bool single_block = (log_n - k - g_log) == 0;
        // a pass never runs in place across blocks: everything but a one-block transform and the second pass goes
        // through the static scratch
        uint32_t* dst = (last && (np == 2 || single_block)) ? d : st->scratch;
        PassArgs a;
        a.in = src;
        a.out = dst;
        a.w = st->twiddles;
        a.pre = nullptr;
        a.post = nullptr;
        a.log_w = VSP_NOALLOC_FFT_MAX_LOG;
        a.log_n = log_n;
        a.k = k;
        a.ns_log = ns_log;
        a.g_log = g_log;
        a.in_stride = m;
        a.out_stride = m;
        a.inverse = inverse ? 1 : 0;
        a.scale_log = (inverse && last) ? (int)log_n : 0;
        const unsigned blocks = 1u << (log_n - k - g_log);
        const size_t smem = ((size_t)32 << (k + g_log)) + ((size_t)16 << k);  // <= 48 KiB for every log_n <= 16
        if (k >= 2)
            k_ntt_pass4<<<dim3(blocks, 1), 1u << (k - 2 + g_log), smem, s>>>(a);
        else
            k_ntt_pass<<<dim3(blocks, 1), 1u << (k - 1 + g_log), smem, s>>>(a);
        ++ctx->launches;
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        if (last && dst != d) VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d, dst, m * 32, cudaMemcpyDeviceToDevice, s));
        src = dst;
        ns_log += k;
    }
    return VSP_OK;
}
