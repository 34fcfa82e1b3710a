// msm.cuh — internal interface of the bucket multiexp (msm_impl.cuh), shared with the prover.
#pragma once
#include "common.cuh"

namespace vsp {

// Device-resident bases of one multiexp (a proving-key query): affine points in Montgomery form.  With `table`,
// d_pts holds `windows` levels of n points: level w = 2^(c w) * P_i (built once by bases_finalize).
struct BasesDev {
    vsp_ctx* ctx = nullptr;
    uint64_t n = 0;
    bool g2 = false;
    bool table = false;
    unsigned c = 0, windows = 0;
    unsigned force_c = 0;  // non-zero: fixed window width (the 8-bit fixed-base tables of the prover)
    uint32_t* d_pts = nullptr;  // n * (table ? windows : 1) * (24 | 48) limbs
};

// upload canonical affine points, convert to Montgomery form and (table = true) build the window table
int bases_upload(vsp_ctx* ctx, bool g2, uint64_t n, const uint64_t* host_bases, BasesDev* out, bool table);
// the same for points already on the device in Montgomery form (d_pts set, n set)
int bases_finalize(vsp_ctx* ctx, BasesDev* b, bool table);
void bases_free(BasesDev* b);

// out[b] = sum_i scalars[b][i] * bases[i], i < n_scalars <= bases.n, for b < batch ballots.
// d_scalars: canonical Fr on the device, ballot b starting at element b * scalar_stride.
// d_out + b * out_stride (u64 limbs): canonical affine point (12 / 24 limbs), or with raw_xyzz the Montgomery-form
// XYZZ accumulator (24 / 48 limbs; infinity <=> ZZ == 0).  Entirely stream-ordered on `s`.
int msm_run(vsp_ctx* ctx, cudaStream_t s, const BasesDev& bases, uint64_t n_scalars, const uint32_t* d_scalars,
            uint64_t* d_out, bool raw_xyzz = false, unsigned batch = 1, uint64_t scalar_stride = 0,
            uint64_t out_stride = 0);

// 32x32->64 multiply-adds k_msm_accumulate<G1> issues per bucket addition: the XYZZ mixed addition is 6 products
// (2 * 12^2 + 12 = 300 each), 2 squares (12 * 11 / 2 + 12 + 12^2 + 12 = 234 each) and Y3 = R (Q - X3) - Y1 PPP as two
// double-width products on one reduction (2 * 144 + 156 = 444).  bench.py's roofline quotes it beside the reference
// algorithm's 11 x 300.
constexpr double ACC_G1_MADS_PER_ADD = 6 * 300.0 + 2 * 234.0 + 444.0;

// out = sum of k affine canonical points (device pointers)
int points_sum_dev(vsp_ctx* ctx, cudaStream_t s, bool g2, uint64_t k, const uint64_t* d_points, uint64_t* d_out);

}  // namespace vsp
