// msm.cuh — internal interface of the bucket multiexp (msm.cu), shared with the prover.
#pragma once
#include "common.cuh"

namespace vsp {

// Device-resident bases of one multiexp (a proving-key query): affine points in Montgomery form.
struct BasesDev {
    vsp_ctx* ctx = nullptr;
    uint64_t n = 0;
    bool g2 = false;
    uint32_t* d_pts = nullptr;  // n * (24 | 48) limbs
};

int bases_upload(vsp_ctx* ctx, bool g2, uint64_t n, const uint64_t* host_bases, BasesDev* out);
void bases_free(BasesDev* b);

// sum_i scalars[i] * bases[i], i < n_scalars <= bases.n.  d_scalars: canonical Fr on the device.
// d_out: canonical affine point on the device (12 / 24 u64 limbs), or with raw_xyzz the Montgomery-form XYZZ
// accumulator (24 / 48 u64 limbs; infinity <=> ZZ == 0).  Entirely stream-ordered on `s`.
int msm_run(vsp_ctx* ctx, cudaStream_t s, const BasesDev& bases, uint64_t n_scalars, const uint32_t* d_scalars,
            uint64_t* d_out, bool raw_xyzz = false);

// out = sum of k affine canonical points (device pointers)
int points_sum_dev(vsp_ctx* ctx, cudaStream_t s, bool g2, uint64_t k, const uint64_t* d_points, uint64_t* d_out);

}  // namespace vsp
