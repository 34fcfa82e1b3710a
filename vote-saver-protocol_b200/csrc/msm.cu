// msm.cu — curve-independent part of the multiexp layer: resident bases and G1 / G2 dispatch.
#include "msm.cuh"

namespace vsp {

int msm_run_g1(vsp_ctx*, cudaStream_t, const BasesDev&, uint64_t, const uint32_t*, uint64_t, unsigned, uint64_t*, uint64_t, bool);
int msm_run_g2(vsp_ctx*, cudaStream_t, const BasesDev&, uint64_t, const uint32_t*, uint64_t, unsigned, uint64_t*, uint64_t, bool);
int bases_finalize_g1(vsp_ctx*, cudaStream_t, BasesDev*, bool);
int bases_finalize_g2(vsp_ctx*, cudaStream_t, BasesDev*, bool);
int points_sum_g1(vsp_ctx*, cudaStream_t, uint64_t, const uint64_t*, uint64_t*);
int points_sum_g2(vsp_ctx*, cudaStream_t, uint64_t, const uint64_t*, uint64_t*);
int bases_to_mont_g1(vsp_ctx*, cudaStream_t, uint32_t*, uint64_t);
int bases_to_mont_g2(vsp_ctx*, cudaStream_t, uint32_t*, uint64_t);

int bases_finalize(vsp_ctx* ctx, BasesDev* b, bool table) {
    return b->g2 ? bases_finalize_g2(ctx, ctx->stream, b, table) : bases_finalize_g1(ctx, ctx->stream, b, table);
}

int bases_upload(vsp_ctx* ctx, bool g2, uint64_t n, const uint64_t* host_bases, BasesDev* out, bool table) {
    out->ctx = ctx;
    out->n = n;
    out->g2 = g2;
    const size_t bytes = (size_t)n * (g2 ? 192 : 96);
    VSP_CUDA_CHECK(ctx, cudaMalloc(&out->d_pts, bytes ? bytes : 16));
    if (n) {
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out->d_pts, host_bases, bytes, cudaMemcpyHostToDevice, ctx->stream));
        VSP_TRY(g2 ? bases_to_mont_g2(ctx, ctx->stream, out->d_pts, n) : bases_to_mont_g1(ctx, ctx->stream, out->d_pts, n));
    }
    VSP_TRY(bases_finalize(ctx, out, table));
    // host_bases may be pageable and reused by the caller; the table must be visible to every stream
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

void bases_free(BasesDev* b) {
    if (b->d_pts) cudaFree(b->d_pts);
    b->d_pts = nullptr;
}

int msm_run(vsp_ctx* ctx, cudaStream_t s, const BasesDev& bases, uint64_t n_scalars, const uint32_t* d_scalars,
            uint64_t* d_out, bool raw_xyzz, unsigned batch, uint64_t scalar_stride, uint64_t out_stride) {
    return bases.g2 ? msm_run_g2(ctx, s, bases, n_scalars, d_scalars, scalar_stride, batch, d_out, out_stride, raw_xyzz)
                    : msm_run_g1(ctx, s, bases, n_scalars, d_scalars, scalar_stride, batch, d_out, out_stride, raw_xyzz);
}

int points_sum_dev(vsp_ctx* ctx, cudaStream_t s, bool g2, uint64_t k, const uint64_t* d_points, uint64_t* d_out) {
    return g2 ? points_sum_g2(ctx, s, k, d_points, d_out) : points_sum_g1(ctx, s, k, d_points, d_out);
}

}  // namespace vsp
