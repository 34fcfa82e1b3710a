// msm_g2.cu — G2 (Fq2) instantiation of the bucket multiexp (msm_impl.cuh), plus the synthetic G2 point generator
// used to build shape-only proving keys for the measurement workloads.
#include "msm_impl.cuh"
namespace vsp {
VSP_MSM_INSTANTIATE(g2, Fq2Ops)

namespace {
struct G2Gen {
    uint64_t xy[24];
};
// BLS12-381 G2 generator, canonical little-endian limbs: x.c0 | x.c1 | y.c0 | y.c1
constexpr G2Gen G2_GENERATOR = {{
    0xd48056c8c121bdb8ULL, 0x0bac0326a805bbefULL, 0xb4510b647ae3d177ULL, 0xc6e47ad4fa403b02ULL, 0x260805272dc51051ULL, 0x024aa2b2f08f0a91ULL,
    0xe5ac7d055d042b7eULL, 0x334cf11213945d57ULL, 0xb5da61bbdc7f5049ULL, 0x596bd0d09920b61aULL, 0x7dacd3a088274f65ULL, 0x13e02b6052719f60ULL,
    0xe193548608b82801ULL, 0x923ac9cc3baca289ULL, 0x6d429a695160d12cULL, 0xadfd9baa8cbdd3a7ULL, 0x8cc9cdc6da2e351aULL, 0x0ce5d527727d6e11ULL,
    0xaaa9075ff05f79beULL, 0x3f370d275cec1da1ULL, 0x267492ab572e99abULL, 0xcb3e287e85a763afULL, 0x32acd2b02bc28b99ULL, 0x0606c4a02ea734ccULL}};

__device__ inline uint64_t splitmix64_g2(uint64_t x) {
    uint64_t z = x + 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

// out[i] = k_i * G2 (canonical affine), k_i = splitmix64(seed + i) | 1
__global__ void __launch_bounds__(64) k_g2_generate(uint64_t* out, uint64_t n, uint64_t seed, G2Gen gen) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    G2Affine g = G2Affine::load(gen.xy);
    g.x = Fq2Ops::to_mont(g.x);
    g.y = Fq2Ops::to_mont(g.y);
    const uint64_t k = splitmix64_g2(seed + i) | 1ULL;
    G2XYZZ acc = G2XYZZ::inf();
    for (int bit = 63 - __clzll((long long)k); bit >= 0; --bit) {
        acc = ec_dbl<Fq2Ops>(acc);
        if ((k >> bit) & 1ULL) ec_madd<Fq2Ops>(acc, g);
    }
    store_canonical_affine<Fq2Ops>(out + i * 24, acc);
}
}  // namespace

int points_generate_g2(vsp_ctx* ctx, cudaStream_t s, uint64_t* d_out, uint64_t n, uint64_t seed) {
    VSP_LAUNCH(ctx, k_g2_generate, (unsigned)((n + 63) / 64), 64, 0, s, d_out, n, seed, G2_GENERATOR);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}
}  // namespace vsp
