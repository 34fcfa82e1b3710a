// msm_g1.cu — G1 (Fq) instantiation of the bucket multiexp (msm_impl.cuh), plus the synthetic-bases generator used
// by the measurement sweeps (bases = k_i * G for seeded 64-bit k_i, built on the device).
#include "msm_impl.cuh"
namespace vsp {
VSP_MSM_INSTANTIATE(g1, FqOps)

namespace {
struct G1Gen {
    uint64_t xy[12];
};
// BLS12-381 G1 generator, canonical little-endian limbs
constexpr G1Gen G1_GENERATOR = {{0xfb3af00adb22c6bbULL, 0x6c55e83ff97a1aefULL, 0xa14e3a3f171bac58ULL,
                                 0xc3688c4f9774b905ULL, 0x2695638c4fa9ac0fULL, 0x17f1d3a73197d794ULL,
                                 0x0caa232946c5e7e1ULL, 0xd03cc744a2888ae4ULL, 0x00db18cb2c04b3edULL,
                                 0xfcf5e095d5d00af6ULL, 0xa09e30ed741d8ae4ULL, 0x08b3f481e3aaa0f1ULL}};

__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
    uint64_t z = x + 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

// pts[i] = k_i * G (Montgomery affine), k_i = splitmix64(seed + i) | 1
__global__ void __launch_bounds__(128) k_g1_generate(uint32_t* pts, uint64_t n, uint64_t seed, G1Gen gen) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    G1Affine g = G1Affine::load(gen.xy);
    g.x = fp_to_mont(g.x);
    g.y = fp_to_mont(g.y);
    const uint64_t k = splitmix64(seed + i) | 1ULL;
    G1XYZZ acc = G1XYZZ::inf();
    for (int bit = 63 - __clzll((long long)k); bit >= 0; --bit) {
        acc = ec_dbl<FqOps>(acc);
        if ((k >> bit) & 1ULL) ec_madd<FqOps>(acc, g);
    }
    G1Affine a = ec_to_affine<FqOps>(acc);
    a.store(pts + i * (G1Affine::BYTES / 4));
}

__global__ void k_g1_from_mont(const uint32_t* pts, uint64_t first, uint64_t count, uint64_t* out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= count) return;
    G1Affine a = G1Affine::load(pts + (first + i) * (G1Affine::BYTES / 4));
    a.x = fp_from_mont(a.x);
    a.y = fp_from_mont(a.y);
    a.store(out + i * 12);
}
}  // namespace

int bases_generate_g1(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_pts, uint64_t n, uint64_t seed) {
    VSP_LAUNCH(ctx, k_g1_generate, (unsigned)((n + 127) / 128), 128, 0, s, d_pts, n, seed, G1_GENERATOR);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}
int bases_read_g1(vsp_ctx* ctx, cudaStream_t s, const uint32_t* d_pts, uint64_t first, uint64_t count, uint64_t* d_out) {
    VSP_LAUNCH(ctx, k_g1_from_mont, (unsigned)((count + 127) / 128), 128, 0, s, d_pts, first, count, d_out);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}
}  // namespace vsp
