// sass_fp.cu — one fp_mul<FqCfg> / fp_sqr<FqCfg> / fp_mul<FrCfg> / fp_sqr<FrCfg> instance each, for counting the SASS
// instruction mix (tools/sass_fp.sh -> profiles/sass_fp_mul.txt).  Not part of the library.
#include "../vote-saver-protocol_b200/csrc/fp.cuh"
using namespace vsp;
extern "C" __global__ void sass_fq_mul(const uint32_t* a, const uint32_t* b, uint32_t* o) {
    fp_store(o, fp_mul(fp_load<FqCfg>(a), fp_load<FqCfg>(b)));
}
extern "C" __global__ void sass_fq_sqr(const uint32_t* a, uint32_t* o) { fp_store(o, fp_sqr(fp_load<FqCfg>(a))); }
extern "C" __global__ void sass_fr_mul(const uint32_t* a, const uint32_t* b, uint32_t* o) {
    fp_store(o, fp_mul(fp_load<FrCfg>(a), fp_load<FrCfg>(b)));
}
extern "C" __global__ void sass_fr_sqr(const uint32_t* a, uint32_t* o) { fp_store(o, fp_sqr(fp_load<FrCfg>(a))); }
