// msm_g1.cu — G1 (Fq) instantiation of the bucket multiexp (msm_impl.cuh)
#include "msm_impl.cuh"
namespace vsp {
VSP_MSM_INSTANTIATE(g1, FqOps)
}
