// msm_g2.cu — G2 (Fq2) instantiation of the bucket multiexp (msm_impl.cuh)
#include "msm_impl.cuh"
namespace vsp {
VSP_MSM_INSTANTIATE(g2, Fq2Ops)
}
