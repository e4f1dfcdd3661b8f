// explicit instantiation of the step / moments kernels for D3Q19 f32
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE(D3Q19, float)
}
