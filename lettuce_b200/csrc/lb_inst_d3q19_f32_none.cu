// explicit instantiation of the step kernels for D3Q19 f32, collision none
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D3Q19, float, 0)
LB_INSTANTIATE_MOMENTS(D3Q19, float)
}
