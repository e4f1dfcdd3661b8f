// explicit instantiation of the step kernels for D3Q19 f32, collision BGK with a forcing scheme
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D3Q19, float, 6)
}
