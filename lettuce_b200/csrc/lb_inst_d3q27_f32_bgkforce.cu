// explicit instantiation of the step kernels for D3Q27 f32, collision BGK with a forcing scheme
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D3Q27, float, 6)
}
