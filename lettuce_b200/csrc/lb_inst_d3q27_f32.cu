// explicit instantiation of the step / moments kernels for D3Q27 f32
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE(D3Q27, float)
}
