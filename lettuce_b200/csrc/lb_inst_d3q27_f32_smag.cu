// explicit instantiation of the step kernels for D3Q27 f32, collision smagorinsky
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D3Q27, float, 5)
}
