// explicit instantiation of the step / moments kernels for D2Q9 f32
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE(D2Q9, float)
}
