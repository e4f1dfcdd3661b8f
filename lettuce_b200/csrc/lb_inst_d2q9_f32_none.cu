// explicit instantiation of the step kernels for D2Q9 f32, collision none
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D2Q9, float, 0)
LB_INSTANTIATE_MOMENTS(D2Q9, float)
}
