// explicit instantiation of the step kernels for D2Q9 f32, collision regularized
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D2Q9, float, 4)
}
