// explicit instantiation of the step kernels for D2Q9 f64, collision BGK with a forcing scheme
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE_COLL(D2Q9, double, 6)
}
