// explicit instantiation of the step / moments kernels for D2Q9 f64
#include "lb_kernels.cuh"
namespace lb {
LB_INSTANTIATE(D2Q9, double)
}
