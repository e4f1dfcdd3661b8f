// Multi-GPU halo plumbing shared by lb_capi.cu and lb_halo.cu.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace lb {
// signal both neighbours with `value`, then wait until both local slots reached it
int launch_halo_sync(uint32_t* signal_up, uint32_t* signal_down, uint32_t* wait_slots,
                     uint32_t value, cudaStream_t st);
}  // namespace lb
