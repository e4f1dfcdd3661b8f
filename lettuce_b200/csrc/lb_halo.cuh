// Multi-GPU halo plumbing shared by lb_capi.cu and lb_halo.cu.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

namespace lb {
// Copy `bytes` from src_hi -> dst_hi and src_lo -> dst_lo (dst_* peer-mapped, 16-byte vector
// stores; either pair may be null), then signal both neighbours with `value` and wait until
// both local slots reached it (signals may be null: copy only).
int launch_halo_push(const void* src_hi, void* dst_hi, const void* src_lo, void* dst_lo, size_t bytes,
                     uint32_t* signal_up, uint32_t* signal_down, uint32_t* wait_slots, uint32_t value,
                     cudaStream_t st);
}  // namespace lb
