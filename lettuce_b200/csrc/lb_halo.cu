// Neighbour handshake over NVLink peer memory, packed-halo scatter, CUDA IPC helpers.
//
// Push mode (see lb_halo in include/lettuce_b200.h): the step kernel packs the populations that
// cross its z faces; halo_push_kernel copies them into the neighbours' receive buffers through
// peer-mapped pointers and then performs the handshake:
//   "I finished step t"  ->  st.release.sys into both neighbours' wait slots,
//   spin (ld.acquire.sys) until both neighbours said the same.
// After it, (a) the neighbours' stores into our planes are visible, so step t+1 may read them,
// and (b) the neighbours have finished reading the buffer step t+1 will overwrite remotely.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/lettuce_b200.h"
#include "lb_halo.cuh"
#include "lb_kernels.cuh"

namespace lb {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// counters only grow; compare with wrap-around safe signed difference
__device__ __forceinline__ bool reached(uint32_t have, uint32_t want) { return int32_t(have - want) >= 0; }

constexpr int PUSH_BLOCK = 256;
constexpr int PUSH_GRID = 32;

// One launch per time step: (1) all blocks copy the packed halos into the neighbours' receive
// buffers with 16-byte stores (coalesced NVLink writes), (2) the last block to finish signals
// both neighbours and waits for theirs.  wait_slots[3] is the block counter.
__global__ void __launch_bounds__(PUSH_BLOCK)
halo_push_kernel(const uint4* __restrict__ src_hi, uint4* __restrict__ dst_hi,
                 const uint4* __restrict__ src_lo, uint4* __restrict__ dst_lo, size_t n_vec, size_t tail_bytes,
                 uint32_t* signal_up, uint32_t* signal_down, uint32_t* wait_slots, uint32_t value,
                 long long timeout_cycles) {
  // Programmatic dependent launch, as in the step kernel: this grid is scheduled while the step kernel
  // in front of it drains and waits here until that kernel has completed and its stores are visible;
  // the step kernel behind it is scheduled likewise and waits for THIS grid (its prerequisite).
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const size_t stride = size_t(gridDim.x) * PUSH_BLOCK;
  if (src_hi != nullptr)
    for (size_t i = size_t(blockIdx.x) * PUSH_BLOCK + threadIdx.x; i < n_vec; i += stride) dst_hi[i] = src_hi[i];
  if (src_lo != nullptr)
    for (size_t i = size_t(blockIdx.x) * PUSH_BLOCK + threadIdx.x; i < n_vec; i += stride) dst_lo[i] = src_lo[i];
  if (tail_bytes != 0 && blockIdx.x == 0 && threadIdx.x < tail_bytes / 4) {  // < 16 trailing bytes
    if (src_hi) reinterpret_cast<uint32_t*>(dst_hi + n_vec)[threadIdx.x] = reinterpret_cast<const uint32_t*>(src_hi + n_vec)[threadIdx.x];
    if (src_lo) reinterpret_cast<uint32_t*>(dst_lo + n_vec)[threadIdx.x] = reinterpret_cast<const uint32_t*>(src_lo + n_vec)[threadIdx.x];
  }
  if (signal_up == nullptr && signal_down == nullptr) return;
  __shared__ bool is_last;
  __threadfence_system();   // this thread's peer stores before the block's arrival
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(&wait_slots[3], 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  wait_slots[3] = 0u;
  __threadfence_system();   // every block's stores (observed through the counter) before the signal
  if (signal_up) st_release_sys(signal_up, value);
  if (signal_down) st_release_sys(signal_down, value);
  const long long t0 = clock64();
  while (!(reached(ld_acquire_sys(&wait_slots[0]), value) && reached(ld_acquire_sys(&wait_slots[1]), value))) {
    if (clock64() - t0 > timeout_cycles) {  // never hang the GPU: flag the error and carry on
      wait_slots[2] = 1u;
      break;
    }
    __nanosleep(64);
  }
}

int launch_halo_push(const void* src_hi, void* dst_hi, const void* src_lo, void* dst_lo, size_t bytes,
                     uint32_t* signal_up, uint32_t* signal_down, uint32_t* wait_slots, uint32_t value,
                     cudaStream_t st) {
  const size_t n_vec = bytes / 16, tail = bytes % 16;
  const bool copy = bytes != 0 && (src_hi != nullptr || src_lo != nullptr);
  const int grid = copy ? PUSH_GRID : 1;
  // ~10 s at 2 GHz: far beyond any legitimate skew between neighbours inside one node
  static const bool no_pdl = getenv("LB_NO_PDL") != nullptr;   // A/B timing only
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(PUSH_BLOCK);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  const cudaError_t e = cudaLaunchKernelEx(
      &cfg, halo_push_kernel, copy ? static_cast<const uint4*>(src_hi) : nullptr, static_cast<uint4*>(dst_hi),
      copy ? static_cast<const uint4*>(src_lo) : nullptr, static_cast<uint4*>(dst_lo), n_vec, tail,
      signal_up, signal_down, wait_slots, value, 20000000000ll);
  return e == cudaSuccess ? 0 : -1;
}

// packed [zface][lines] -> f[q][line * nz + z_edge].  sbits: the receiver's packed no-streaming bits -- an element
// whose population the node keeps is dropped (destination-masked streaming, _simulation.py:171-174).  LL: the
// buffers hold tagged words; every element is polled until its tag is `tag`.
template <class S, typename T, bool LL>
__global__ void unpack_halo_kernel(T* __restrict__ f, const void* __restrict__ from_down,
                                   const void* __restrict__ from_up, const uint32_t* __restrict__ sbits,
                                   unsigned lines, unsigned nz, size_t n_nodes, uint32_t tag, uint32_t* err) {
  const unsigned line = blockIdx.x * blockDim.x + threadIdx.x;
  if (line >= lines) return;
  constexpr int NF = zface_count<S>();
  T lo[NF], hi[NF];
  if constexpr (LL) {
    if (from_down) ll_gather<NF, T>(from_down, line, lines, tag, err, lo);
    if (from_up) ll_gather<NF, T>(from_up, line, lines, tag, err, hi);
  } else {
#pragma unroll
    for (int k = 0; k < NF; ++k) {
      if (from_down) lo[k] = static_cast<const T*>(from_down)[size_t(k) * lines + line];
      if (from_up) hi[k] = static_cast<const T*>(from_up)[size_t(k) * lines + line];
    }
  }
  const uint32_t b_lo = sbits ? sbits[size_t(line) * nz] : 0u;
  const uint32_t b_hi = sbits ? sbits[size_t(line) * nz + (nz - 1)] : 0u;
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    constexpr int ez = S::e(q, 2);
    if constexpr (ez == 1) {
      if (from_down && !((b_lo >> q) & 1u)) f[q * n_nodes + size_t(line) * nz] = lo[zface_index<S>(q)];
    } else if constexpr (ez == -1) {
      if (from_up && !((b_hi >> q) & 1u)) f[q * n_nodes + size_t(line) * nz + (nz - 1)] = hi[zface_index<S>(q)];
    }
  });
}

template <class S, typename T>
int unpack_typed(void* f, const void* d, const void* u, const uint32_t* sbits, unsigned lines, unsigned nz, size_t n,
                 uint32_t tag, uint32_t* err, cudaStream_t st) {
  if (tag != 0u)
    unpack_halo_kernel<S, T, true><<<(lines + 127) / 128, 128, 0, st>>>(static_cast<T*>(f), d, u, sbits, lines, nz, n, tag, err);
  else
    unpack_halo_kernel<S, T, false><<<(lines + 255) / 256, 256, 0, st>>>(static_cast<T*>(f), d, u, sbits, lines, nz, n, 0u, nullptr);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace lb

extern "C" {

int lb_zface_count(int stencil) {
  switch (stencil) {
    case LB_D2Q9: return lb::zface_count<lb::D2Q9>();
    case LB_D3Q19: return lb::zface_count<lb::D3Q19>();
    case LB_D3Q27: return lb::zface_count<lb::D3Q27>();
  }
  return -1;
}

int lb_unpack_halo_ex(const lb_desc* d, void* f, const void* from_down, const void* from_up, const uint32_t* sbits,
                      uint32_t ll_tag, uint32_t* ll_err, void* stream) {
  if (!d || !f) return LB_EINVAL;
  if (ll_tag != 0u && !ll_err) return LB_EINVAL;
  const bool two_d = d->stencil == LB_D2Q9;
  const unsigned nz = unsigned(two_d ? d->dims[1] : d->dims[2]);
  const unsigned lines = unsigned(two_d ? d->dims[0] : d->dims[0] * d->dims[1]);
  const size_t n = size_t(lines) * nz;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = -1;
  if (d->dtype == LB_F32) {
    if (d->stencil == LB_D2Q9) rc = lb::unpack_typed<lb::D2Q9, float>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
    if (d->stencil == LB_D3Q19) rc = lb::unpack_typed<lb::D3Q19, float>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
    if (d->stencil == LB_D3Q27) rc = lb::unpack_typed<lb::D3Q27, float>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
  } else if (d->dtype == LB_F64) {
    if (d->stencil == LB_D2Q9) rc = lb::unpack_typed<lb::D2Q9, double>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
    if (d->stencil == LB_D3Q19) rc = lb::unpack_typed<lb::D3Q19, double>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
    if (d->stencil == LB_D3Q27) rc = lb::unpack_typed<lb::D3Q27, double>(f, from_down, from_up, sbits, lines, nz, n, ll_tag, ll_err, st);
  }
  return rc == 0 ? LB_OK : LB_ECUDA;
}

int lb_unpack_halo(const lb_desc* d, void* f, const void* from_down, const void* from_up, void* stream) {
  return lb_unpack_halo_ex(d, f, from_down, from_up, nullptr, 0u, nullptr, stream);
}

size_t lb_halo_ll_bytes(const lb_desc* d) {
  if (!d) return 0;
  const int nf = lb_zface_count(d->stencil);
  if (nf < 0) return 0;
  const bool two_d = d->stencil == LB_D2Q9;
  const size_t lines = size_t(two_d ? d->dims[0] : d->dims[0] * d->dims[1]);
  return size_t(nf) * lines * (d->dtype == LB_F64 ? 16 : 8);
}

int lb_ipc_alloc(size_t bytes, void** dptr) {
  if (!dptr) return LB_EINVAL;
  if (cudaMalloc(dptr, bytes) != cudaSuccess) return LB_ECUDA;
  if (cudaMemset(*dptr, 0, bytes) != cudaSuccess) return LB_ECUDA;
  return LB_OK;
}
int lb_ipc_free(void* dptr) { return cudaFree(dptr) == cudaSuccess ? LB_OK : LB_ECUDA; }
int lb_ipc_export(void* dptr, unsigned char handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, dptr) != cudaSuccess) return LB_ECUDA;
  memcpy(handle, &h, 64);
  return LB_OK;
}
int lb_ipc_open(const unsigned char handle[64], void** dptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  return cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess ? LB_OK : LB_ECUDA;
}
int lb_ipc_close(void* dptr) { return cudaIpcCloseMemHandle(dptr) == cudaSuccess ? LB_OK : LB_ECUDA; }

}  // extern "C"
