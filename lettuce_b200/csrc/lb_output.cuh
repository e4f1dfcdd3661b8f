// Field extraction for the output side: pressure and velocity in PHYSICAL units of a
// state f, written in the point order of a VTK file (lattice x fastest) -- what
// `VTKReporter.__call__` (reference/lettuce/ext/_reporter/vtk_reporter.py:29-47) obtains from
// `flow.p_pu`, `flow.u_pu` (lettuce/_flow.py:136-172, 183-190; _unit.py:52-53, 82-83) followed by pyevtk's
// Fortran-order packing.  One pass: every population is read once (coalesced along the contiguous lattice
// axis), the 1 + d fields of a 32 x 32 tile are transposed through shared memory and stored coalesced along
// the file's fastest axis.  Algorithmic traffic: (Q + 1 + d) values per node.
#pragma once
#include "lb_physics.cuh"

namespace lb {

constexpr int OUT_TILE = 32;   // tile edge (nodes), blockDim = (32, 8)
constexpr int OUT_ROWS = 8;

template <typename T>
struct OutputParams {
  const T* f;      // [Q][n_nodes], kernel layout (Z contiguous)
  T* out;          // [1 + D][n_nodes]: p, then u in lattice axis order; node index = (z * nm + m) * na + a
  unsigned nz;     // contiguous axis of f = SLOWEST axis of the file
  unsigned na;     // fastest axis of the file: lattice x (3-D: kernel X, 2-D: kernel Y)
  unsigned nm;     // the axis in between (3-D: kernel Y; 2-D: 1)
  unsigned n_nodes;
  T rho0;          // p = (rho - rho0) * p_scale,  p_scale = cs^2 * p_char_pu / p_char_lu
  T p_scale;
  T u_scale;       // u_pu = u * u_scale,          u_scale = u_char_pu / u_char_lu
};

template <class S, typename T>
__global__ void __launch_bounds__(OUT_TILE * OUT_ROWS)
output_fields_kernel(const __grid_constant__ OutputParams<T> p) {
  constexpr int NF = 1 + S::D;
  __shared__ T tile[NF][OUT_TILE][OUT_TILE + 1];
  const unsigned tx = threadIdx.x, ty = threadIdx.y;
  const unsigned z0 = blockIdx.x * OUT_TILE, a0 = blockIdx.y * OUT_TILE, m = blockIdx.z;
  const size_t N = p.n_nodes;
  // in: node = (a * nm + m) * nz + z  (a slowest, z contiguous)
  {
    const unsigned z = z0 + tx;
#pragma unroll
    for (int j = 0; j < OUT_TILE / OUT_ROWS; ++j) {
      const unsigned al = ty + OUT_ROWS * j, a = a0 + al;
      if (z < p.nz && a < p.na) {
        const size_t node = (size_t(a) * p.nm + m) * p.nz + z;
        T g[S::Q];
        static_for<S::Q>([&](auto qc) {
          constexpr int q = decltype(qc)::value;
          g[q] = p.f[q * N + node];
        });
        T rho, u[3];
        moments<S, T>(g, rho, u);
        tile[0][al][tx] = (rho - p.rho0) * p.p_scale;
#pragma unroll
        for (int c = 0; c < S::D; ++c) tile[1 + c][al][tx] = u[S::axis(c)] * p.u_scale;
      }
    }
  }
  __syncthreads();
  // out: index = (z * nm + m) * na + a  (a contiguous)
  {
    const unsigned a = a0 + tx;
#pragma unroll
    for (int j = 0; j < OUT_TILE / OUT_ROWS; ++j) {
      const unsigned zl = ty + OUT_ROWS * j, z = z0 + zl;
      if (z < p.nz && a < p.na) {
        const size_t o = (size_t(z) * p.nm + m) * p.na + a;
#pragma unroll
        for (int c = 0; c < NF; ++c) p.out[c * N + o] = tile[c][tx][zl];
      }
    }
  }
}

template <class S, typename T>
cudaError_t launch_output_fields(const OutputParams<T>& p, cudaStream_t st) {
  const dim3 block(OUT_TILE, OUT_ROWS);
  // the middle axis rides in gridDim.z (<= 65535); larger lattices are walked in chunks
  for (unsigned m0 = 0; m0 < p.nm; m0 += 65535u) {
    OutputParams<T> q = p;
    const unsigned cnt = (p.nm - m0 < 65535u) ? p.nm - m0 : 65535u;
    // shift the views so that blockIdx.z = 0 is plane m0: in-node += m0 * nz, out-index += m0 * na
    q.f = p.f + size_t(m0) * p.nz;
    q.out = p.out + size_t(m0) * p.na;
    const dim3 grid((p.nz + OUT_TILE - 1) / OUT_TILE, (p.na + OUT_TILE - 1) / OUT_TILE, cnt);
    output_fields_kernel<S, T><<<grid, block, 0, st>>>(q);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace lb
