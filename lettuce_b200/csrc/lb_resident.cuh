// Small 2-D lattices: the state of a whole call stays ON CHIP.
//
// A lattice of a few 10^4 nodes (BASELINE config 1: 256^2 D2Q9 fp64, 4.7 MB) is not HBM-bound at all: with one
// launch -- or one grid barrier -- per time step, a step costs the barrier plus an L2 round trip of every
// population (step_loop_kernel: 3.0 us per step at 256^2 whatever the arithmetic).  Here a CTA keeps a TILE of the
// lattice in shared memory and advances it h time steps between two global exchanges ("temporal blocking"): it loads
// its tile plus a halo of h nodes on every side, steps the region h times -- after step k the outermost k rings are
// invalid (their neighbours were not loaded) and are not used any more --, writes the tile back and meets the other
// CTAs at ONE grid barrier per h steps.  Inside the region a step is the same push as in the big kernels: node x
// reads its Q populations at x, collides (the same collide_node, so the same bits), and writes population q to
// x + e_q of the other shared-memory buffer.  Halo nodes are computed redundantly by the neighbouring CTAs, with
// the same inputs and the same arithmetic: every copy of a node carries the same bits.
#pragma once
#include <cooperative_groups.h>

namespace lb {

#ifndef LB_RES_BLOCK
#define LB_RES_BLOCK 512
#endif
constexpr int RES_BLOCK = LB_RES_BLOCK;

template <typename T>
struct ResidentParams {
  T* buf[2];           // global ping-pong: round r reads buf[r & 1] and writes buf[(r + 1) & 1]
  unsigned ny, nz;     // the lattice in kernel axes (Z contiguous)
  unsigned n_nodes;
  unsigned ty, tz;     // tile of a CTA; ty | ny, tz | nz; grid = (ny / ty) * (nz / tz) CTAs, all co-resident
  unsigned plane_max;  // elements per population of the largest padded region: (ty + 2 h_max + 2) * (tz + 2 h_max + 2)
  int rounds;          // global exchanges of the call
  int steps_base;      // round r advances steps_base + (r < steps_extra ? 1 : 0) time steps
  int steps_extra;
  CollisionParams<T> cp;
};

template <class S, typename T, int COLL>
__global__ void __launch_bounds__(RES_BLOCK, 1) step_resident_kernel(const __grid_constant__ ResidentParams<T> p) {
  static_assert(S::D == 2, "tile-resident stepping is for the 2-D stencil");
  constexpr int Q = S::Q;
  extern __shared__ __align__(16) unsigned char res_smem[];
  T* A = reinterpret_cast<T*>(res_smem);
  T* B = A + size_t(Q) * p.plane_max;
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const unsigned tiles_z = p.nz / p.tz;
  const int ty0 = int((blockIdx.x / tiles_z) * p.ty), tz0 = int((blockIdx.x % tiles_z) * p.tz);
  const size_t N = p.n_nodes;
  int h_prev = -1;
  for (int r = 0; r < p.rounds; ++r) {
    const int h = p.steps_base + (r < p.steps_extra ? 1 : 0);   // time steps of this round = width of the halo
    // The buffers carry one ring of padding around the region: a push from the region's edge lands in the padding
    // instead of needing a bounds test (18 compare / predicate instructions per node update).  PZ = pitch of a row,
    // at(ry, rz) = element of region node (ry, rz).
    const unsigned RY = p.ty + 2u * unsigned(h), RZ = p.tz + 2u * unsigned(h), PZ = RZ + 2u, plane = (RY + 2u) * PZ;
    auto at = [&](unsigned ry, unsigned rz) { return (ry + 1u) * PZ + rz + 1u; };
    const T* __restrict__ src = p.buf[r & 1];
    T* __restrict__ dst = p.buf[(r + 1) & 1];
    // ---- tile + halo: global (L2) -> shared, periodic wrap.  ld.cg: the lines were written by other CTAs.
    // From the second round on the tile itself is still in shared memory (this CTA computed it) at the same place of
    // the region as long as the halo width does not change: only the FRAME around it -- 2 h RZ + 2 h ty nodes, the
    // neighbours' results -- is loaded, at most one node per thread, i.e. one L2 round trip for the whole load.
    const bool frame_only = r > 0 && h == h_prev;
    h_prev = h;
    const unsigned uh = unsigned(h), n_band = 2u * uh * RZ, n_load = frame_only ? n_band + 2u * uh * p.ty : RY * RZ;
    for (unsigned j = threadIdx.x; j < n_load; j += RES_BLOCK) {
      unsigned ry, rz;
      if (!frame_only) {
        ry = j / RZ; rz = j - ry * RZ;
      } else if (j < n_band) {                 // the h rows above and the h rows below the tile
        ry = j / RZ; rz = j - ry * RZ;
        if (ry >= uh) ry += p.ty;
      } else {                                 // the h columns left and right of the tile
        const unsigned m = j - n_band, row = m / (2u * uh), c = m - row * (2u * uh);
        ry = uh + row;
        rz = c < uh ? c : c + p.tz;
      }
      int gy = (ty0 - h + int(ry)) % int(p.ny), gz = (tz0 - h + int(rz)) % int(p.nz);
      if (gy < 0) gy += int(p.ny);
      if (gz < 0) gz += int(p.nz);
      const unsigned g = unsigned(gy) * p.nz + unsigned(gz);
      const unsigned i = at(ry, rz);
      T v[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) v[q] = __ldcg(src + size_t(q) * N + g);
#pragma unroll
      for (int q = 0; q < Q; ++q) A[q * plane + i] = v[q];
    }
    __syncthreads();
    // ---- h time steps in shared memory: step k updates the region shrunk by k - 1 rings.
    // A thread takes nodes i = tid, tid + RES_BLOCK, ... of the cy x cz sub-region; (y, z) advance incrementally --
    // one division per thread and step.  (Two nodes per trip for more independent arithmetic in flight was measured
    // no faster at 512 threads: the step is bound by instruction issue and the barrier skew, not by one chain.)
    for (int k = 1; k <= h; ++k) {
      const unsigned o = unsigned(k - 1), cy = RY - 2u * o, cz = RZ - 2u * o, n_sub = cy * cz;
      const unsigned qy = unsigned(RES_BLOCK) / cz, qz = unsigned(RES_BLOCK) - qy * cz;
      unsigned y = threadIdx.x / cz, z = threadIdx.x - y * cz;
      auto advance = [&]() { y += qy; z += qz; if (z >= cz) { z -= cz; ++y; } };
      auto load = [&](T (&g)[Q], unsigned idx) {
        static_for<Q>([&](auto qc) { g[decltype(qc)::value] = A[decltype(qc)::value * plane + idx]; });
      };
      auto collide = [&](T (&g)[Q]) {
        if constexpr (COLL != COLL_NONE) {
          T a[N_SLOTS], b[N_SLOTS], rho = T(0), u[3] = {T(0), T(0), T(0)};
          pair_split<S, T>(g, a, b);
          moments_from_pairs<S, T>(g[0], a, b, rho, u);
          collide_node<S, T, COLL, T>(g, a, b, rho, u, p.cp);
        }
      };
      auto push = [&](const T (&g)[Q], unsigned ry, unsigned rz) {
        static_for<Q>([&](auto qc) {
          constexpr int q = decltype(qc)::value;
          B[q * plane + at(ry + unsigned(S::e(q, 1)), rz + unsigned(S::e(q, 2)))] = g[q];   // edge pushes: into the padding
        });
      };
      for (unsigned i = threadIdx.x; i < n_sub; i += RES_BLOCK) {
        const unsigned ry0 = y + o, rz0 = z + o;
        advance();
        T g0[Q];
        load(g0, at(ry0, rz0));
        collide(g0);
        push(g0, ry0, rz0);
      }
      __syncthreads();
      T* t = A; A = B; B = t;
    }
    // ---- the tile (the part of the region that is still valid) -> global
    for (unsigned i = threadIdx.x; i < p.ty * p.tz; i += RES_BLOCK) {
      const unsigned y = i / p.tz, z = i - y * p.tz;
      const unsigned g = (unsigned(ty0) + y) * p.nz + unsigned(tz0) + z;
      const unsigned idx = at(y + unsigned(h), z + unsigned(h));
#pragma unroll
      for (int q = 0; q < Q; ++q) dst[size_t(q) * N + g] = A[q * plane + idx];
    }
    if (r + 1 < p.rounds) grid.sync();   // also orders this CTA's shared-memory reads before the next round's loads
  }
}

// Plan of a call of n_steps time steps on an ny x nz lattice: tile, rounds, steps per round.  Cost model per round:
// a fixed part (load + store + grid barrier) + the node updates of the shrinking region at the
// measured rate per CTA; the best (tile, h) minimises the time per step.  Constraints: all CTAs co-resident (one per
// SM: the tile needs most of the shared memory), two region buffers within the shared memory of an SM.
struct ResidentPlan {
  bool ok = false;
  unsigned ty = 0, tz = 0;
  int h = 0, rounds = 0, steps_base = 0, steps_extra = 0;
  unsigned plane_max = 0;
  size_t smem = 0;
};

inline ResidentPlan plan_resident(unsigned ny, unsigned nz, int q, size_t elem, int n_steps, int sms, size_t smem_max,
                                  int force_h) {
  ResidentPlan best;
  if (n_steps < 2) return best;
  double best_cost = 1e30;
  // measured on B200 (h sweep at 256^2, profiles/r2_bench_small_lattices_resident_h_sweep.jsonl): ~4.5 us per round
  // (load + store + grid barrier) and 1.7 (fp64) / 1.2 (fp32) ns per node update of ONE CTA
  const double fixed_ns = 4500.0, node_ns = elem == 8 ? 1.7 : 1.2;
  // (divisors first: the search runs on the host inside every multi-step call)
  unsigned dy[128], dz[128];
  int ndy = 0, ndz = 0;
  const unsigned t_max = 4096;   // a tile edge beyond that cannot fit: (ty + 2)(tz + 2) 2 q elem <= 227 KB
  for (unsigned t = 1; t <= ny && t <= t_max && ndy < 128; ++t) if (ny % t == 0) dy[ndy++] = t;
  for (unsigned t = 1; t <= nz && t <= t_max && ndz < 128; ++t) if (nz % t == 0) dz[ndz++] = t;
  for (int iy = 0; iy < ndy; ++iy) {
    const unsigned ty = dy[iy];
    for (int iz = 0; iz < ndz; ++iz) {
      const unsigned tz = dz[iz];
      const unsigned tiles = (ny / ty) * (nz / tz);
      if (tiles > unsigned(sms)) continue;
      for (int h = 1; h <= 16 && h <= n_steps; ++h) {
        if (force_h > 0 && h != force_h) continue;
        if (unsigned(h) > ny || unsigned(h) > nz) break;
        const size_t plane = size_t(ty + 2 * h + 2) * (tz + 2 * h + 2);   // region + one ring of padding
        const size_t smem = 2 * size_t(q) * plane * elem;
        if (smem > smem_max) break;
        double updates = 0;
        for (int j = 1; j <= h; ++j) updates += double(ty + 2 * j) * double(tz + 2 * j);
        const double cost = (fixed_ns + node_ns * updates + 0.15 * double(plane)) / h;
        if (cost < best_cost) {
          best_cost = cost;
          best.ok = true; best.ty = ty; best.tz = tz; best.h = h;
        }
      }
    }
  }
  if (!best.ok) return best;
  // rounds: ceil(n / h), made to have the parity of n so that the state ends where n separate steps would leave it
  int rounds = (n_steps + best.h - 1) / best.h;
  if ((rounds & 1) != (n_steps & 1)) ++rounds;
  best.rounds = rounds;
  best.steps_base = n_steps / rounds;
  best.steps_extra = n_steps % rounds;
  const int hmax = best.steps_base + (best.steps_extra ? 1 : 0);
  best.plane_max = (best.ty + 2 * hmax + 2) * (best.tz + 2 * hmax + 2);
  best.smem = 2 * size_t(q) * best.plane_max * elem;
  return best;
}

// ONE cooperative launch for n_steps time steps; *launched = false (and nothing enqueued) where the lattice does not
// fit.  a = the state, b = the other buffer: after the call the state is in b when n_steps is odd, in a when even.
template <class S, typename T, int COLL>
cudaError_t launch_resident(T* a, T* b, unsigned ny, unsigned nz, const CollisionParams<T>& cp, int n_steps,
                            cudaStream_t st, bool* launched) {
  *launched = false;
  if constexpr (S::D != 2) {
    return cudaSuccess;
  } else {
    auto kernel = step_resident_kernel<S, T, COLL>;
    static int sms = 0, smem_optin = 0;
    if (sms == 0) {
      int dev = 0;
      cudaError_t e = cudaGetDevice(&dev);
      if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (e == cudaSuccess) e = cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin);
      if (e != cudaSuccess) { sms = 0; (void)cudaGetLastError(); return cudaSuccess; }
    }
    static const int force_h = getenv("LB_RES_H") ? atoi(getenv("LB_RES_H")) : 0;   // A/B timing only
    const ResidentPlan pl = plan_resident(ny, nz, S::Q, sizeof(T), n_steps, sms, size_t(smem_optin), force_h);
    if (!pl.ok) return cudaSuccess;
    ResidentParams<T> p{};
    p.buf[0] = a; p.buf[1] = b;
    p.ny = ny; p.nz = nz; p.n_nodes = ny * nz;
    p.ty = pl.ty; p.tz = pl.tz; p.plane_max = pl.plane_max;
    p.rounds = pl.rounds; p.steps_base = pl.steps_base; p.steps_extra = pl.steps_extra;
    p.cp = cp;
    const unsigned grid = (ny / pl.ty) * (nz / pl.tz);
    void* args[] = {&p};
    const cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3(grid), dim3(RES_BLOCK), args,
                                                      pl.smem, st);
    if (e == cudaSuccess) *launched = true;
    else (void)cudaGetLastError();   // e.g. cooperative launch not supported: the caller falls back
    return cudaSuccess;
  }
}

}  // namespace lb
