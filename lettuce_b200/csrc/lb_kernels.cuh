// Fused collide -> boundary -> stream kernel for sm_100a (B200).
//
// One launch = one reference time step (reference/lettuce/_simulation.py:92-96):
//   f*      = where(ncm == 0, collision(f), f); f* = where(ncm == i, boundary_i(f*), f*)
//   f'[q,x] = nsm[q,x] ? f*[q,x] : f*[q, x - e_q]          (periodic, _simulation.py:160-175)
//
// Scheme.  "Aligned push": a thread owns V consecutive nodes along the contiguous z axis
// (V = 16 bytes / sizeof(T)), loads its Q x V populations with 128-bit loads from the node's
// own (aligned) location, does moments / equilibrium / relaxation in registers, and writes
// every population to its destination line (x + ex, y + ey) -- a whole-line offset, so still
// 16-byte aligned.  The +-1 shift along z is done in registers: each lane takes one element
// from its neighbour lane with a warp shuffle, so the z-shifted populations also leave as
// aligned 128-bit stores.  Only the two elements per warp that cross a warp/line edge are
// written as 4(8)-byte scalars (they merge with the neighbouring vector store in L2).
// Every load and store instruction is therefore fully coalesced and the kernel moves
// exactly 2*Q*sizeof(T) bytes per node through HBM: the roofline figure of BASELINE.md.
// The state stays in the reference's canonical form (post-streaming f_t) after every launch.
//
// With a packed streaming mask (MODE 2) a store is suppressed where the destination's mask bit is
// set and the destination node itself writes its kept value, so every (q, node) of the output is
// written exactly once (no race, unlike the reference's generated kernel, SURVEY section 3.3
// "semantic trap").  The kernel keeps its vector width: a hint bit in the packed mask tells a warp
// whether any of its nodes is near a kept population at all; only those warps run the masked stores.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "lb_physics.cuh"

namespace lb {

constexpr int MAX_BND = 8;
constexpr int STEP_BLOCK = 128;
constexpr int N_OBS_PART = 3;  // energy, max|u|^2, mass
constexpr int HALO_STAGE_LINES = 64;  // most lines of which a block can stage the halo elements

enum { BND_BOUNCE_BACK = 1, BND_EQUILIBRIUM = 2, BND_ABB_OUTLET = 3, BND_EQ_OUTLET_P = 4 };
enum { MODE_PLAIN = 0, MODE_NCM = 1, MODE_NCM_SBITS = 2 };
constexpr uint32_t SBITS_NEAR = 0x80000000u;   // streaming_bits: neighbourhood hint, see lb_pack_streaming_mask

// Scratch of the fused reductions (lb_obs_scratch_bytes): ticket counters and per-block / per-group
// partial results.  All counters are zero between launches (each is reset by its last user).
constexpr int OBS_GROUP = 256;   // blocks per first-level group
struct ObsScratch {
  unsigned* global_counter;   // groups finished
  unsigned* group_counters;   // [n_groups] blocks finished per group
  double* block_partials;     // [grid][N_OBS_PART]
  double* group_partials;     // [n_groups][N_OBS_PART]
};

__host__ __device__ inline size_t obs_groups(size_t blocks) { return (blocks + OBS_GROUP - 1) / OBS_GROUP; }
// bytes of the counter region in front (header + group counters, padded to 64)
__host__ __device__ inline size_t obs_counter_bytes(size_t max_blocks) {
  return 64 + ((obs_groups(max_blocks) * sizeof(unsigned) + 63) / 64) * 64;
}
__host__ __device__ inline size_t obs_scratch_bytes(size_t max_blocks) {
  return obs_counter_bytes(max_blocks) + (max_blocks + obs_groups(max_blocks)) * N_OBS_PART * sizeof(double);
}
inline ObsScratch obs_layout(void* scratch, size_t max_blocks) {
  ObsScratch o;
  char* base = static_cast<char*>(scratch);
  o.global_counter = reinterpret_cast<unsigned*>(base);
  o.group_counters = reinterpret_cast<unsigned*>(base + 64);
  o.block_partials = reinterpret_cast<double*>(base + obs_counter_bytes(max_blocks));
  o.group_partials = o.block_partials + max_blocks * N_OBS_PART;
  return o;
}

template <typename T>
struct BoundaryDev {
  int kind;
  const T* rho;
  const T* u;
  long long rs[3];  // strides (elements) along kernel axes X, Y, Z
  long long us[3];
  long long ucs;    // stride between velocity components (lattice order)
  int axis;         // outlet kinds: kernel axis (0 X, 1 Y, 2 Z) normal to the plane ...
  int sign;         // ... and the side: +1 = the last plane of that axis, -1 = the first
  T value;          // BND_EQ_OUTLET_P: rho_outlet
};

template <typename T>
struct StepParams {
  const T* f_in;
  T* f_out;
  const uint8_t* ncm;
  const uint32_t* sbits;
  unsigned nx, ny, nz, nzv;    // kernel-axis sizes; nzv = nz / V
  unsigned n_nodes;            // nx * ny * nz  (< 2^31, checked by the host)
  unsigned n_threads;          // nx * ny * nzv
  CollisionParams<T> cp;
  int n_bnd;
  int has_outlet;              // some boundary is an outlet plane: boundaries are applied in index order
  BoundaryDev<T> bnd[MAX_BND];
  ObsScratch obs;              // scratch of the fused reductions
  double* obs_out;             // [LB_OBS_COUNT]
  T* u_out;                    // [D][n_nodes] or nullptr
  // z-slab decomposition (multi-GPU): when set, populations leaving the slab through its
  // z = nz-1 face (ez = +1) / z = 0 face (ez = -1) are stored to
  //   halo_hi / halo_lo [qidx * halo_qs + dest_line * halo_ls + halo_{hi,lo}_off]
  // instead of wrapping around locally.  qidx = q (direct: the base is the neighbour's
  // population buffer, peer-mapped) or the compact z-face index (packed send buffer).
  // halo_in_*: packed [zface][lines] populations that ARRIVED through the z = 0 face (ez = +1,
  // from the -z neighbour) / the z = nz-1 face (ez = -1, from the +z neighbour) in the previous
  // step; when set they replace the corresponding boundary-plane values at load time, so the
  // received halo never has to be scattered into f between steps.
  const T* halo_in_lo;
  const T* halo_in_hi;
  T* halo_hi;
  T* halo_lo;
  long long halo_qs, halo_ls;
  unsigned halo_hi_off, halo_lo_off;
  int halo_compact;
  int halo_stage_shift;   // see step_body: 0 = isolated stores, else log2(slots per (face, population))
  // "LL" transport (LB_HALO_LL): every packed halo element travels as (payload, tag) words written with ONE
  // 8-byte store, tag = the global index of the time step that produced it.  The receiver's loads of the
  // incoming halo see an element either complete or not at all, so the data is its own flag: no handshake
  // kernel, no fences, no wait on the critical path (only the first blocks of a step ever re-poll).
  // halo_in_* / halo_hi / halo_lo then point to LLWord<T> arrays instead of T arrays.
  int halo_ll;
  uint32_t ll_tag_in, ll_tag_out;
  uint32_t* ll_err;       // set to 1 when a poll gives up (neighbour never delivered): state is then invalid
  // Block b of the grid works on tile (b + block_rot) mod grid.  The host rotates the tiles of an outlet
  // plane at the END of the lattice to the front: they run the long boundary path (neighbour loads,
  // masked stores), and as the last blocks of the grid that path was a tail of 6-9 us per step.
  unsigned block_rot;
  unsigned pf_chunk;   // prefetch kernel: > 0 = consecutive tiles per CTA, scheduled by the hardware (see step_pf_kernel)
  // prefetch kernel: the base pointer of every population, f_in + q N / f_out + q N, computed by the host.  Read from
  // the constant bank they cost nothing; derived in the kernel they were ~170 64-bit integer adds per thread in vector
  // registers (of 1300 instructions of the D3Q27 KBC kernel: the compiler keeps them in the uniform datapath only in
  // the leanest variants)
  const T* in_q[27];
  T* out_q[27];
};

// number of directions q' < q that leave through the same z face as q
template <class S>
__host__ __device__ constexpr int zface_index(int q) {
  int n = 0;
  for (int k = 0; k < q; ++k) n += (S::e(k, 2) == S::e(q, 2)) ? 1 : 0;
  return n;
}
template <class S>
__host__ __device__ constexpr int zface_count() {
  int n = 0;
  for (int k = 0; k < S::Q; ++k) n += (S::e(k, 2) == 1) ? 1 : 0;
  return n;
}

// bit q set for the populations q >= 1 whose velocity is (ex, ey, ez) in kernel axes
template <class S>
__host__ __device__ constexpr uint32_t line_populations(int ex, int ey, int ez) {
  uint32_t m = 0;
  for (int k = 1; k < S::Q; ++k)
    if (S::e(k, 0) == ex && S::e(k, 1) == ey && S::e(k, 2) == ez) m |= (1u << k);
  return m;
}

template <typename T, int V>
struct alignas(sizeof(T) * V) Vec {
  T v[V];
};

template <typename T, int V>
__device__ __forceinline__ Vec<T, V> ld_vec(const T* p) {
  return *reinterpret_cast<const Vec<T, V>*>(p);
}
template <typename T, int V>
__device__ __forceinline__ void st_vec(T* p, const Vec<T, V>& v) {
  *reinterpret_cast<Vec<T, V>*>(p) = v;
}

template <int V>
__device__ __forceinline__ uint32_t ld_mask(const uint8_t* p) {
  if constexpr (V == 4) return *reinterpret_cast<const uint32_t*>(p);
  else if constexpr (V == 2) return *reinterpret_cast<const uint16_t*>(p);
  else return *p;
}

// ---- LL halo words: (payload, tag) pairs with 8-byte single-copy atomicity (the protocol of NCCL's LL
// algorithm: st/ld.volatile.global.v2/v4.u32 over NVLink peer mappings) ------------------------------
template <typename T> struct LLWord;
template <> struct LLWord<float> { using type = uint2; };    // {bits, tag}
template <> struct LLWord<double> { using type = uint4; };   // {lo, tag, hi, tag}: two atomic halves

__device__ __forceinline__ void ll_store(uint2* p, float v, uint32_t tag) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(tag) : "memory");
}
__device__ __forceinline__ void ll_store(uint4* p, double v, uint32_t tag) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(uint32_t(b)), "r"(tag),
               "r"(uint32_t(b >> 32)), "r"(tag) : "memory");
}
__device__ __forceinline__ bool ll_load(const uint2* p, uint32_t tag, float& v) {
  uint32_t d, t;
  asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d), "=r"(t) : "l"(p) : "memory");
  v = __uint_as_float(d);
  return t == tag;
}
__device__ __forceinline__ bool ll_load(const uint4* p, uint32_t tag, double& v) {
  uint32_t d0, t0, d1, t1;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(d0), "=r"(t0), "=r"(d1), "=r"(t1) : "l"(p) : "memory");
  v = __longlong_as_double((long long)(((unsigned long long)d1 << 32) | d0));
  return t0 == tag && t1 == tag;
}
constexpr long long LL_TIMEOUT_CYCLES = 20000000000ll;   // ~10 s: far beyond any skew between ranks of one node

// slow path of ll_gather, out of line so that it costs the step kernels no registers: poll until all n
// elements carry `tag` (bounded: sets *err and gives up)
template <typename W>
__device__ __noinline__ void ll_wait(const W* src, size_t stride, int n, uint32_t tag, uint32_t* err) {
  const long long t0 = clock64();
  for (;;) {
    bool ok = true;
    for (int k = 0; k < n; ++k) {
      typename std::conditional<sizeof(W) == 8, float, double>::type v;
      ok &= ll_load(src + k * stride, tag, v);
    }
    if (ok) return;
    if (clock64() - t0 > LL_TIMEOUT_CYCLES) { *err = 1u; return; }   // never hang the GPU
    __nanosleep(100);
  }
}

// first attempt: an ordinary load (the compiler schedules it with the population loads; a stale cached copy
// would carry a stale tag and only send us to the slow path, never to a wrong value: word and tag are one
// 8-byte unit)
__device__ __forceinline__ bool ll_peek(const uint2* p, uint32_t tag, float& v) {
  const uint2 w = *p;
  v = __uint_as_float(w.x);
  return w.y == tag;
}
__device__ __forceinline__ bool ll_peek(const uint4* p, uint32_t tag, double& v) {
  const uint4 w = *p;
  v = __longlong_as_double((long long)(((unsigned long long)w.z << 32) | w.x));
  return w.y == tag && w.w == tag;
}

// NF tagged elements src[k * stride], k < NF, all loads in flight together; where a tag is not there yet (only
// the first blocks of a step ever see that) wait out of line and load again
template <int NF, typename T>
__device__ __forceinline__ void ll_gather(const void* base, size_t index, size_t stride, uint32_t tag, uint32_t* err,
                                          T (&out)[NF]) {
  using W = typename LLWord<T>::type;
  const W* src = static_cast<const W*>(base) + index;
  bool ok = true;
#pragma unroll
  for (int k = 0; k < NF; ++k) ok &= ll_peek(src + k * stride, tag, out[k]);
  if (ok) return;
  ll_wait<W>(src, stride, NF, tag, err);
#pragma unroll
  for (int k = 0; k < NF; ++k) ll_load(src + k * stride, tag, out[k]);
}

// (sum, max, sum) over the block; the result is valid in thread 0.  Fixed order: lanes by a
// shuffle tree in the accumulator type A (float partials of an fp32 lattice stay float inside the
// warp: 64 nodes, half the shuffles of a double), then warps in order in double.
template <int BLOCK, typename A>
__device__ __forceinline__ void block_reduce3(A wa, A wb, A wc, double& a, double& b, double& c) {
  __shared__ double sm[N_OBS_PART][BLOCK / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    wa += __shfl_down_sync(0xffffffffu, wa, o);
    wb = max_(wb, __shfl_down_sync(0xffffffffu, wb, o));
    wc += __shfl_down_sync(0xffffffffu, wc, o);
  }
  __syncthreads();   // sm may still be read by the previous use
  if (lane == 0) { sm[0][warp] = double(wa); sm[1][warp] = double(wb); sm[2][warp] = double(wc); }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = sm[0][0]; b = sm[1][0]; c = sm[2][0];
#pragma unroll
    for (int w = 1; w < BLOCK / 32; ++w) { a += sm[0][w]; b = fmax(b, sm[1][w]); c += sm[2][w]; }
  }
}
template <int BLOCK>
__device__ __forceinline__ void block_reduce3(double& a, double& b, double& c) {
  block_reduce3<BLOCK, double>(a, b, c, a, b, c);
}

// Deterministic grid reduction of (sum, max, sum) without value atomics, in two levels so that no
// single block ever walks more than OBS_GROUP or grid / OBS_GROUP partials (a 512^3 lattice has
// 5e5 blocks): the last block of each group of OBS_GROUP consecutive blocks to finish (atomic ticket)
// sums that group's block partials in a fixed order; the last group to finish sums the group partials.
// Which block does the summing varies from run to run, what it sums and in which order does not:
// bitwise reproducible.
template <int BLOCK>
__device__ __forceinline__ void reduce_observables(double e, double m2, double mass, const ObsScratch& sc,
                                                   double* out, int n_out = N_OBS_PART) {
  __shared__ bool is_last;
  const unsigned grid = gridDim.x, group = blockIdx.x / OBS_GROUP;
  const unsigned n_groups = (grid + OBS_GROUP - 1) / OBS_GROUP;
  const unsigned first = group * OBS_GROUP;
  const unsigned group_size = min(unsigned(OBS_GROUP), grid - first);
  block_reduce3<BLOCK>(e, m2, mass);
  if (threadIdx.x == 0) {
    double* bp = sc.block_partials + size_t(blockIdx.x) * N_OBS_PART;
    bp[0] = e; bp[1] = m2; bp[2] = mass;
    __threadfence();
    is_last = (atomicAdd(&sc.group_counters[group], 1u) == group_size - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  double a = 0.0, b = 0.0, c = 0.0;
  for (unsigned i = threadIdx.x; i < group_size; i += BLOCK) {
    const double* bp = sc.block_partials + size_t(first + i) * N_OBS_PART;
    a += __ldcg(bp + 0);
    b = fmax(b, __ldcg(bp + 1));
    c += __ldcg(bp + 2);
  }
  block_reduce3<BLOCK>(a, b, c);
  if (threadIdx.x == 0) {
    double* gp = sc.group_partials + size_t(group) * N_OBS_PART;
    gp[0] = a; gp[1] = b; gp[2] = c;
    sc.group_counters[group] = 0u;
    __threadfence();
    is_last = (atomicAdd(sc.global_counter, 1u) == n_groups - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  a = 0.0; b = 0.0; c = 0.0;
  for (unsigned i = threadIdx.x; i < n_groups; i += BLOCK) {
    const double* gp = sc.group_partials + size_t(i) * N_OBS_PART;
    a += __ldcg(gp + 0);
    b = fmax(b, __ldcg(gp + 1));
    c += __ldcg(gp + 2);
  }
  block_reduce3<BLOCK>(a, b, c);
  if (threadIdx.x == 0) {
    out[0] = a;                  // LB_OBS_ENERGY (or the single sum when n_out == 1)
    if (n_out > 1) out[1] = b;   // LB_OBS_MAXU2
    if (n_out > 2) out[2] = c;   // LB_OBS_MASS
    *sc.global_counter = 0u;
  }
}

// equilibrium_boundary_pu.py:27-32: prescribed (rho, u) of boundary `bd` at node (x, y, z), lattice units
template <class S, typename T>
__device__ __forceinline__ void boundary_values(const BoundaryDev<T>& bd, unsigned x, unsigned y, unsigned z, T& rb,
                                                T (&ub)[3]) {
  const long long zz = z;
  rb = bd.rho[x * bd.rs[0] + y * bd.rs[1] + zz * bd.rs[2]];
  ub[0] = ub[1] = ub[2] = T(0);
  const long long off = x * bd.us[0] + y * bd.us[1] + zz * bd.us[2];
#pragma unroll
  for (int c = 0; c < S::D; ++c) ub[S::axis(c)] = bd.u[c * bd.ucs + off];
}

// Velocity the reference's outlet boundary `i` (1-based) sees on the node next to its plane, i.e. of
// that node's state after the collision and the boundaries of lower index.  Collisions conserve rho
// and u, so it is taken from the node's incoming populations and corrected for a bounce-back
// (u -> -u) or equilibrium boundary (u -> prescribed u) of lower index on that node.
template <class S, typename T>
__device__ __forceinline__ void outlet_neighbour_velocity(const StepParams<T>& p, const T* __restrict__ f_in, int i,
                                                          unsigned xn, unsigned yn, unsigned zn, T (&un)[3]) {
  const size_t N = p.n_nodes;
  const unsigned nb = (xn * p.ny + yn) * p.nz + zn;
  T h[S::Q], rn;
  static_for<S::Q>([&](auto qc) { h[decltype(qc)::value] = f_in[decltype(qc)::value * N + nb]; });
  moments<S, T>(h, rn, un);
  const int mn = p.ncm[nb];
  if (mn > 0 && mn < i) {
    const BoundaryDev<T>& bn = p.bnd[mn - 1];
    if (bn.kind == BND_BOUNCE_BACK) {
      un[0] = -un[0]; un[1] = -un[1]; un[2] = -un[2];
    } else if (bn.kind == BND_EQUILIBRIUM) {
      T rb;
      boundary_values<S, T>(bn, xn, yn, zn, rb, un);
    }
  }
}

// The outlet boundaries: in place on their plane (anti_bounce_back_outlet.py:71-86,
// equilibrium_outlet_p.py:63-71)
template <class S, typename T>
__device__ __forceinline__ void outlet_node(T (&g)[S::Q], const BoundaryDev<T>& bd, const T (&un)[3]) {
  if (bd.kind == BND_EQ_OUTLET_P) {
    equilibrium<S, T>(bd.value, un, g);
    return;
  }
  T rho, u[3];
  moments<S, T>(g, rho, u);
  const T uw[3] = {u[0] + T(0.5) * (u[0] - un[0]), u[1] + T(0.5) * (u[1] - un[1]), u[2] + T(0.5) * (u[2] - un[2])};
  const T uw2 = fma_(uw[0], uw[0], fma_(uw[1], uw[1], uw[2] * uw[2]));
  const T inv_cs2 = T(1.0 / LB_CS2), inv_cs4 = T(1.0 / (LB_CS2 * LB_CS2));
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    const int ed = (bd.axis == 0 ? S::e(q, 0) : (bd.axis == 1 ? S::e(q, 1) : S::e(q, 2))) * bd.sign;
    if (ed == 1) {   // leaves through the plane: its opposite is rebuilt
      const T eu = T(S::e(q, 0)) * uw[0] + T(S::e(q, 1)) * uw[1] + T(S::e(q, 2)) * uw[2];
      g[S::opp(q)] = T(S::w(q)) * rho * (T(2) + eu * eu * inv_cs4 - uw2 * inv_cs2) - g[q];
    }
  });
}

// u [D][n_nodes] of one node for the enstrophy kernel.  Out of line ON PURPOSE: only the reporting steps of an
// Enstrophy reporter take it, but inlined the compiler predicates the D stores and their 64-bit address arithmetic into
// the hot path of every OBS kernel; a call is a branch that the other steps skip.  c0, c1, c2 = components in lattice
// order (c2 ignored in 2-D).
template <int D, typename T>
__device__ __noinline__ void store_velocity(T* u_out, unsigned n_nodes, unsigned node, T c0, T c1, T c2) {
  u_out[node] = c0;
  u_out[size_t(n_nodes) + node] = c1;
  if constexpr (D == 3) u_out[2 * size_t(n_nodes) + node] = c2;
}

// One node: moments -> (fused observables) -> collision where ncm == 0, boundary i where ncm == i
// OUTLETS: compile the in-order boundary loop that outlet planes need (the masked-streaming kernels --
// outlets always come with a streaming mask -- and the one-node-per-thread kernels)
template <class S, typename T, int COLL, int MODE, bool OBS, bool OUTLETS>
__device__ __forceinline__ void node_update(T (&g)[S::Q], unsigned m, const StepParams<T>& p,
                                            const T* __restrict__ f_in, T* u_out, unsigned x,
                                            unsigned y, unsigned z, unsigned node, bool active, T& acc_e,
                                            T& acc_m2, T& acc_mass) {
  T a[N_SLOTS], b[N_SLOTS];
  T rho = T(0), u[3] = {T(0), T(0), T(0)};
  if (OBS || (COLL != COLL_NONE && m == 0)) {
    pair_split<S, T>(g, a, b);
    moments_from_pairs<S, T>(g[0], a, b, rho, u);
  }
  if constexpr (OBS) {
    const T u2 = fma_(u[0], u[0], fma_(u[1], u[1], u[2] * u[2]));
    acc_e += T(0.5) * u2;
    acc_m2 = max_(acc_m2, u2);
    acc_mass += rho;
    if (u_out != nullptr && active)    // only on steps that feed the enstrophy kernel
      store_velocity<S::D, T>(u_out, p.n_nodes, node, u[S::axis(0)], u[S::axis(1)], u[S::axis(S::D - 1)]);
  }
  if (m == 0) {
    collide_node<S, T, COLL, T>(g, a, b, rho, u, p.cp);
  } else if constexpr (MODE != MODE_PLAIN) {
    if (!OUTLETS || !p.has_outlet) {
      if (m <= unsigned(p.n_bnd)) {
        const BoundaryDev<T>& bd = p.bnd[m - 1];
        if (bd.kind == BND_BOUNCE_BACK) {
          bounce_back_node<S, T>(g);
        } else if (bd.kind == BND_EQUILIBRIUM) {
          T rb, ub[3];
          boundary_values<S, T>(bd, x, y, z, rb, ub);
          equilibrium<S, T>(rb, ub, g);
        }
      }
    } else {
      // with outlet planes the boundaries act one after the other, in index order (_simulation.py:186-188):
      // a mask boundary where ncm == i, an outlet on every node of its plane (a plane node always has ncm != 0)
      for (int i = 1; i <= p.n_bnd; ++i) {
        const BoundaryDev<T>& bd = p.bnd[i - 1];
        if (bd.kind == BND_ABB_OUTLET || bd.kind == BND_EQ_OUTLET_P) {
          const unsigned c = bd.axis == 0 ? x : (bd.axis == 1 ? y : z);
          const unsigned n = bd.axis == 0 ? p.nx : (bd.axis == 1 ? p.ny : p.nz);
          if (c == (bd.sign > 0 ? n - 1u : 0u)) {
            const unsigned cn = bd.sign > 0 ? n - 2u : 1u;
            T un[3];
            outlet_neighbour_velocity<S, T>(p, f_in, i, bd.axis == 0 ? cn : x, bd.axis == 1 ? cn : y,
                                            bd.axis == 2 ? cn : z, un);
            outlet_node<S, T>(g, bd, un);
          }
        } else if (m == unsigned(i)) {
          if (bd.kind == BND_BOUNCE_BACK) {
            bounce_back_node<S, T>(g);
          } else if (bd.kind == BND_EQUILIBRIUM) {
            T rb, ub[3];
            boundary_values<S, T>(bd, x, y, z, rb, ub);
            equilibrium<S, T>(rb, ub, g);
          }
        }
      }
    }
  }
}

// Two z-adjacent fp32 fluid nodes (both ncm == 0) as one packed value per population
template <class S, int COLL, bool OBS>
__device__ __forceinline__ void packed_update(F2 (&g)[S::Q], const StepParams<float>& p, float* u_out, unsigned node,
                                              bool active, float& acc_e, float& acc_m2, float& acc_mass) {
  if constexpr (!OBS && COLL == COLL_NONE) return;
  F2 a[N_SLOTS], b[N_SLOTS], rho, u[3];
  pair_split<S, F2>(g, a, b);
  moments_from_pairs<S, F2>(g[0], a, b, rho, u);
  if constexpr (OBS) {
    const F2 u2 = fma_(u[0], u[0], fma_(u[1], u[1], u[2] * u[2]));
    acc_e += 0.5f * u2.x;
    acc_e += 0.5f * u2.y;
    acc_m2 = max_(acc_m2, hmax(u2));
    acc_mass += rho.x;
    acc_mass += rho.y;
    if (u_out != nullptr && active) {
      store_velocity<S::D, float>(u_out, p.n_nodes, node, u[S::axis(0)].x, u[S::axis(1)].x, u[S::axis(S::D - 1)].x);
      store_velocity<S::D, float>(u_out, p.n_nodes, node + 1, u[S::axis(0)].y, u[S::axis(1)].y, u[S::axis(S::D - 1)].y);
    }
  }
  collide_node<S, F2, COLL, float>(g, a, b, rho, u, p.cp);
}

// collide / boundaries of the V nodes of a thread, all in registers.
// fp32 threads that own an even number of nodes run them two at a time as packed F2 values
// (FADD2 / FMUL2 / FFMA2); a pair that contains a boundary node, and every other case, goes node
// by node.  Both routes round identically (lb_physics.cuh), so the result does not depend on V.
template <class S, typename T, int V, int COLL, int MODE, bool OBS>
__device__ __forceinline__ void collide_thread(T (&f)[S::Q][V], uint32_t mword, const StepParams<T>& p,
                                               const T* __restrict__ f_in, T* u_out, unsigned x, unsigned y, unsigned z0,
                                               unsigned node, bool active, T& acc_e, T& acc_m2, T& acc_mass) {
  constexpr int Q = S::Q;
  constexpr bool PACK = (sizeof(T) == 4) && (V % 2 == 0);
  // the outlet kinds are compiled into the kernels that can meet them: outlets come with a
  // no-streaming mask (MODE_NCM_SBITS); collide-only calls and odd layouts run V == 1
  constexpr bool OUTL = (V == 1) || (MODE == MODE_NCM_SBITS);
  if constexpr (PACK) {
#pragma unroll
    for (int kk = 0; kk < V / 2; ++kk) {
      const unsigned m2 = (MODE == MODE_PLAIN) ? 0u : ((mword >> (16 * kk)) & 0xffffu);
      if (m2 == 0) {
        F2 g[Q];
        static_for<Q>([&](auto qc) {
          constexpr int q = decltype(qc)::value;
          g[q] = F2(f[q][2 * kk], f[q][2 * kk + 1]);
        });
        packed_update<S, COLL, OBS>(g, p, u_out, node + 2 * kk, active, acc_e, acc_m2, acc_mass);
        static_for<Q>([&](auto qc) {
          constexpr int q = decltype(qc)::value;
          f[q][2 * kk] = g[q].x;
          f[q][2 * kk + 1] = g[q].y;
        });
      } else {
#pragma unroll
        for (int k = 2 * kk; k < 2 * kk + 2; ++k) {
          T g[Q];
          static_for<Q>([&](auto qc) { g[decltype(qc)::value] = f[decltype(qc)::value][k]; });
          node_update<S, T, COLL, MODE, OBS, OUTL>(g, (m2 >> (8 * (k - 2 * kk))) & 0xffu, p, f_in, u_out, x, y, z0 + k, node + k, active,
                                             acc_e, acc_m2, acc_mass);
          static_for<Q>([&](auto qc) { f[decltype(qc)::value][k] = g[decltype(qc)::value]; });
        }
      }
    }
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      T g[Q];
      static_for<Q>([&](auto qc) { g[decltype(qc)::value] = f[decltype(qc)::value][k]; });
      const unsigned m = (MODE == MODE_PLAIN) ? 0u : ((mword >> (8 * k)) & 0xffu);
      node_update<S, T, COLL, MODE, OBS, OUTL>(g, m, p, f_in, u_out, x, y, z0 + k, node + k, active, acc_e, acc_m2, acc_mass);
      static_for<Q>([&](auto qc) { f[decltype(qc)::value][k] = g[decltype(qc)::value]; });
    }
  }
}

// Resident blocks asked of the register allocator.  D3Q19 fp32 with 4 nodes per thread: 76 population
// registers, 4 blocks (<= 128 registers) as measured at the HBM roofline -- except TRT, whose packed
// pair form keeps a_p, b_p, S_p, D_p of two nodes live and spills below 168; D3Q27 fp32 with 2 or 4
// nodes per thread (54 / 108 population registers): 3 blocks (<= 168 registers).
template <class S, typename T, int V, int COLL, int MODE = MODE_PLAIN>
constexpr int step_min_blocks() {
  // masked streaming keeps a few more registers per thread (destination bits): one block less
  if (sizeof(T) == 4 && V >= 2 && MODE == MODE_NCM_SBITS) return (S::Q == 27) ? 2 : 3;
  // D3Q27 KBC / TRT / Smagorinsky fp32 with four nodes per thread: up to 255 registers, no spills, two CTAs per SM
  // (see vector_nodes)
  if (sizeof(T) == 4 && V == 4 && S::Q == 27 && (COLL == COLL_KBC || COLL == COLL_TRT || COLL == COLL_SMAGORINSKY) &&
      MODE != MODE_NCM_SBITS)
    return 2;
  if (sizeof(T) == 4 && V >= 2) return (S::Q == 27 || ((COLL == COLL_TRT || COLL >= COLL_REGULARIZED) && V == 4)) ? 3 : 4;
  return 1;
}

// One time step of the lattice for this thread and tile `blk` (STEP_BLOCK threads x V nodes): f_in -> f_out, and the
// thread's observables of f_in (sum 0.5 |u|^2, max |u|^2, sum rho over its V nodes; zero for a thread beyond the
// lattice) into t_e / t_m2 / t_mass when OBS.  SMEM_IN != 0: the tile's populations are already in shared memory
// (smem_in[q][STEP_BLOCK * V]) instead of global memory --
//   SMEM_TMA:     brought in by bulk copies for the whole block; the block hands the stage back with a barrier;
//   SMEM_PRIVATE: every thread copied ITS OWN V values per population (cp.async) and reads only those: no barrier,
//                 the tile's mask word arrives in `mword_pf` (loaded into a register one tile ahead).
enum { SMEM_NONE = 0, SMEM_TMA = 1, SMEM_PRIVATE = 2 };
struct NoRefill { __device__ __forceinline__ void operator()() const {} };

template <class S, typename T, int V, int COLL, int MODE, bool OBS, bool STREAM, bool HALO, int SMEM_IN = SMEM_NONE,
          class Refill = NoRefill>
__device__ __forceinline__ void step_tile(const StepParams<T>& p, const T* __restrict__ f_in, T* __restrict__ f_out,
                                          T* u_out, unsigned blk, const T* smem_in, T& t_e, T& t_m2, T& t_mass,
                                          Refill smem_tile_consumed = Refill{}, uint32_t mword_pf = 0u) {
  constexpr int Q = S::Q;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned tid = blk * STEP_BLOCK + threadIdx.x;
  const bool active = tid < p.n_threads;
  const unsigned t = active ? tid : p.n_threads - 1u;  // clamp: keeps address math valid
  const unsigned nzv = p.nzv, nz = p.nz, ny = p.ny, nx = p.nx;
  const unsigned line = t / nzv;
  const unsigned zv = t - line * nzv;
  const unsigned x = line / ny;
  const unsigned y = line - x * ny;
  const unsigned z0 = zv * V;
  const unsigned node = line * nz + z0;
  const size_t N = p.n_nodes;

  constexpr int NF = zface_count<S>();

  // ---- masked streaming (no_streaming_mask packed as one bit per population and node, p.sbits) ----
  // Bit 31 of a node's word says that the node or one of its 26 neighbours keeps a population
  // (lb_pack_streaming_mask).  Loaded HERE with the populations; a warp none of whose nodes has it set --
  // almost all of them -- streams with the plain vector stores.
  constexpr bool SB = (MODE == MODE_NCM_SBITS) && STREAM;
  [[maybe_unused]] uint32_t near_kept = 0, own_first = 0, own_last = 0;
  if constexpr (SB) {
    const Vec<uint32_t, V> ow = ld_vec<uint32_t, V>(p.sbits + node);
#pragma unroll
    for (int k = 0; k < V; ++k) near_kept |= ow.v[k];
    own_first = ow.v[0];
    own_last = ow.v[V - 1];
  }
  [[maybe_unused]] auto kept = [](uint32_t bits, int q) { return ((bits >> q) & 1u) != 0u; };

  // ---- load: Q aligned vector loads ------------------------------------------------
  T f[Q][V];
  {
    static_for<Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      const Vec<T, V> v = SMEM_IN ? ld_vec<T, V>(smem_in + q * (STEP_BLOCK * V) + threadIdx.x * V)
                                  : ld_vec<T, V>(f_in + size_t(q) * N + node);
#pragma unroll
      for (int k = 0; k < V; ++k) f[q][k] = v.v[k];
    });
  }
  [[maybe_unused]] uint32_t smem_mword = 0;
  if constexpr (SMEM_IN == SMEM_TMA && MODE != MODE_PLAIN)
    smem_mword = ld_mask<V>(reinterpret_cast<const uint8_t*>(smem_in + Q * (STEP_BLOCK * V)) + threadIdx.x * V);
  if constexpr (SMEM_IN == SMEM_TMA) {
    // the stage is free as soon as every thread holds its values: the producer refills it while this tile
    // computes and stores (the caller issues the next bulk copies right behind this barrier)
    __syncthreads();
    smem_tile_consumed();
  }
  if constexpr (SMEM_IN == SMEM_PRIVATE) smem_tile_consumed();   // the thread's own slots: refilled at once
  uint32_t mword = 0;
  if constexpr (MODE != MODE_PLAIN) {
    if constexpr (SMEM_IN == SMEM_TMA) mword = smem_mword;     // the tile's mask bytes came with the populations
    else if constexpr (SMEM_IN == SMEM_PRIVATE) mword = mword_pf;
    else mword = ld_mask<V>(p.ncm + node);
  }
  // ---- incoming halo (multi-GPU): issued right behind the population loads so that both are in flight
  // together (LL transport: the tag test then waits for these few words only, not in front of the population
  // loads); a thread is on at most one z face unless the line is a single vector.
  // A received element is dropped where the node keeps its own value of that population (the sender does not
  // know the receiver's no-streaming bits: it always sends, the destination decides -- _simulation.py:171-174)
  const bool in_lo = HALO && (p.halo_in_lo != nullptr) && (zv == 0);
  const bool edge_in = HALO && (in_lo || ((p.halo_in_hi != nullptr) && (zv == nzv - 1)));
  const bool edge_in2 = HALO && in_lo && (p.halo_in_hi != nullptr) && (nzv == 1);
  if (edge_in) {
    T hin[NF] = {};
    const size_t lines = size_t(nx) * ny;
    if (p.halo_ll) {
      ll_gather<NF, T>(in_lo ? p.halo_in_lo : p.halo_in_hi, line, lines, p.ll_tag_in, p.ll_err, hin);
    } else {
      const T* hsrc = (in_lo ? p.halo_in_lo : p.halo_in_hi) + line;
#pragma unroll
      for (int k = 0; k < NF; ++k) hin[k] = hsrc[k * lines];
    }
    static_for<Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      if constexpr (S::e(q, 2) == 1) { if (in_lo && !(SB && kept(own_first, q))) f[q][0] = hin[zface_index<S>(q)]; }
      if constexpr (S::e(q, 2) == -1) { if (!in_lo && !(SB && kept(own_last, q))) f[q][V - 1] = hin[zface_index<S>(q)]; }
    });
  }
  if (edge_in2) {   // one-vector lines (nzv == 1): the thread is on both faces (rare: loaded here, not up front)
    T hin2[NF] = {};
    const size_t lines = size_t(nx) * ny;
    if (p.halo_ll) {
      ll_gather<NF, T>(p.halo_in_hi, line, lines, p.ll_tag_in, p.ll_err, hin2);
    } else {
#pragma unroll
      for (int k = 0; k < NF; ++k) hin2[k] = p.halo_in_hi[k * lines + line];
    }
    static_for<Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      if constexpr (S::e(q, 2) == -1) { if (!(SB && kept(own_last, q))) f[q][V - 1] = hin2[zface_index<S>(q)]; }
    });
  }

  T acc_e = T(0), acc_mass = T(0), acc_m2 = T(0);   // per-thread partials over V nodes; widened below

  // ---- collide / boundaries, all in registers -----------------------------------------
  collide_thread<S, T, V, COLL, MODE, OBS>(f, mword, p, f_in, u_out, x, y, z0, node, active, acc_e, acc_m2, acc_mass);

  // ---- stream: push to (x + ex, y + ey, z + ez), periodic -----------------------------
  if constexpr (!STREAM) {
    if (active) {
      static_for<Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        Vec<T, V> v;
#pragma unroll
        for (int k = 0; k < V; ++k) v.v[k] = f[q][k];
        st_vec<T, V>(f_out + q * N + node, v);
      });
    }
  } else {
    // Addressing: one uniform 64-bit base per population (f_out + q N) plus a 32-bit element index
    // per thread, so an address costs one wide multiply-add.  All index arithmetic sits outside the
    // `active` / edge predicates: the guarded statements are single stores, which the compiler
    // predicates instead of branching around.
    const unsigned xp = (x + 1 == nx) ? 0u : x + 1, xm = (x == 0) ? nx - 1 : x - 1;
    const unsigned yp = (y + 1 == ny) ? 0u : y + 1, ym = (y == 0) ? ny - 1 : y - 1;
    const unsigned xs[3] = {xm * ny, x * ny, xp * ny};
    const unsigned ys[3] = {ym, y, yp};
    const bool first_in_line = (zv == 0), last_in_line = (zv == nzv - 1);
    const bool has_left = (lane > 0) && !first_in_line;
    const bool has_right = (lane < 31) && !last_in_line;
    const bool st_left = active && has_left, st_noleft = active && !has_left;
    const bool st_right = active && has_right, st_noright = active && !has_right;
    const bool edge_up = active && (lane == 31 || last_in_line);   // owns the element that leaves the warp upwards
    const bool edge_dn = active && (lane == 0 || first_in_line);
    // z-slabs: this thread's upward / downward element leaves through the slab's z face (to the neighbour rank)
    const bool cross_up = HALO && active && last_in_line && p.halo_hi != nullptr;
    const bool cross_dn = HALO && active && first_in_line && p.halo_lo != nullptr;
    // element offset of that edge element relative to this thread's own vector in the destination line
    const unsigned d_up = last_in_line ? 0u - z0 : unsigned(V);
    const unsigned d_dn = first_in_line ? nz - 1u : 0u - 1u;

    // z-slab halos in the PACKED layout [z-face population][line]: the one element per line and
    // population that leaves through a z face is staged in shared memory and written by consecutive
    // lanes for consecutive lines, i.e. as contiguous runs (full sectors for >= 8 lines per block)
    // instead of one isolated 4-byte store per warp -- isolated remote stores over NVLink were
    // measured to cost up to 20 % of a step on 1024^2 faces.
    // Slot of (face, population k, line j of this block): ((face * NF + k) << halo_stage_shift) | j; the
    // host sets halo_stage_shift = log2 of the next power of two above the lines a block can touch
    // (0 = do not stage: more than HALO_STAGE_LINES lines per block, or a non-packed layout).
    __shared__ T stage_val[(HALO && V > 1) ? 2 * NF * HALO_STAGE_LINES : 1];
    __shared__ unsigned stage_dst[(HALO && V > 1) ? 2 * NF * HALO_STAGE_LINES : 1];
    const unsigned sshift = HALO ? unsigned(p.halo_stage_shift) : 0u;
    const bool staged = HALO && V > 1 && sshift != 0u;
    const unsigned first_line_blk = (blk * unsigned(STEP_BLOCK)) / nzv;
    [[maybe_unused]] const unsigned lline = line - first_line_blk;
    // one halo element to the neighbour (or the local send buffer): a plain store, or a tagged LL word
    [[maybe_unused]] auto halo_store = [&](T* base, long long index, T val) {
      if (p.halo_ll) ll_store(reinterpret_cast<typename LLWord<T>::type*>(base) + index, val, p.ll_tag_out);
      else base[index] = val;
    };

    // The masked variant of the loop below (element-by-element stores under the no-streaming bits) is a
    // SECOND copy of the code, entered by whole warps: a warp none of whose lanes saw a set bit -- almost
    // all of them -- runs exactly the unmasked stores.  (One copy with per-population tests was measured
    // at 1640 instructions per warp and 0.70 of the roofline: the compiler predicates the rare path in.)
    auto stream_all = [&](auto masked_c) {
    constexpr bool MASKED = SB && decltype(masked_c)::value;
    // masked copy: ALL the bits it will test in one batch of independent loads (tested one load at a time,
    // the blocks on an outlet plane were a 9 us tail of the step):
    //   own_m[k]   bit q: this thread's node k keeps its own value of population q
    //   dst_m[k]   bit q: element k of the aligned vector this thread writes for population q falls on a
    //              kept node of the DESTINATION line (one vector load of bits per (ex, ey) neighbour line)
    //   edge_m_up / edge_m_dn: the same for the single element that leaves the thread's vector upwards /
    //              downwards (ez = +1 / -1; with V == 1 every destination of ez != 0 is such an element)
    [[maybe_unused]] uint32_t own_m[V], dst_m[V];
    [[maybe_unused]] uint32_t own_any = 0, dst_any = 0, edge_m_up = 0, edge_m_dn = 0;
    if constexpr (MASKED) {
      const bool ld_up = active && (V == 1 || edge_up), ld_dn = active && (V == 1 || edge_dn);
      const Vec<uint32_t, V> ow = ld_vec<uint32_t, V>(p.sbits + node);
#pragma unroll
      for (int k = 0; k < V; ++k) { own_m[k] = ow.v[k]; own_any |= ow.v[k]; dst_m[k] = 0u; }
      static_for<9>([&](auto cc) {
        constexpr int c = decltype(cc)::value;
        constexpr int ex = c / 3 - 1, ey = c % 3 - 1;
        constexpr uint32_t m0 = line_populations<S>(ex, ey, 0), mu = line_populations<S>(ex, ey, 1),
                           md = line_populations<S>(ex, ey, -1);
        constexpr uint32_t in_vec = (V == 1) ? m0 : (m0 | mu | md);
        if constexpr ((m0 | mu | md) != 0) {
          const unsigned i0 = (xs[ex + 1] + ys[ey + 1]) * nz + z0;
          if constexpr (in_vec != 0) {
            const Vec<uint32_t, V> w = ld_vec<uint32_t, V>(p.sbits + i0);
#pragma unroll
            for (int k = 0; k < V; ++k) dst_m[k] |= w.v[k] & in_vec;
          }
          if constexpr (mu != 0) edge_m_up |= (ld_up ? p.sbits[i0 + d_up] : 0u) & mu;
          if constexpr (md != 0) edge_m_dn |= (ld_dn ? p.sbits[i0 + d_dn] : 0u) & md;
        }
      });
#pragma unroll
      for (int k = 0; k < V; ++k) dst_any |= dst_m[k];
    }
    static_for<Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      constexpr int ex = S::e(q, 0), ey = S::e(q, 1), ez = S::e(q, 2);
      const unsigned dlines = xs[ex + 1] + ys[ey + 1];   // destination line number
      const unsigned dline = dlines * nz;                // its first element
      [[maybe_unused]] const unsigned idx = dline + z0;  // this thread's vector in the destination line
      T* const bq = (SMEM_IN == SMEM_PRIVATE) ? p.out_q[q] : f_out + size_t(q) * N;   // uniform

      // The ONE element of this thread that leaves its vector upwards (ez = +1) / downwards (ez = -1): into the
      // next lane's vector or around the line end (periodic lattice).  `local_ok` = the destination node does not
      // keep its own value (masked streaming).  With z-slabs the element of a line-end thread leaves through the
      // slab's z face instead: that is done for all face populations together behind this loop (send_faces).
      // Packed halo layouts ALSO store to the local wrap position: the value there is never read (the
      // neighbour's element is consumed from the receive buffer at load time and scattered at the end
      // of the call), but without it the boundary sector of every line stays partially written, and
      // every partially written sector costs a read-modify-write in HBM (measured: 10 % of a step).
      [[maybe_unused]] auto push_edge = [&](T val, bool up, bool local_ok) {
        const bool cross = up ? cross_up : cross_dn;
        if ((!cross || p.halo_compact) && local_ok) bq[idx + (up ? d_up : d_dn)] = val;
      };

      if constexpr (MASKED && q != 0) {
        // own values that streaming must not overwrite (the destination tests below keep every other
        // thread from writing there)
        if (active && kept(own_any, q)) {
#pragma unroll
          for (int k = 0; k < V; ++k)
            if (kept(own_m[k], q)) bq[node + k] = f[q][k];
        }
      }
      if constexpr (MASKED && q != 0 && ez == 0) {
        Vec<T, V> v;
#pragma unroll
        for (int k = 0; k < V; ++k) v.v[k] = f[q][k];
        if (active) {
          if (!kept(dst_any, q)) {
            st_vec<T, V>(bq + idx, v);
          } else {
#pragma unroll
            for (int k = 0; k < V; ++k)
              if (!kept(dst_m[k], q)) bq[idx + k] = v.v[k];
          }
        }
      } else if constexpr (MASKED && q != 0 && V == 1) {
        // one node per thread: its destination is the "edge" element (d_up = 1 or the wrap)
        if (active) push_edge(f[q][0], ez > 0, !kept(ez > 0 ? edge_m_up : edge_m_dn, q));
      } else if constexpr (MASKED && q != 0 && ez > 0) {
        const T up = __shfl_up_sync(0xffffffffu, f[q][V - 1], 1);
        Vec<T, V> v;
        v.v[0] = up;
#pragma unroll
        for (int k = 1; k < V; ++k) v.v[k] = f[q][k - 1];
        if (st_left && !kept(dst_any, q)) {
          st_vec<T, V>(bq + idx, v);
        } else if (active) {
#pragma unroll
          for (int k = 0; k < V; ++k)
            if ((k > 0 || has_left) && !kept(dst_m[k], q)) bq[idx + k] = v.v[k];
        }
        if (edge_up) push_edge(f[q][V - 1], true, !kept(edge_m_up, q));
      } else if constexpr (MASKED && q != 0) {
        const T dn = __shfl_down_sync(0xffffffffu, f[q][0], 1);
        Vec<T, V> v;
#pragma unroll
        for (int k = 0; k < V - 1; ++k) v.v[k] = f[q][k + 1];
        v.v[V - 1] = dn;
        if (st_right && !kept(dst_any, q)) {
          st_vec<T, V>(bq + idx, v);
        } else if (active) {
#pragma unroll
          for (int k = 0; k < V; ++k)
            if ((k < V - 1 || has_right) && !kept(dst_m[k], q)) bq[idx + k] = v.v[k];
        }
        if (edge_dn) push_edge(f[q][0], false, !kept(edge_m_dn, q));
      } else if constexpr (ez == 0) {
        Vec<T, V> v;
#pragma unroll
        for (int k = 0; k < V; ++k) v.v[k] = f[q][k];
        if (active) st_vec<T, V>(bq + idx, v);
      } else if constexpr (V == 1) {
        if (active) push_edge(f[q][0], ez > 0, true);
      } else if constexpr (ez > 0) {
        // destination z0+k takes the source value of z0+k-1: element 0 comes from the left lane
        const T up = __shfl_up_sync(0xffffffffu, f[q][V - 1], 1);
        Vec<T, V> v;
        v.v[0] = up;
#pragma unroll
        for (int k = 1; k < V; ++k) v.v[k] = f[q][k - 1];
        if (st_left) st_vec<T, V>(bq + idx, v);
#pragma unroll
        for (int k = 1; k < V; ++k)
          if (st_noleft) bq[idx + k] = f[q][k - 1];
        if (edge_up) push_edge(f[q][V - 1], true, true);
      } else {
        // ez < 0: destination z0+k takes the source value of z0+k+1
        const T dn = __shfl_down_sync(0xffffffffu, f[q][0], 1);
        Vec<T, V> v;
#pragma unroll
        for (int k = 0; k < V - 1; ++k) v.v[k] = f[q][k + 1];
        v.v[V - 1] = dn;
        if (st_right) st_vec<T, V>(bq + idx, v);
#pragma unroll
        for (int k = 0; k < V - 1; ++k)
          if (st_noright) bq[idx + k] = f[q][k + 1];
        if (edge_dn) push_edge(f[q][0], false, true);
      }
    });
    };
    if constexpr (SB) {
      if (__any_sync(0xffffffffu, (near_kept & SBITS_NEAR) != 0u)) stream_all(std::true_type{});
      else stream_all(std::false_type{});
    } else {
      stream_all(std::false_type{});
    }

    // ---- z-slab faces: the line-end threads hand their NF face populations to the neighbour rank -- staged per
    // block (packed layouts) or stored directly.  One pass for all populations, entered by the line-end lanes
    // only: the streaming loop above stays the plain kernel's.  A face element is ALWAYS sent: the receiver
    // drops it where its node keeps the population (masked streaming).
    if constexpr (HALO) {
      if (cross_up || cross_dn) {
        static_for<Q>([&](auto qc) {
          constexpr int q = decltype(qc)::value;
          constexpr int ez = S::e(q, 2);
          if constexpr (ez != 0) {
            constexpr int zf = zface_index<S>(q);
            constexpr bool up = ez > 0;
            if (up ? cross_up : cross_dn) {
              const unsigned dlines = xs[S::e(q, 0) + 1] + ys[S::e(q, 1) + 1];
              const T val = up ? f[q][V - 1] : f[q][0];
              if (staged) {
                const unsigned slot = (unsigned((up ? 0 : NF) + zf) << sshift) | lline;
                stage_val[slot] = val;
                stage_dst[slot] = dlines;
              } else {
                halo_store(up ? p.halo_hi : p.halo_lo,
                           (long long)(p.halo_compact ? zf : q) * p.halo_qs + (long long)(dlines) * p.halo_ls +
                               (up ? p.halo_hi_off : p.halo_lo_off), val);
              }
            }
          }
        });
      }
    }
    if constexpr (HALO && V > 1) {
      if (staged) {   // block-uniform
        __syncthreads();
        const unsigned t0 = blk * unsigned(STEP_BLOCK);
        const unsigned t1 = min(t0 + unsigned(STEP_BLOCK), p.n_threads);
        for (unsigned e = threadIdx.x; e < (unsigned(2 * NF) << sshift); e += STEP_BLOCK) {
          const unsigned pair = e >> sshift, j = e & ((1u << sshift) - 1u);
          const unsigned face = pair >= unsigned(NF) ? 1u : 0u, k = pair - face * unsigned(NF);
          // the thread of line first_line_blk + j that owns this face's element: its last / first lane
          const unsigned long long owner = (unsigned long long)(first_line_blk + j) * nzv + (face == 0 ? nzv - 1u : 0u);
          T* const base = (face == 0) ? p.halo_hi : p.halo_lo;
          if (owner >= t0 && owner < t1 && base != nullptr)
            halo_store(base, k * p.halo_qs + (long long)(stage_dst[e]) * p.halo_ls +
                                 (face == 0 ? p.halo_hi_off : p.halo_lo_off), stage_val[e]);
        }
      }
    }
  }

  if constexpr (OBS) {
    t_e = active ? acc_e : T(0);
    t_m2 = active ? acc_m2 : T(0);
    t_mass = active ? acc_mass : T(0);
  }
}

// One time step for the tile of this thread block + the fused reduction of its observables
template <class S, typename T, int V, int COLL, int MODE, bool OBS, bool STREAM, bool HALO, bool TICKET>
__device__ __forceinline__ void step_body(const StepParams<T>& p, const T* __restrict__ f_in, T* __restrict__ f_out,
                                          double* obs_out, T* u_out) {
  unsigned blk = blockIdx.x + p.block_rot;
  if (blk >= gridDim.x) blk -= gridDim.x;
  T acc_e = T(0), acc_m2 = T(0), acc_mass = T(0);
  step_tile<S, T, V, COLL, MODE, OBS, STREAM, HALO>(p, f_in, f_out, u_out, blk, nullptr, acc_e, acc_m2, acc_mass);
  // ---- fused observables of f_in ------------------------------------------------------------
  if constexpr (OBS) {
    if constexpr (TICKET) {
      reduce_observables<STEP_BLOCK>(double(acc_e), double(acc_m2), double(acc_mass), p.obs, obs_out);
    } else {
      // one partial per block, no fence and no ticket: with 5e5 blocks that live ~4 us each, a global
      // round trip per block on the critical path cost 18 % of the step.  obs_finish_kernel, launched
      // right behind this grid, adds the partials up (the kernel boundary makes them visible).
      double a, b, c;
      block_reduce3<STEP_BLOCK, T>(acc_e, acc_m2, acc_mass, a, b, c);
      if (threadIdx.x == 0) {
        double* bp = p.obs.block_partials + size_t(blockIdx.x) * N_OBS_PART;
        bp[0] = a; bp[1] = b; bp[2] = c;
      }
    }
  }
}

template <class S, typename T, int V, int COLL, int MODE, bool OBS, bool STREAM, bool HALO>
__global__ void __launch_bounds__(STEP_BLOCK, step_min_blocks<S, T, V, COLL, MODE>())
step_kernel(const __grid_constant__ StepParams<T> p) {
  // Programmatic dependent launch: let the next step's grid be scheduled while this one drains, and
  // wait here -- before the first global access -- until the previous step's grid has completed and
  // its stores are visible.  Both are no-ops for a launch without the PDL attribute.
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  step_body<S, T, V, COLL, MODE, OBS, STREAM, HALO, false>(p, p.f_in, p.f_out, p.obs_out, p.u_out);
}

// --------------------------------------------------------------------------------------
// Persistent variant with TMA prefetch (plain lattices: no masks, no halos).  One CTA per resident slot of the
// device walks over the tiles blockIdx.x, blockIdx.x + gridDim.x, ...; a tile's Q population chunks (each
// STEP_BLOCK * V contiguous values) are brought into shared memory by Q 1-D bulk copies (cp.async.bulk, the TMA
// engine) that complete on an mbarrier, TMA_STAGES tiles deep.  The threads read their values from shared memory,
// hand the stage back (one __syncthreads) and compute / store while the NEXT tiles' copies are in flight: the
// loads of a CTA no longer wait behind its own arithmetic, which is what bounded the register-heavy kernels
// (D3Q27 KBC: 160 registers, 3 CTAs per SM, top stall long_scoreboard at 0.91 of the HBM roofline).
// --------------------------------------------------------------------------------------
constexpr int TMA_STAGES = 2;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (bytes and both addresses multiples of 16), completion counted on `bar`
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// a stage: Q population chunks, then (MODE_NCM) the tile's no_collision_mask bytes, padded to 128 bytes
template <class S, typename T, int V, int MODE>
constexpr size_t tma_stage_bytes() {
  return (size_t(S::Q) * STEP_BLOCK * V * sizeof(T) + (MODE != MODE_PLAIN ? size_t(STEP_BLOCK) * V : 0) + 127) / 128 * 128;
}
template <class S, typename T, int V, int MODE>
constexpr size_t tma_smem_bytes() { return size_t(TMA_STAGES) * tma_stage_bytes<S, T, V, MODE>(); }

template <class S, typename T, int V, int COLL, int MODE, bool OBS>
__global__ void __launch_bounds__(STEP_BLOCK, step_min_blocks<S, T, V, COLL, MODE>())
step_tma_kernel(const __grid_constant__ StepParams<T> p) {
  static_assert(MODE == MODE_PLAIN || MODE == MODE_NCM, "no masked streaming in the TMA kernel");
  constexpr unsigned TILE = STEP_BLOCK * V;                    // nodes per tile
  constexpr unsigned TILE_BYTES = TILE * sizeof(T);            // one population chunk
  constexpr unsigned STAGE_ELEMS = unsigned(tma_stage_bytes<S, T, V, MODE>() / sizeof(T));
  constexpr bool MASK = MODE != MODE_PLAIN;
  static_assert(!MASK || (TILE % 16 == 0), "mask chunk is a bulk copy: multiple of 16 bytes");
  extern __shared__ __align__(128) unsigned char tma_smem[];
  T* const stages = reinterpret_cast<T*>(tma_smem);
  __shared__ uint64_t full[TMA_STAGES];
  asm volatile("griddepcontrol.launch_dependents;");
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < TMA_STAGES; ++s) mbar_init(&full[s], 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const unsigned n_tiles = p.n_threads / STEP_BLOCK;           // whole tiles only (host-checked)
  const size_t N = p.n_nodes;
  // warp 0 issues a tile: lane 0 arms the barrier with the byte count, then lane q issues the copy of population q
  // (one thread issuing all Q copies put ~150 serial instructions on the critical path of every tile)
  static_assert(S::Q + (MASK ? 1 : 0) <= 32, "one lane per chunk");
  auto issue = [&](unsigned tile, int s) {                     // all lanes of warp 0
    if (threadIdx.x == 0) mbar_arrive_expect_tx(&full[s], S::Q * TILE_BYTES + (MASK ? TILE : 0u));
    __syncwarp();
    if (threadIdx.x < unsigned(S::Q)) {
      const T* src = p.f_in + size_t(tile) * TILE + size_t(threadIdx.x) * N;
      T* dst = stages + size_t(s) * STAGE_ELEMS + threadIdx.x * TILE;
      bulk_copy_g2s(dst, src, TILE_BYTES, &full[s]);
    } else if (MASK && threadIdx.x == unsigned(S::Q)) {
      bulk_copy_g2s(stages + size_t(s) * STAGE_ELEMS + S::Q * TILE, p.ncm + size_t(tile) * TILE, TILE, &full[s]);
    }
  };
  if (threadIdx.x < 32u) {
#pragma unroll
    for (int s = 0; s < TMA_STAGES; ++s) {
      const unsigned tile = blockIdx.x + unsigned(s) * gridDim.x;
      if (tile < n_tiles) issue(tile, s);
    }
  }
  double e = 0.0, m2 = 0.0, mass = 0.0;
  unsigned it = 0;
  for (unsigned tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
    const int s = int(it % TMA_STAGES);
    mbar_wait(&full[s], (it / TMA_STAGES) & 1u);
    T te = T(0), tm2 = T(0), tmass = T(0);
    step_tile<S, T, V, COLL, MODE, OBS, true, false, SMEM_TMA>(
        p, p.f_in, p.f_out, p.u_out, tile, stages + size_t(s) * STAGE_ELEMS, te, tm2, tmass, [&]() {
          const unsigned next = tile + unsigned(TMA_STAGES) * gridDim.x;
          if (threadIdx.x < 32u && next < n_tiles) issue(next, s);
        });
    if constexpr (OBS) { e += double(te); m2 = fmax(m2, double(tm2)); mass += double(tmass); }
  }
  if constexpr (OBS) {   // one partial per CTA; obs_finish_kernel adds them up (fixed order: deterministic)
    block_reduce3<STEP_BLOCK>(e, m2, mass);
    if (threadIdx.x == 0) {
      double* bp = p.obs.block_partials + size_t(blockIdx.x) * N_OBS_PART;
      bp[0] = e; bp[1] = m2; bp[2] = mass;
    }
  }
}

// --------------------------------------------------------------------------------------
// Persistent variant with PER-THREAD prefetch (cp.async / LDGSTS).  For the register-heavy collisions (D3Q27 KBC:
// 160 registers, three CTAs per SM, ncu: long_scoreboard, 0.91 of the HBM roofline at 512^3) the classic kernel is
// latency-bound: a CTA's loads are in flight only while it waits for them, not while it computes and stores, and
// there are no registers for a second tile.  Here every thread copies the Q x V values of ITS OWN part of the next
// tile into its own shared-memory slots while it works on the current one -- asynchronously, without registers.
// A thread reads only what it copied itself, so there is no block barrier and no mbarrier anywhere (the TMA
// variant above runs its CTAs in lock step per tile): cp.async.wait_group orders the thread's own copies, the
// occupancy and the body are the classic kernel's.  One stage: Q * V * sizeof(T) bytes per thread.
// --------------------------------------------------------------------------------------
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gmem_src) {
  static_assert(BYTES == 4 || BYTES == 8 || BYTES == 16, "cp.async sizes");
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <class S, typename T, int V>
constexpr size_t pf_smem_bytes() { return size_t(S::Q) * STEP_BLOCK * V * sizeof(T); }

template <class S, typename T, int V, int COLL, int MODE, bool OBS>
__global__ void __launch_bounds__(STEP_BLOCK, step_min_blocks<S, T, V, COLL, MODE>())
step_pf_kernel(const __grid_constant__ StepParams<T> p) {
  static_assert(MODE == MODE_PLAIN || MODE == MODE_NCM, "no masked streaming in the prefetch kernel");
  constexpr unsigned TILE = STEP_BLOCK * V;                    // nodes per tile
  extern __shared__ __align__(128) unsigned char pf_smem[];
  T* const stage = reinterpret_cast<T*>(pf_smem);
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const unsigned n_tiles = p.n_threads / STEP_BLOCK;           // whole tiles only (host-checked)
  const size_t N = p.n_nodes;
  T* const mine = stage + threadIdx.x * V;                     // + q * TILE: this thread's V values of population q
  uint32_t mword_next = 0;
  auto prefetch = [&](unsigned tile) {
    const unsigned node = tile * TILE + threadIdx.x * V;   // < 2^31 nodes: a 32-bit index, one wide multiply-add per address
    static_for<S::Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      cp_async<int(V * sizeof(T))>(mine + q * TILE, p.in_q[q] + node);
    });
    cp_async_commit();
    if constexpr (MODE != MODE_PLAIN) mword_next = ld_mask<V>(p.ncm + node);   // needed one tile from now
  };
  // Which tiles a CTA works on (kernel-uniform):
  //   p.pf_chunk = K > 0 (the default, K = 2 .. 4, see pf_chunk()): the K consecutive tiles blockIdx.x * K ...; the grid has n_tiles / K
  //     CTAs and the hardware block scheduler balances the SMs exactly as it does for the classic kernel.  A CTA's
  //     first tile is not prefetched, the others are;
  //   p.pf_chunk = 0: persistent, one CTA per resident slot, tiles blockIdx.x + k * gridDim.x.  Measured 15 - 30 %
  //     SLOWER than the classic kernel (profiles/r2_bench_prefetch_persistent_by_stride.jsonl): the SMs do not all see
  //     the same memory bandwidth and the call ends when the slowest CTA has done its fixed share.  With the tiles
  //     handed out dynamically (atomic counter, experiment) the same kernel was 4 % FASTER than the classic one; the
  //     chunked schedule gets the same effect from the hardware scheduler, needs no counter, and its tile -> CTA map
  //     is fixed, which keeps the fused reductions bitwise reproducible.
  const unsigned chunk = p.pf_chunk;
  unsigned tile = chunk ? blockIdx.x * chunk : blockIdx.x;
  const unsigned tile_end = chunk ? min(n_tiles, tile + chunk) : n_tiles;
  const unsigned stride = chunk ? 1u : gridDim.x;
  if (tile < tile_end) prefetch(tile);
  // the thread's running (sum, max, sum) over its tiles live in shared memory, not in six registers across the tile
  // body: the D3Q27 KBC kernel with four nodes per thread sits at 255 registers
  __shared__ double s_obs[OBS ? N_OBS_PART : 1][OBS ? STEP_BLOCK : 1];
  if constexpr (OBS) { s_obs[0][threadIdx.x] = 0.0; s_obs[1][threadIdx.x] = 0.0; s_obs[2][threadIdx.x] = 0.0; }
  for (; tile < tile_end; tile += stride) {
    cp_async_wait_all();
    const uint32_t mword = mword_next;
    T te = T(0), tm2 = T(0), tmass = T(0);
    step_tile<S, T, V, COLL, MODE, OBS, true, false, SMEM_PRIVATE>(
        p, p.f_in, p.f_out, p.u_out, tile, stage, te, tm2, tmass, [&]() {
          if (tile + stride < tile_end) prefetch(tile + stride);
        }, mword);
    if constexpr (OBS) {
      s_obs[0][threadIdx.x] += double(te);
      s_obs[1][threadIdx.x] = fmax(s_obs[1][threadIdx.x], double(tm2));
      s_obs[2][threadIdx.x] += double(tmass);
    }
  }
  if constexpr (OBS) {   // one partial per CTA (fixed tiles per CTA); obs_finish_kernel adds them up in a fixed order
    double e = s_obs[0][threadIdx.x], m2 = s_obs[1][threadIdx.x], mass = s_obs[2][threadIdx.x];
    block_reduce3<STEP_BLOCK>(e, m2, mass);
    if (threadIdx.x == 0) {
      double* bp = p.obs.block_partials + size_t(blockIdx.x) * N_OBS_PART;
      bp[0] = e; bp[1] = m2; bp[2] = mass;
    }
  }
}

// Small lattices (the whole grid co-resident on the 148 SMs, state in L2): ONE cooperative launch
// advances n_loop time steps, ping-ponging f_in / f_out with a grid-wide barrier between steps, so
// a step costs a barrier (~1 us) instead of a kernel launch.  Step s writes its observables to
// obs_out + s * OBS_STRIDE; u_out belongs to the last step.  After the call the state is in f_out
// when n_loop is odd, in f_in when it is even -- the same convention as n_loop separate launches.
constexpr int OBS_STRIDE = 4;   // == LB_OBS_COUNT (include/lettuce_b200.h)

template <class S, typename T, int V, int COLL, int MODE, bool OBS>
__global__ void __launch_bounds__(STEP_BLOCK, step_min_blocks<S, T, V, COLL, MODE>())
step_loop_kernel(const __grid_constant__ StepParams<T> p, int n_loop) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  for (int s = 0; s < n_loop; ++s) {
    const T* fi = (s & 1) ? p.f_out : p.f_in;
    T* fo = (s & 1) ? const_cast<T*>(p.f_in) : p.f_out;
    step_body<S, T, V, COLL, MODE, OBS, true, false, true>(p, fi, fo, OBS ? p.obs_out + size_t(s) * OBS_STRIDE : nullptr,
                                                     (s == n_loop - 1) ? p.u_out : nullptr);
    if (s + 1 < n_loop) grid.sync();
  }
}

// --------------------------------------------------------------------------------------
// moments / standalone observables: rho [N], u [D][N] and/or the reductions of a state
// (reference/lettuce/_flow.py:136-172, observable_reporter.py:27-42)
// --------------------------------------------------------------------------------------
template <typename T>
struct MomentsParams {
  const T* f;
  T* rho_out;   // [n_nodes] or nullptr
  T* u_out;     // [D][n_nodes] or nullptr
  unsigned n_nodes;
  ObsScratch obs;
  double* obs_out;  // nullptr: no reduction
};

// V consecutive nodes per thread (16-byte loads and stores, like the step kernel): Q * V * sizeof(T) bytes in
// flight per thread instead of Q scalars
template <class S, typename T, int V, bool OBS>
__global__ void __launch_bounds__(STEP_BLOCK)
moments_kernel(const __grid_constant__ MomentsParams<T> p) {
  const unsigned n_threads = p.n_nodes / V;
  const unsigned tid = blockIdx.x * STEP_BLOCK + threadIdx.x;
  const bool active = tid < n_threads;
  const unsigned node = (active ? tid : n_threads - 1u) * V;
  const size_t N = p.n_nodes;
  T g[S::Q][V];
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    const Vec<T, V> v = ld_vec<T, V>(p.f + q * N + node);
#pragma unroll
    for (int k = 0; k < V; ++k) g[q][k] = v.v[k];
  });
  Vec<T, V> rho, u[3];
  T acc_e = T(0), acc_m2 = T(0), acc_mass = T(0);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    T h[S::Q], r, uu[3];
    static_for<S::Q>([&](auto qc) { h[decltype(qc)::value] = g[decltype(qc)::value][k]; });
    moments<S, T>(h, r, uu);
    rho.v[k] = r;
    u[0].v[k] = uu[0]; u[1].v[k] = uu[1]; u[2].v[k] = uu[2];
    if constexpr (OBS) {
      const T u2 = fma_(uu[0], uu[0], fma_(uu[1], uu[1], uu[2] * uu[2]));
      acc_e += T(0.5) * u2;
      acc_m2 = max_(acc_m2, u2);
      acc_mass += r;
    }
  }
  if (active) {
    if (p.rho_out) st_vec<T, V>(p.rho_out + node, rho);
    if (p.u_out) {
#pragma unroll
      for (int c = 0; c < S::D; ++c) st_vec<T, V>(p.u_out + c * N + node, u[S::axis(c)]);
    }
  }
  if constexpr (OBS) {
    // one partial per block, no fence and no ticket (as in the step kernels: a global round trip per block on the
    // critical path held this kernel at 0.82 of the roofline); obs_finish_kernel, launched behind this grid, adds
    // the partials up in a fixed order
    if (!active) { acc_e = T(0); acc_m2 = T(0); acc_mass = T(0); }
    double a, b, c;
    block_reduce3<STEP_BLOCK, T>(acc_e, acc_m2, acc_mass, a, b, c);
    if (threadIdx.x == 0) {
      double* bp = p.obs.block_partials + size_t(blockIdx.x) * N_OBS_PART;
      bp[0] = a; bp[1] = b; bp[2] = c;
    }
  }
}

// *obs_blocks = thread blocks that wrote a partial of the reduction (when p.obs_out is set): the caller launches
// obs_finish_kernel over them
template <class S, typename T>
cudaError_t launch_moments(const MomentsParams<T>& p, cudaStream_t st, unsigned* obs_blocks);

template <class S, typename T>
cudaError_t launch_moments_impl(const MomentsParams<T>& p, cudaStream_t st, unsigned* obs_blocks) {
  constexpr int VEC = 16 / int(sizeof(T));
  auto aligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  const bool vec = p.n_nodes % VEC == 0 && aligned(p.f) && aligned(p.rho_out) && aligned(p.u_out) &&
                   (size_t(p.n_nodes) * sizeof(T)) % 16 == 0;
  const unsigned n_threads = vec ? p.n_nodes / VEC : p.n_nodes;
  const unsigned grid = (n_threads + STEP_BLOCK - 1) / STEP_BLOCK;
  *obs_blocks = grid;
  if (vec) {
    if (p.obs_out) moments_kernel<S, T, VEC, true><<<grid, STEP_BLOCK, 0, st>>>(p);
    else moments_kernel<S, T, VEC, false><<<grid, STEP_BLOCK, 0, st>>>(p);
  } else {
    if (p.obs_out) moments_kernel<S, T, 1, true><<<grid, STEP_BLOCK, 0, st>>>(p);
    else moments_kernel<S, T, 1, false><<<grid, STEP_BLOCK, 0, st>>>(p);
  }
  return cudaGetLastError();
}

}  // namespace lb
#include "lb_resident.cuh"
namespace lb {

// --------------------------------------------------------------------------------------
// host-side dispatch (runtime descriptor -> template instantiation), one per (stencil, T)
// --------------------------------------------------------------------------------------
struct StepConfig {
  int loop_steps = 0;   // > 1: try ONE cooperative launch that advances this many steps (small lattices)
  bool looped = false;  // out: the cooperative launch was used (all loop_steps steps are enqueued)
  int vec;        // nodes per thread: 16 / 8 / 4(8) bytes
  int collision;  // COLL_*
  int mode;       // MODE_*
  bool obs;
  bool stream;
  bool halo;      // z-slab halo plumbing compiled in (multi-GPU)
  int tma = -1;   // persistent TMA-prefetch kernel: -1 = the compiled-in default, 0 = never, 1 = wherever it applies
  int pf = -1;    // persistent per-thread prefetch kernel (cp.async): the same three values
  bool resident = true;   // small 2-D lattices: the tile-resident multi-step kernel (lb_resident.cuh) where it fits
  unsigned obs_blocks = 0;   // out: thread blocks that wrote a partial of the fused reduction
};

// where the persistent TMA kernel is the default (measured, DESIGN.md section 3.1 and profiles/r2_bench_tma.jsonl):
// 2-D lattices whose populations fit in L2 -- D2Q9 fp32 2048 x 1024 (75 MB): BGK / TRT 1.07x, KBC 1.15x of the
// classic kernel.  Beyond L2 (4096^2 fp32: 0.85x; 2048 x 1024 fp64 BGK: 0.86x) and on the 3-D lattices (0.84 - 0.91x)
// the classic kernel's four (three) CTAs per SM already sit at the HBM ceiling and the persistent variant loses.
template <class S, typename T, int COLL>
bool tma_default(size_t n_nodes) {
  return S::Q == 9 && n_nodes * S::Q * sizeof(T) <= (size_t(96) << 20);
}

// one explicit specialisation per (stencil, dtype, collision), each in its own translation unit
// (lb_inst_*.cu) so that `make -j` spreads the ~400 kernel variants over all cores
template <class S, typename T, int COLL>
cudaError_t launch_step_coll(StepConfig& c, const StepParams<T>& p, cudaStream_t st);

template <class S, typename T>
cudaError_t launch_step(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  switch (c.collision) {
    case COLL_NONE: return launch_step_coll<S, T, COLL_NONE>(c, p, st);
    case COLL_BGK: return launch_step_coll<S, T, COLL_BGK>(c, p, st);
    case COLL_TRT: return launch_step_coll<S, T, COLL_TRT>(c, p, st);
    case COLL_KBC:
      if constexpr (S::ID != 1) return launch_step_coll<S, T, COLL_KBC>(c, p, st);
      else return cudaErrorNotSupported;
    case COLL_REGULARIZED: return launch_step_coll<S, T, COLL_REGULARIZED>(c, p, st);
    case COLL_SMAGORINSKY: return launch_step_coll<S, T, COLL_SMAGORINSKY>(c, p, st);
    case COLL_BGK_FORCE: return launch_step_coll<S, T, COLL_BGK_FORCE>(c, p, st);
  }
  return cudaErrorInvalidValue;
}

// Nodes per thread along the contiguous axis that are COMPILED for (stencil, dtype, collision, mode): one
// node per thread (any shape, any alignment) and one vector width -- 16 bytes, except where the register
// budget says otherwise (measured):
//   D3Q27 KBC fp64 one node per thread;
//   D3Q27 KBC / TRT / Smagorinsky fp32: FOUR nodes per thread like everything else (up to 255 registers without
//     spills, two blocks per SM) -- TRT 512^3 0.97 -> 1.03 (classic), Smagorinsky 0.98 -> 1.02 (prefetch kernel),
//     profiles/r2_bench_trt_smagorinsky_four_nodes_per_thread.jsonl.  KBC: round 1 had
//     measured two nodes faster (0.92 vs 0.82); with the packed pair arithmetic of today and the prefetch kernel
//     hiding the load latency the balance is the other way: the index / edge-store overhead per node halves, and
//     the kernel is issue- and clock-bound -- config 4 (512^3) 0.955 -> 0.995 in the same run, classic kernel
//     0.92 -> 0.94 (profiles/r2_bench_kbc_four_nodes_per_thread.jsonl);
//   D3Q27 with masked streaming: 212 registers (2 blocks per SM) with four nodes, 161 (3 blocks) with two;
//     0.89 vs 0.92 on the stock obstacle.
// Lines whose length or alignment does not fit the vector run one node per thread.
template <class S, typename T>
__host__ __device__ constexpr int vector_nodes(int coll, int mode) {
  int v = 16 / int(sizeof(T));
  if (S::Q == 27 && coll == COLL_KBC && sizeof(T) == 8) v = 1;   // (two nodes: 255 registers, classic 0.77, prefetch 0.84 vs 0.865)
  if (S::Q == 27 && sizeof(T) == 4 && mode == MODE_NCM_SBITS && v > 2) v = 2;
  return v;
}

// blocks of a kernel that fit on the device at once (cached per instantiation)
template <class K>
int coresident_blocks(K kernel) {
  static int cached = -1;
  if (cached < 0) {
    int dev = 0, sms = 0, per_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, STEP_BLOCK, 0) != cudaSuccess) {
      (void)cudaGetLastError();
      return 0;
    }
    cached = sms * per_sm;
  }
  return cached;
}

template <class S, typename T, int V, int COLL, int MODE, bool OBS>
cudaError_t launch_tma(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  auto kernel = step_tma_kernel<S, T, V, COLL, MODE, OBS>;
  constexpr size_t smem = tma_smem_bytes<S, T, V, MODE>();
  static int resident = -1;   // CTAs the device holds at once
  if (resident < 0) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, STEP_BLOCK, smem);
    if (e != cudaSuccess) return e;
    resident = sms * per_sm;
    if (resident <= 0) return cudaErrorLaunchOutOfResources;
  }
  const unsigned n_tiles = p.n_threads / STEP_BLOCK;
  const unsigned grid = n_tiles < unsigned(resident) ? n_tiles : unsigned(resident);
  c.obs_blocks = grid;
  static const bool no_pdl = getenv("LB_NO_PDL") != nullptr;   // A/B timing only
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(STEP_BLOCK);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

// The per-thread prefetch kernel is compiled for the stencil whose collisions are register-bound (D3Q27: three CTAs
// per SM or fewer) and for D2Q9, and is the default where it was measured to win (profiles/r2_bench_prefetch.jsonl,
// D3Q27 fp32 with four nodes per thread, fraction of the HBM roofline classic -> prefetch: KBC 512^3 0.94 -> 0.98 -
// 1.02, Smagorinsky 512^3 0.97 -> 1.02; TRT 1.03 -> 1.01, BGK 1.03 -> 1.01, regularized 1.00 -> 1.00 and every D3Q27
// fp64 variant -- KBC fp64, one node per thread: 0.84 -> 0.81 -- stay with the classic kernel).
template <class S, typename T, int COLL>
constexpr bool pf_compiled() { return (S::Q == 27 && COLL != COLL_NONE) || S::Q == 9; }
template <class S, typename T, int COLL>
bool pf_default(size_t n_nodes) {
  if (S::Q == 27) return sizeof(T) == 4 && (COLL == COLL_KBC || COLL == COLL_SMAGORINSKY);   // TRT: classic 1.03 vs 1.01
  // D2Q9 KBC beyond the L2-sized lattices that the TMA kernel takes (fp32 4096^2: 0.95 -> 1.03, TMA 0.80; fp64
  // 2048 x 1024: 0.84 -> 0.90, TMA 0.89)
  return S::Q == 9 && COLL == COLL_KBC && !tma_default<S, T, COLL>(n_nodes);
}
// Tiles per CTA.  More tiles per CTA = more of them prefetched, fewer CTAs = a coarser tail: measured on D3Q27 KBC
// fp32, fraction of the HBM roofline, at 256^3 (65 536 tiles, 444 resident CTAs): 2 -> 1.00, 4 -> 0.99, 8 -> 0.96,
// 16 -> 0.92; at 512^3 (524 288 tiles): 2 -> 0.92, 3 -> 0.93, 4 -> 0.94, 8 -> 0.92.  So: 4 where that still leaves
// >= 64 waves of CTAs, else 3, else 2.
inline unsigned pf_chunk(unsigned n_tiles, unsigned resident) {
  for (unsigned k = 4; k > 2; --k)
    if (n_tiles / k >= 64u * resident) return k;
  return 2;
}

template <class S, typename T, int V, int COLL, int MODE, bool OBS>
cudaError_t launch_pf(StepConfig& c, const StepParams<T>& p_in, cudaStream_t st) {
  auto kernel = step_pf_kernel<S, T, V, COLL, MODE, OBS>;
  constexpr size_t smem = pf_smem_bytes<S, T, V>();
  static int resident = -1;   // CTAs the device holds at once
  if (resident < 0) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, STEP_BLOCK, smem);
    if (e != cudaSuccess) return e;
    resident = sms * per_sm;
    if (resident <= 0) return cudaErrorLaunchOutOfResources;
  }
  static const int chunk_env = getenv("LB_PF_CHUNK") ? atoi(getenv("LB_PF_CHUNK")) : -1;   // A/B timing only
  StepParams<T> p = p_in;
  for (int q = 0; q < S::Q; ++q) {
    p.in_q[q] = p.f_in + size_t(q) * p.n_nodes;
    p.out_q[q] = p.f_out + size_t(q) * p.n_nodes;
  }
  const unsigned n_tiles = p.n_threads / STEP_BLOCK;
  p.pf_chunk = chunk_env >= 0 ? unsigned(chunk_env) : pf_chunk(n_tiles, unsigned(resident));
  unsigned grid = n_tiles < unsigned(resident) ? n_tiles : unsigned(resident);
  if (p.pf_chunk) grid = (n_tiles + p.pf_chunk - 1u) / p.pf_chunk;
  c.obs_blocks = grid;
  static const bool no_pdl = getenv("LB_NO_PDL") != nullptr;   // A/B timing only
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(STEP_BLOCK);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

// the persistent TMA kernel applies to: lattices without masked streaming, outlets and halos, made of whole tiles
// whose population (and mask) chunks are 16-byte aligned in global memory
template <typename T, int V>
bool tma_applicable(const StepConfig& c, const StepParams<T>& p) {
  return V > 1 && c.stream && !c.halo && (c.mode == MODE_PLAIN || (c.mode == MODE_NCM && !p.has_outlet)) &&
         c.loop_steps <= 1 && p.n_threads % STEP_BLOCK == 0 &&
         (size_t(p.n_nodes) * sizeof(T)) % 16 == 0 && (reinterpret_cast<uintptr_t>(p.f_in) & 15u) == 0 &&
         (reinterpret_cast<uintptr_t>(p.ncm) & 15u) == 0 && p.n_threads / STEP_BLOCK >= 4u * 148u;
}

template <class S, typename T, int V, int COLL, int MODE, bool OBS, bool STREAM, bool HALO>
cudaError_t launch_one(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  const unsigned grid = (p.n_threads + STEP_BLOCK - 1) / STEP_BLOCK;
  c.obs_blocks = grid;
  if constexpr (STREAM && !HALO && (MODE == MODE_PLAIN || MODE == MODE_NCM) && V > 1) {
    if ((c.tma == 1 || (c.tma < 0 && tma_default<S, T, COLL>(p.n_nodes))) && tma_applicable<T, V>(c, p))
      return launch_tma<S, T, V, COLL, MODE, OBS>(c, p, st);
  }
  if constexpr (STREAM && !HALO && (MODE == MODE_PLAIN || MODE == MODE_NCM) && pf_compiled<S, T, COLL>()) {
    // whole tiles, a grid of several waves, no outlet planes (they read neighbours from global memory)
    const bool applies = (c.mode == MODE_PLAIN || !p.has_outlet) && c.loop_steps <= 1 && p.n_threads % STEP_BLOCK == 0 &&
                         p.n_threads / STEP_BLOCK >= 4u * 148u &&
                         (V == 1 || ((size_t(p.n_nodes) * sizeof(T)) % 16 == 0 && (reinterpret_cast<uintptr_t>(p.f_in) & 15u) == 0));
    if ((c.pf == 1 || (c.pf < 0 && pf_default<S, T, COLL>(p.n_nodes))) && applies)
      return launch_pf<S, T, V, COLL, MODE, OBS>(c, p, st);
  }
  if (c.loop_steps > 1) {
    // small 2-D lattice without masks and reductions: the state stays in shared memory for several steps
    if constexpr (STREAM && !HALO && MODE == MODE_PLAIN && !OBS && S::D == 2) {
      if (c.resident) {
        bool done = false;
        const cudaError_t e = launch_resident<S, T, COLL>(const_cast<T*>(p.f_in), p.f_out, p.ny, p.nz, p.cp, c.loop_steps,
                                                          st, &done);
        if (e != cudaSuccess) return e;
        if (done) { c.looped = true; return cudaSuccess; }
      }
    }
    // request for ONE launch covering loop_steps steps: honoured if the grid is co-resident, otherwise
    // nothing is launched and the caller (c.looped == false) issues one launch per step
    if constexpr (STREAM && !HALO && V > 1) {
      auto kernel = step_loop_kernel<S, T, V, COLL, MODE, OBS>;
      if (int(grid) <= coresident_blocks(kernel)) {
        int n_loop = c.loop_steps;
        void* args[] = {const_cast<StepParams<T>*>(&p), &n_loop};
        const cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3(grid), dim3(STEP_BLOCK),
                                                          args, 0, st);
        if (e == cudaSuccess) c.looped = true;
        else (void)cudaGetLastError();   // e.g. cooperative launch not supported
      }
    }
    return cudaSuccess;
  }
  static const bool no_pdl = getenv("LB_NO_PDL") != nullptr;   // A/B timing only
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(STEP_BLOCK);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, step_kernel<S, T, V, COLL, MODE, OBS, STREAM, HALO>, p);
}

template <class S, typename T, int V, int COLL, int MODE>
cudaError_t launch_mode(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  if (!c.stream) {
    if constexpr (V == 1 && MODE != MODE_NCM_SBITS)
      return launch_one<S, T, V, COLL, MODE, false, false, false>(c, p, st);
    else
      return cudaErrorInvalidValue;
  }
  if (c.halo) {
    if (c.obs) return launch_one<S, T, V, COLL, MODE, true, true, true>(c, p, st);
    return launch_one<S, T, V, COLL, MODE, false, true, true>(c, p, st);
  }
  if (c.obs) return launch_one<S, T, V, COLL, MODE, true, true, false>(c, p, st);
  return launch_one<S, T, V, COLL, MODE, false, true, false>(c, p, st);
}

template <class S, typename T, int COLL, int MODE>
cudaError_t launch_vec(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  constexpr int VEC = vector_nodes<S, T>(COLL, MODE);
  if (c.vec == 1) return launch_mode<S, T, 1, COLL, MODE>(c, p, st);
  if constexpr (VEC > 1) {
    if (c.vec == VEC) return launch_mode<S, T, VEC, COLL, MODE>(c, p, st);
  }
  return cudaErrorInvalidValue;
}

template <class S, typename T, int COLL>
cudaError_t launch_step_coll_impl(StepConfig& c, const StepParams<T>& p, cudaStream_t st) {
  switch (c.mode) {
    case MODE_PLAIN: return launch_vec<S, T, COLL, MODE_PLAIN>(c, p, st);
    case MODE_NCM: return launch_vec<S, T, COLL, MODE_NCM>(c, p, st);
    case MODE_NCM_SBITS: return launch_vec<S, T, COLL, MODE_NCM_SBITS>(c, p, st);
  }
  return cudaErrorInvalidValue;
}

#if defined(LB_STUB)   // development builds (Makefile, DEV=...): this unit's kernels are left out
#define LB_INSTANTIATE_COLL(S, T, COLL)                                                   \
  template <>                                                                             \
  cudaError_t lb::launch_step_coll<S, T, COLL>(StepConfig&, const StepParams<T>&, cudaStream_t) { \
    return cudaErrorNotSupported;                                                         \
  }
#else
#define LB_INSTANTIATE_COLL(S, T, COLL)                                                   \
  template <>                                                                             \
  cudaError_t lb::launch_step_coll<S, T, COLL>(StepConfig& c, const StepParams<T>& p,       \
                                               cudaStream_t st) {                         \
    return launch_step_coll_impl<S, T, COLL>(c, p, st);                                   \
  }
#endif

#define LB_INSTANTIATE_MOMENTS(S, T)                                                      \
  template <>                                                                             \
  cudaError_t lb::launch_moments<S, T>(const MomentsParams<T>& p, cudaStream_t st, unsigned* obs_blocks) { \
    return launch_moments_impl<S, T>(p, st, obs_blocks);                                  \
  }

}  // namespace lb
