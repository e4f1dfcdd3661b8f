// C ABI of the engine (include/lettuce_b200.h) and the small stencil-independent kernels
// (mask packing, enstrophy).  Translates the runtime descriptor into the template
// instantiations of lb_kernels.cuh; launches on the caller's stream, never synchronises
// (except lb_run_host, which is a host-buffer convenience entry).
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/lettuce_b200.h"
#include "lb_kernels.cuh"
#include "lb_halo.cuh"
#include "lb_output.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define LB_CUDA(call)                                                                 \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess)                                                            \
      return fail(LB_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                  __FILE__, __LINE__);                                                \
  } while (0)

struct Geometry {
  unsigned nx, ny, nz;  // kernel axes (Z contiguous); 2-D runs as nx = 1
  unsigned n_nodes;
  int q, d;
};

int stencil_q(int s) { return s == LB_D2Q9 ? 9 : s == LB_D3Q19 ? 19 : s == LB_D3Q27 ? 27 : -1; }
int stencil_d(int s) { return s == LB_D2Q9 ? 2 : (s == LB_D3Q19 || s == LB_D3Q27) ? 3 : -1; }

int check_desc(const lb_desc* d, Geometry* g) {
  if (!d) return fail(LB_EINVAL, "null descriptor");
  const int q = stencil_q(d->stencil), dim = stencil_d(d->stencil);
  if (q < 0) return fail(LB_EINVAL, "unknown stencil id %d", d->stencil);
  if (d->ndim != dim) return fail(LB_EINVAL, "stencil is %d-D but ndim = %d", dim, d->ndim);
  if (d->dtype != LB_F32 && d->dtype != LB_F64) return fail(LB_EINVAL, "unknown dtype %d", d->dtype);
  if (d->collision < LB_COLL_NONE || d->collision > LB_COLL_SMAGORINSKY)
    return fail(LB_EINVAL, "unknown collision id %d", d->collision);
  if (d->collision == LB_COLL_KBC && d->stencil == LB_D3Q19)
    return fail(LB_EUNSUPPORTED, "KBC collision is only implemented for D2Q9 and D3Q27");
  long long n = 1;
  for (int a = 0; a < dim; ++a) {
    if (d->dims[a] < 1) return fail(LB_EINVAL, "dims[%d] = %lld", a, (long long)d->dims[a]);
    n *= d->dims[a];
    if (n >= (1ll << 31)) return fail(LB_EINVAL, "more than 2^31-1 nodes per device");
  }
  if (d->collision != LB_COLL_NONE && !(d->tau > 0.0)) return fail(LB_EINVAL, "tau must be > 0");
  if (d->collision == LB_COLL_TRT && !(d->tau_minus > 0.0)) return fail(LB_EINVAL, "tau_minus must be > 0");
  if (d->collision == LB_COLL_SMAGORINSKY && !(d->tau_minus >= 0.0))
    return fail(LB_EINVAL, "the Smagorinsky constant (tau_minus) must be >= 0");
  if (d->force < LB_FORCE_NONE || d->force > LB_FORCE_SHAN_CHEN) return fail(LB_EINVAL, "unknown force id %d", d->force);
  if (d->force != LB_FORCE_NONE) {
    if (d->collision != LB_COLL_BGK)
      return fail(LB_EUNSUPPORTED, "forcing schemes are implemented for BGKCollision only (collision id %d)", d->collision);
    if (!(d->force_tau > 0.0)) return fail(LB_EINVAL, "force_tau must be > 0");
  }
  g->q = q;
  g->d = dim;
  if (dim == 2) { g->nx = 1; g->ny = (unsigned)d->dims[0]; g->nz = (unsigned)d->dims[1]; }
  else { g->nx = (unsigned)d->dims[0]; g->ny = (unsigned)d->dims[1]; g->nz = (unsigned)d->dims[2]; }
  g->n_nodes = (unsigned)n;
  return LB_OK;
}

template <typename T>
lb::CollisionParams<T> make_cp(const lb_desc* d) {
  lb::CollisionParams<T> cp{};
  if (d->collision == LB_COLL_NONE) return cp;
  cp.omega = T(1.0 / d->tau);
  cp.inv_2tau_p = T(1.0 / (2.0 * d->tau));
  cp.inv_2tau_m = T(1.0 / (2.0 * (d->collision == LB_COLL_TRT ? d->tau_minus : 1.0)));
  const double beta = 1. / (2 * d->tau);
  cp.beta = T(beta);
  cp.inv_beta = T(1. / beta);
  cp.reg_c = T(1. - 1. / d->tau);
  cp.smag_tau = T(d->tau);
  cp.smag_nu = T((d->tau - 0.5) / 3.0);
  cp.smag_c2 = T(d->tau_minus * d->tau_minus);   // LB_COLL_SMAGORINSKY: tau_minus = the Smagorinsky constant
  if (d->force != LB_FORCE_NONE) {
    // ueq_scaling_factor (guo.py:35-37: 0.5; shan_chen.py:21-23: tau * 1) times the acceleration, and the
    // prefactor of the source term (guo.py:29: 1 - 1 / (2 tau); shan_chen.py:13-14: no source);
    // lattice axes -> kernel axes (2-D runs as X = 1)
    const double scale = d->force == LB_FORCE_GUO ? 0.5 : d->force_tau * 1;
    const int dim = stencil_d(d->stencil);
    for (int a = 0; a < dim; ++a) {
      const int ka = (dim == 2) ? a + 1 : a;
      cp.force_a[ka] = T(d->acceleration[a]);
      cp.force_ueq[ka] = T(scale * d->acceleration[a]);
    }
    cp.force_src = d->force == LB_FORCE_GUO ? T(1 - 1 / (2 * d->force_tau)) : T(0);
  }
  return cp;
}

bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

// the largest grid any kernel launches over this lattice: one node per thread
size_t max_blocks(const Geometry& g) { return (size_t(g.n_nodes) + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK; }

// lay out the reduction scratch and zero its ticket counters
int obs_prepare(const Geometry& g, void* scratch, lb::ObsScratch* out, cudaStream_t st) {
  *out = lb::obs_layout(scratch, max_blocks(g));
  LB_CUDA(cudaMemsetAsync(scratch, 0, lb::obs_counter_bytes(max_blocks(g)), st));
  return LB_OK;
}

// nodes per thread along the contiguous axis: the vector width compiled for this (stencil, dtype, collision,
// mode) -- lb::vector_nodes -- when it divides the line and matches the buffers' alignment, else one.
template <class S, typename T>
int vector_width(int coll, int mode, const Geometry& g, const void* a, const void* b, const void* ncm, const void* sbits,
                 bool scalar_only) {
  if (scalar_only) return 1;
  const int v = lb::vector_nodes<S, T>(coll, mode);
  static const bool force_scalar = getenv("LB_VEC") != nullptr && atoi(getenv("LB_VEC")) == 1;   // A/B timing only
  if (v > 1 && !force_scalar && g.nz % unsigned(v) == 0 && aligned_to(a, v * sizeof(T)) && aligned_to(b, v * sizeof(T)) &&
      aligned_to(ncm, v) && aligned_to(sbits, v * sizeof(uint32_t)))
    return v;
  return 1;
}

// Second stage of the step kernel's fused reduction: n_blocks per-block partials -> out[0..n_out).
// OBS_FINISH_CHUNKS blocks each sum a contiguous chunk in a fixed order; the last to finish (ticket;
// these few blocks are all resident, nothing waits on them) sums the chunk results.  Deterministic:
// the chunking depends on n_blocks only.
constexpr int OBS_FINISH_CHUNKS = 128;
constexpr int OBS_FINISH_BLOCK = 256;
constexpr int OBS_FINISH_SINGLE = 4096;

__global__ void __launch_bounds__(OBS_FINISH_BLOCK)
obs_finish_kernel(lb::ObsScratch sc, unsigned n_blocks, double* out, int n_out) {
  // programmatic dependent launch, as in the step kernels: this grid sits between two time steps of a call with an
  // every-step reporter, and without it each of its two kernel boundaries cost a launch latency (measured: + 11 us
  // per step, + 50 % on a 2048 x 1024 lattice).  The next step's blocks may become resident now; the partials are
  // read only once the step grid that wrote them has completed.
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __shared__ bool is_last;
  const unsigned per = (n_blocks + gridDim.x - 1) / gridDim.x;
  const unsigned lo = blockIdx.x * per, hi = min(n_blocks, lo + per);
  double a = 0.0, b = 0.0, c = 0.0;
  for (unsigned i = lo + threadIdx.x; i < hi; i += OBS_FINISH_BLOCK) {
    const double* bp = sc.block_partials + size_t(i) * lb::N_OBS_PART;
    a += bp[0];
    b = fmax(b, bp[1]);
    c += bp[2];
  }
  lb::block_reduce3<OBS_FINISH_BLOCK>(a, b, c);
  if (gridDim.x == 1) {   // few partials (persistent kernels, 2-D lattices): one block, no fence, no ticket
    if (threadIdx.x == 0) {
      out[0] = a;
      if (n_out > 1) out[1] = b;
      if (n_out > 2) out[2] = c;
    }
    return;
  }
  if (threadIdx.x == 0) {
    double* gp = sc.group_partials + size_t(blockIdx.x) * lb::N_OBS_PART;
    gp[0] = a; gp[1] = b; gp[2] = c;
    __threadfence();
    is_last = (atomicAdd(sc.global_counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  a = 0.0; b = 0.0; c = 0.0;
  if (threadIdx.x < gridDim.x) {
    const double* gp = sc.group_partials + size_t(threadIdx.x) * lb::N_OBS_PART;
    a = __ldcg(gp + 0); b = __ldcg(gp + 1); c = __ldcg(gp + 2);
  }
  lb::block_reduce3<OBS_FINISH_BLOCK>(a, b, c);
  if (threadIdx.x == 0) {
    out[0] = a;
    if (n_out > 1) out[1] = b;
    if (n_out > 2) out[2] = c;
    *sc.global_counter = 0u;
  }
}

cudaError_t launch_obs_finish(const lb::ObsScratch& sc, unsigned n_blocks, double* out, int n_out, cudaStream_t st) {
  // up to OBS_FINISH_SINGLE partials: ONE block (measured: the three-block ticket version took 5 us to add up the 740
  // partials of a persistent kernel -- two fences and an atomic round trip -- between two 22 us time steps)
  const unsigned chunks = n_blocks <= unsigned(OBS_FINISH_SINGLE) ? 1u : (n_blocks + OBS_FINISH_BLOCK - 1) / OBS_FINISH_BLOCK;
  const unsigned grid = chunks < unsigned(OBS_FINISH_CHUNKS) ? (chunks ? chunks : 1u) : unsigned(OBS_FINISH_CHUNKS);
  static const bool no_pdl = getenv("LB_NO_PDL") != nullptr;   // A/B timing only
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(OBS_FINISH_BLOCK);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, obs_finish_kernel, sc, n_blocks, out, n_out);
}

template <typename T>
int fill_boundaries(const lb_desc* d, const Geometry& g, const lb_boundary* b, int nb,
                    lb::StepParams<T>* p) {
  if (nb < 0 || nb > LB_MAX_BOUNDARIES) return fail(LB_EINVAL, "n_boundaries = %d (max %d)", nb, LB_MAX_BOUNDARIES);
  if (nb > 0 && !b) return fail(LB_EINVAL, "boundaries is null");
  p->n_bnd = nb;
  for (int i = 0; i < nb; ++i) {
    lb::BoundaryDev<T>& o = p->bnd[i];
    o.kind = b[i].kind;
    o.rho = static_cast<const T*>(b[i].rho);
    o.u = static_cast<const T*>(b[i].u);
    o.ucs = b[i].u_comp_stride;
    if (o.kind < LB_BND_BOUNCE_BACK || o.kind > LB_BND_EQ_OUTLET_P)
      return fail(LB_EINVAL, "boundary %d: unknown kind %d", i, o.kind);
    o.axis = 0; o.sign = 0; o.value = T(b[i].value);
    if (o.kind == LB_BND_ABB_OUTLET || o.kind == LB_BND_EQ_OUTLET_P) {
      const int dir = b[i].direction, la = (dir < 0 ? -dir : dir) - 1;   // lattice axis
      if (dir == 0 || la >= g.d) return fail(LB_EINVAL, "boundary %d: outlet direction %d on a %d-D lattice", i, dir, g.d);
      o.axis = (g.d == 2) ? la + 1 : la;   // lattice axis -> kernel axis (2-D runs as X = 1)
      o.sign = dir > 0 ? 1 : -1;
      const unsigned n = o.axis == 0 ? g.nx : (o.axis == 1 ? g.ny : g.nz);
      if (n < 2) return fail(LB_EINVAL, "boundary %d: the outlet needs at least 2 planes along its axis", i);
      if (!p->ncm) return fail(LB_EINVAL, "boundary %d: outlet boundaries need no_collision_mask", i);
      // the kernel takes the velocity next to the plane from the neighbour's INCOMING populations, which is
      // only right when the collision conserves momentum; a forced collision adds the force's increment
      // (the reference's outlets read flow.u() after it, anti_bounce_back_outlet.py:71-86)
      if (d->force != LB_FORCE_NONE)
        return fail(LB_EUNSUPPORTED, "boundary %d: outlet boundaries are not supported together with a forcing scheme", i);
      p->has_outlet = 1;
    }
    if (o.kind == LB_BND_EQUILIBRIUM && (!o.rho || !o.u))
      return fail(LB_EINVAL, "boundary %d: equilibrium boundary needs rho and u", i);
    // lattice axes -> kernel axes (2-D: X unused)
    for (int a = 0; a < 3; ++a) { o.rs[a] = 0; o.us[a] = 0; }
    for (int a = 0; a < g.d; ++a) {
      const int ka = (g.d == 2) ? a + 1 : a;
      o.rs[ka] = b[i].rho_stride[a];
      o.us[ka] = b[i].u_stride[a];
    }
  }
  return LB_OK;
}

template <class S, typename T>
int step_typed(const lb_desc* d, const Geometry& g, const lb_fields* fl, int n_steps, bool do_stream,
               cudaStream_t st) {
  lb::StepParams<T> p{};
  p.nx = g.nx; p.ny = g.ny; p.nz = g.nz; p.n_nodes = g.n_nodes;
  p.cp = make_cp<T>(d);
  p.ncm = fl->no_collision_mask;
  p.sbits = fl->streaming_bits;
  int rc = fill_boundaries<T>(d, g, fl->boundaries, fl->n_boundaries, &p);
  if (rc) return rc;
  if (p.sbits && !p.ncm) return fail(LB_EINVAL, "streaming_bits requires no_collision_mask (lettuce creates both, _simulation.py:71-75)");
  // an outlet node reads ALL populations of the node next to it from f_in: in place, other threads would be
  // overwriting them meanwhile
  if (p.has_outlet && fl->f == fl->f_next)
    return fail(LB_EINVAL, "outlet boundaries read neighbouring nodes: f_out must not alias f_in");

  lb::StepConfig c{};
  c.collision = (d->collision == LB_COLL_BGK && d->force != LB_FORCE_NONE) ? int(lb::COLL_BGK_FORCE) : d->collision;
  c.mode = p.sbits ? lb::MODE_NCM_SBITS : (p.ncm ? lb::MODE_NCM : lb::MODE_PLAIN);
  c.stream = do_stream;
  c.halo = fl->halo != nullptr;
  static const char* tma_env = getenv("LB_TMA");   // A/B timing: 0 = never, 1 = wherever the TMA kernel applies
  if (tma_env) c.tma = atoi(tma_env) != 0 ? 1 : 0;
  if (d->kernel_flags & LB_KF_NO_TMA) c.tma = 0;
  if (d->kernel_flags & LB_KF_FORCE_TMA) c.tma = 1;
  static const char* pf_env = getenv("LB_PF");     // A/B timing: the per-thread prefetch kernel, 0 = never, 1 = wherever
  if (pf_env) c.pf = atoi(pf_env) != 0 ? 1 : 0;
  if (d->kernel_flags & LB_KF_NO_PREFETCH) c.pf = 0;
  if (d->kernel_flags & LB_KF_FORCE_PREFETCH) { c.pf = 1; c.tma = 0; }
  static const bool no_resident = getenv("LB_NO_RESIDENT") != nullptr;   // A/B timing only
  c.resident = !no_resident && !(d->kernel_flags & LB_KF_NO_RESIDENT);
  c.vec = do_stream ? vector_width<S, T>(c.collision, c.mode, g, fl->f, fl->f_next, p.ncm, p.sbits,
                                            p.has_outlet && p.sbits == nullptr) : 1;
  p.nzv = g.nz / c.vec;
  p.n_threads = g.nx * g.ny * p.nzv;
  // an outlet on the far side of the slowest axis: its plane is the last tiles of the grid -> run them first
  for (int i = 0; i < p.n_bnd && do_stream; ++i) {
    const lb::BoundaryDev<T>& o = p.bnd[i];
    const bool outlet = o.kind == LB_BND_ABB_OUTLET || o.kind == LB_BND_EQ_OUTLET_P;
    const int slowest = g.nx > 1 ? 0 : 1;
    if (outlet && o.sign > 0 && o.axis == slowest) {
      const unsigned grid = (p.n_threads + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK;
      const unsigned plane_threads = p.n_threads / (slowest == 0 ? g.nx : g.ny);
      const unsigned plane_blocks = plane_threads / lb::STEP_BLOCK + 2;
      if (plane_blocks < grid) p.block_rot = grid - plane_blocks;
    }
  }

  if (fl->obs_out) {
    if (!fl->obs_scratch) return fail(LB_EINVAL, "obs_out given without obs_scratch");
    rc = obs_prepare(g, fl->obs_scratch, &p.obs, st);
    if (rc) return rc;
  }
  const lb_halo* h = fl->halo;
  bool halo_sync = false;
  const size_t face_elems = size_t(lb::zface_count<S>()) * g.nx * g.ny;
  if (h) {
    if (!do_stream) return fail(LB_EINVAL, "halo given to a collide-only call");
    if (p.sbits && h->mode == LB_HALO_DIRECT)
      return fail(LB_EUNSUPPORTED, "masked streaming (no_streaming_mask) across slabs needs a packed halo transport: "
                                   "LB_HALO_DIRECT stores into planes whose no-streaming bits the sender does not have");
    static const bool debug_nosync = getenv("LB_DEBUG_HALO_NOSYNC") != nullptr;  // timing experiments only
    switch (h->mode) {
      case LB_HALO_PACKED:
        if (n_steps > 1) return fail(LB_EINVAL, "LB_HALO_PACKED advances one step per call");
        if (!h->send_hi[0] || !h->send_lo[0]) return fail(LB_EINVAL, "halo send buffer is null");
        break;
      case LB_HALO_PUSH:
        for (int k = 0; k < 2; ++k)
          if (!h->send_hi[k] || !h->send_lo[k] || !h->recv_lo[k] || !h->recv_hi[k] ||
              !h->peer_up_recv_lo[k] || !h->peer_down_recv_hi[k])
            return fail(LB_EINVAL, "LB_HALO_PUSH needs both parities of every halo buffer");
        if (!h->wait_slots || !h->signal_up || !h->signal_down) return fail(LB_EINVAL, "LB_HALO_PUSH needs the handshake slots");
        halo_sync = !debug_nosync;
        break;
      case LB_HALO_REMOTE:
        for (int k = 0; k < 2; ++k)
          if (!h->recv_lo[k] || !h->recv_hi[k] || !h->peer_up_recv_lo[k] || !h->peer_down_recv_hi[k])
            return fail(LB_EINVAL, "LB_HALO_REMOTE needs both parities of the receive buffers, local and peer-mapped");
        if (!h->wait_slots || !h->signal_up || !h->signal_down) return fail(LB_EINVAL, "LB_HALO_REMOTE needs the handshake slots");
        halo_sync = !debug_nosync;
        break;
      case LB_HALO_LL:
        for (int k = 0; k < 3; ++k)
          if (!h->recv_lo[k] || !h->recv_hi[k] || !h->peer_up_recv_lo[k] || !h->peer_down_recv_hi[k])
            return fail(LB_EINVAL, "LB_HALO_LL needs three receive buffers per face, local and peer-mapped");
        if (!h->wait_slots) return fail(LB_EINVAL, "LB_HALO_LL needs wait_slots (error flag)");
        p.halo_ll = 1;
        p.ll_err = h->wait_slots + 2;
        break;
      case LB_HALO_DIRECT:
        if (!h->peer_up_f[0] || !h->peer_down_f[0] || (n_steps > 1 && (!h->peer_up_f[1] || !h->peer_down_f[1])))
          return fail(LB_EINVAL, "LB_HALO_DIRECT needs the neighbours' population buffers");
        if (!h->wait_slots || !h->signal_up || !h->signal_down) return fail(LB_EINVAL, "LB_HALO_DIRECT needs the handshake slots");
        halo_sync = !debug_nosync;
        break;
      default:
        return fail(LB_EINVAL, "unknown halo mode %d", h->mode);
    }
    if (h->mode == LB_HALO_DIRECT) {
      p.halo_qs = g.n_nodes; p.halo_ls = g.nz; p.halo_hi_off = 0; p.halo_lo_off = g.nz - 1; p.halo_compact = 0;
    } else {
      p.halo_qs = (long long)g.nx * g.ny; p.halo_ls = 1; p.halo_hi_off = 0; p.halo_lo_off = 0; p.halo_compact = 1;
      // packed layout: stage the z-face elements per block (lb_kernels.cuh).  A block of STEP_BLOCK threads
      // touches at most STEP_BLOCK / nzv + 2 lines; slots per (face, population) = next power of two.
      static const bool no_stage = getenv("LB_HALO_NO_STAGE") != nullptr;   // A/B timing only
      const unsigned lines_per_block = lb::STEP_BLOCK / p.nzv + 2;
      if (c.vec > 1 && lines_per_block <= unsigned(lb::HALO_STAGE_LINES) && !no_stage) {
        int sh = 1;
        while ((1u << sh) < lines_per_block) ++sh;
        p.halo_stage_shift = sh;
      }
    }
    if (halo_sync) {  // everybody has entered the call: the neighbours' buffers are ours to write
      rc = lb::launch_halo_push(nullptr, nullptr, nullptr, nullptr, 0, h->signal_up, h->signal_down,
                                h->wait_slots, h->epoch + 1u, st);
      if (rc) return fail(LB_ECUDA, "halo sync launch failed");
    }
  }
  T* a = static_cast<T*>(fl->f);
  T* b = static_cast<T*>(fl->f_next);
  static_assert(lb::OBS_STRIDE == LB_OBS_COUNT, "observable record size");
  if (!h && do_stream && n_steps > 1) {
    // small lattice: all n_steps in ONE cooperative launch (grid-wide barrier between steps);
    // launch_one falls back to the per-step path when the grid is not co-resident
    static const bool no_loop = getenv("LB_NO_LOOP") != nullptr;   // A/B timing only
    if (!no_loop && !(d->kernel_flags & LB_KF_NO_LOOP)) {
      c.loop_steps = n_steps;
      c.obs = fl->obs_out != nullptr;
      p.f_in = a; p.f_out = b;
      p.obs_out = fl->obs_out;
      p.u_out = c.obs ? static_cast<T*>(fl->u_out) : nullptr;
      const unsigned grid = (p.n_threads + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK;
      if (grid <= 148u * 16u) {   // cannot be co-resident beyond that; skips the occupancy query
        cudaError_t e = lb::launch_step<S, T>(c, p, st);   // launches nothing unless it can loop
        if (e != cudaSuccess) return fail(LB_ECUDA, "step kernel launch failed: %s", cudaGetErrorString(e));
        if (c.looped) return LB_OK;
      }
      c.loop_steps = 0;
    }
  }
  for (int s = 0; s < n_steps; ++s) {
    if (h) {
      const int k = s & 1;
      if (h->mode == LB_HALO_DIRECT) {
        p.halo_hi = static_cast<T*>(h->peer_up_f[k]); p.halo_lo = static_cast<T*>(h->peer_down_f[k]);
      } else {
        const int ks = h->mode == LB_HALO_PACKED ? 0 : k;
        if (h->mode == LB_HALO_LL) {   // tagged words, slot = global step index mod 3
          const uint32_t gs = h->epoch + uint32_t(s) + 1u;
          p.halo_hi = static_cast<T*>(h->peer_up_recv_lo[gs % 3u]); p.halo_lo = static_cast<T*>(h->peer_down_recv_hi[gs % 3u]);
          p.ll_tag_out = gs;
          if (s > 0 || (h->flags & LB_HALO_F_CONSUME_FIRST)) {
            p.halo_in_lo = static_cast<const T*>(h->recv_lo[(gs - 1u) % 3u]);
            p.halo_in_hi = static_cast<const T*>(h->recv_hi[(gs - 1u) % 3u]);
            p.ll_tag_in = gs - 1u;
          }
        } else if (h->mode == LB_HALO_REMOTE) {   // straight into the neighbours' receive buffers
          p.halo_hi = static_cast<T*>(h->peer_up_recv_lo[k]); p.halo_lo = static_cast<T*>(h->peer_down_recv_hi[k]);
        } else {
          p.halo_hi = static_cast<T*>(h->send_hi[ks]); p.halo_lo = static_cast<T*>(h->send_lo[ks]);
        }
        if ((h->mode == LB_HALO_PUSH || h->mode == LB_HALO_REMOTE) && s > 0) {   // consume what the previous step delivered
          p.halo_in_lo = static_cast<const T*>(h->recv_lo[k ^ 1]);
          p.halo_in_hi = static_cast<const T*>(h->recv_hi[k ^ 1]);
        } else if (h->mode == LB_HALO_PACKED) {   // optional: a halo received but not yet unpacked
          p.halo_in_lo = static_cast<const T*>(h->recv_lo[0]);
          p.halo_in_hi = static_cast<const T*>(h->recv_hi[0]);
        }
      }
    }
    p.f_in = a;
    p.f_out = b;
    c.obs = fl->obs_out != nullptr;
    p.obs_out = c.obs ? fl->obs_out + (size_t)s * LB_OBS_COUNT : nullptr;
    p.u_out = (c.obs && s == n_steps - 1) ? static_cast<T*>(fl->u_out) : nullptr;
    cudaError_t e = lb::launch_step<S, T>(c, p, st);
    if (e != cudaSuccess) return fail(LB_ECUDA, "step kernel launch failed: %s", cudaGetErrorString(e));
    if (c.obs) {   // second stage of the fused reduction: block partials -> obs_out[s]
      e = launch_obs_finish(p.obs, c.obs_blocks, p.obs_out, lb::N_OBS_PART, st);
      if (e != cudaSuccess) return fail(LB_ECUDA, "observable reduction launch failed: %s", cudaGetErrorString(e));
    }
    if (h && h->mode != LB_HALO_PACKED && h->mode != LB_HALO_LL) {
      const int k = s & 1;
      const bool push = h->mode == LB_HALO_PUSH;
      if (push || halo_sync) {
        rc = lb::launch_halo_push(push ? h->send_hi[k] : nullptr, push ? h->peer_up_recv_lo[k] : nullptr,
                                  push ? h->send_lo[k] : nullptr, push ? h->peer_down_recv_hi[k] : nullptr,
                                  push ? face_elems * sizeof(T) : 0, halo_sync ? h->signal_up : nullptr,
                                  halo_sync ? h->signal_down : nullptr, h->wait_slots,
                                  h->epoch + 2u + (unsigned)s, st);
        if (rc) return fail(LB_ECUDA, "halo push launch failed");
      }
    }
    T* t = a; a = b; b = t;
  }
  if (h && (h->mode == LB_HALO_PUSH || h->mode == LB_HALO_REMOTE) && n_steps > 0) {
    // make the population buffer canonical again: scatter the last received halo into it
    const int k = (n_steps - 1) & 1;
    rc = lb_unpack_halo_ex(d, a, h->recv_lo[k], h->recv_hi[k], p.sbits, 0u, nullptr, st);
    if (rc) return fail(LB_ECUDA, "halo unpack launch failed");
  }
  if (h && h->mode == LB_HALO_LL && n_steps > 0 && !(h->flags & LB_HALO_F_NO_UNPACK)) {
    // the same for the LL transport: this is the one place that WAITS for the neighbours (their last step)
    const uint32_t gs = h->epoch + uint32_t(n_steps);
    rc = lb_unpack_halo_ex(d, a, h->recv_lo[gs % 3u], h->recv_hi[gs % 3u], p.sbits, gs, h->wait_slots + 2, st);
    if (rc) return fail(LB_ECUDA, "halo unpack launch failed");
  }
  return LB_OK;
}

template <typename T>
int step_stencil(const lb_desc* d, const Geometry& g, const lb_fields* fl, int n, bool s, cudaStream_t st) {
  switch (d->stencil) {
    case LB_D2Q9: return step_typed<lb::D2Q9, T>(d, g, fl, n, s, st);
    case LB_D3Q19: return step_typed<lb::D3Q19, T>(d, g, fl, n, s, st);
    case LB_D3Q27: return step_typed<lb::D3Q27, T>(d, g, fl, n, s, st);
  }
  return fail(LB_EINVAL, "bad stencil");
}

int step_any(const lb_desc* d, const lb_fields* fl, int n_steps, bool do_stream, void* stream) {
  Geometry g;
  int rc = check_desc(d, &g);
  if (rc) return rc;
  if (!fl || !fl->f || !fl->f_next) return fail(LB_EINVAL, "null population buffer");
  if (fl->f == fl->f_next && do_stream) return fail(LB_EINVAL, "f and f_next must be distinct buffers");
  if (n_steps < 0) return fail(LB_EINVAL, "n_steps = %d", n_steps);
  if (fl->u_out && !fl->obs_out) return fail(LB_EINVAL, "u_out requires obs_out");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return d->dtype == LB_F32 ? step_stencil<float>(d, g, fl, n_steps, do_stream, st)
                            : step_stencil<double>(d, g, fl, n_steps, do_stream, st);
}

// ---- bits[x] = OR_q (nsm[q][x] == 1) << q ------------------------------------------------
__global__ void pack_sbits_kernel(const uint8_t* __restrict__ nsm, uint32_t* __restrict__ bits,
                                  unsigned n_nodes, int q) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  uint32_t w = 0;
  for (int k = 0; k < q; ++k) w |= (nsm[(size_t)k * n_nodes + i] == 1 ? 1u : 0u) << k;  // torch.eq(nsm, 1), _simulation.py:173
  bits[i] = w;
}

// Second pass: bit 31 (LB_SBITS_NEAR) of a node's word = "this node or one of its 26 periodic neighbours
// keeps a population".  The step kernel decides from this one bit, per warp, whether streaming needs the
// masked stores at all.  Runs in place: only bit 31 changes, the reads look at the low bits.
__global__ void mark_sbits_neighbourhood_kernel(uint32_t* bits, unsigned nx, unsigned ny, unsigned nz) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nx * ny * nz) return;
  const unsigned z = i % nz, y = (i / nz) % ny, x = i / (nz * ny);
  uint32_t any = 0;
  for (int dx = -1; dx <= 1; ++dx) {
    const unsigned xx = (x + nx + dx) % nx;
    for (int dy = -1; dy <= 1; ++dy) {
      const unsigned yy = (y + ny + dy) % ny;
      for (int dz = -1; dz <= 1; ++dz) {
        const unsigned zz = (z + nz + dz) % nz;
        any |= bits[(size_t(xx) * ny + yy) * nz + zz] & ~lb::SBITS_NEAR;
      }
    }
  }
  if (any) atomicOr(bits + i, lb::SBITS_NEAR);
}

// ---- enstrophy: sum |curl u|^2 with the 6th-order periodic central difference -------------
// reference/lettuce/util/utility.py:37-99 (weights [-1/60, 3/20, -3/4, 3/4, -3/20, 1/60] on
// roll shifts [+3, +2, +1, -1, -2, -3], i.e. on u[x-3] ... u[x+3]) and
// observable_reporter.py:53-68.  u is [D][n_nodes] in lattice units, dx = 1.
// `ghost` planes of the neighbouring z-slabs on both sides of the LAST axis make that axis non-periodic
// (multi-GPU); every periodic axis needs n >= 3 (one wrap per tap), checked by the host.
template <typename T>
__device__ __forceinline__ T diff6(const T* __restrict__ u, size_t base, unsigned i, unsigned n, size_t stride,
                                   bool periodic) {
  // base = element index with the differentiated coordinate set to 0
  auto at = [&](int s) {
    int j = int(i) + s;
    if (periodic) {
      if (j < 0) j += int(n);
      if (j >= int(n)) j -= int(n);
    }
    return u[base + size_t(j) * stride];
  };
  return T(-1.0 / 60) * at(-3) + T(3.0 / 20) * at(-2) + T(-3.0 / 4) * at(-1) + T(3.0 / 4) * at(1) +
         T(-3.0 / 20) * at(2) + T(1.0 / 60) * at(3);
}

template <typename T, int D>
__global__ void __launch_bounds__(lb::STEP_BLOCK)
enstrophy_kernel(const T* __restrict__ u, unsigned n0, unsigned n1, unsigned n2, unsigned n_nodes, unsigned ghost,
                 lb::ObsScratch scratch, double* out) {
  // lattice axes: D == 2 -> (n0, n1), D == 3 -> (n0, n1, n2); last is contiguous and the only one with ghosts.
  // One thread per INTERIOR node; u is [D][n_ext], n_ext = nodes of the extended slab.
  const unsigned tid = blockIdx.x * lb::STEP_BLOCK + threadIdx.x;
  double acc = 0.0;
  if (tid < n_nodes) {
    const bool per = ghost == 0;
    if constexpr (D == 2) {
      const unsigned e1 = n1 + 2 * ghost;
      const size_t n_ext = size_t(n0) * e1;
      const T* u0 = u;
      const T* u1 = u + n_ext;
      const unsigned i0 = tid / n1, i1 = tid - i0 * n1 + ghost;
      const T d1u0 = diff6<T>(u0, size_t(i0) * e1, i1, e1, 1, per);          // d u0 / d y
      const T d0u1 = diff6<T>(u1, i1, i0, n0, e1, true);                     // d u1 / d x
      const T w = d1u0 - d0u1;
      acc = double(w * w);
    } else {
      const unsigned e2 = n2 + 2 * ghost;
      const size_t n_ext = size_t(n0) * n1 * e2;
      const T* u0 = u;
      const T* u1 = u + n_ext;
      const T* u2 = u + 2 * n_ext;
      const unsigned i0 = tid / (n1 * n2);
      const unsigned r = tid - i0 * n1 * n2;
      const unsigned i1 = r / n2, i2 = r - i1 * n2 + ghost;
      const size_t b0 = size_t(i1) * e2 + i2, s0 = size_t(n1) * e2;        // vary axis 0
      const size_t b1 = size_t(i0) * n1 * e2 + i2, s1 = e2;                // vary axis 1
      const size_t b2 = (size_t(i0) * n1 + i1) * e2;                       // vary axis 2
      const T wz = diff6<T>(u0, b1, i1, n1, s1, true) - diff6<T>(u1, b0, i0, n0, s0, true);
      const T wx = diff6<T>(u2, b1, i1, n1, s1, true) - diff6<T>(u1, b2, i2, e2, 1, per);
      const T wy = diff6<T>(u0, b2, i2, e2, 1, per) - diff6<T>(u2, b0, i0, n0, s0, true);
      acc = double(wz * wz) + double(wx * wx + wy * wy);
    }
  }
  // deterministic reduction; the single sum lands in the enstrophy slot
  lb::reduce_observables<lb::STEP_BLOCK>(acc, 0.0, 0.0, scratch, out + LB_OBS_ENSTROPHY, 1);
}

}  // namespace

extern "C" {

int lb_abi_version(void) { return LB_ABI_VERSION; }
const char* lb_last_error(void) { return g_err; }
int lb_stencil_q(int stencil) { return stencil_q(stencil); }

size_t lb_bytes_per_update(const lb_desc* d) {
  if (!d) return 0;
  const int q = stencil_q(d->stencil);
  return q < 0 ? 0 : size_t(2) * q * (d->dtype == LB_F64 ? 8 : 4);
}

int lb_vector_width(const lb_desc* d, int has_streaming_bits) {
  Geometry g;
  if (check_desc(d, &g)) return -1;
  static char dummy[16] __attribute__((aligned(16)));
  const int coll = (d->collision == LB_COLL_BGK && d->force != LB_FORCE_NONE) ? int(lb::COLL_BGK_FORCE) : d->collision;
  const int mode = has_streaming_bits ? lb::MODE_NCM_SBITS : lb::MODE_PLAIN;
  auto w = [&](auto s, auto t) {
    return vector_width<decltype(s), decltype(t)>(coll, mode, g, dummy, dummy, nullptr, nullptr, false);
  };
  const bool f64 = d->dtype == LB_F64;
  switch (d->stencil) {
    case LB_D2Q9: return f64 ? w(lb::D2Q9{}, double{}) : w(lb::D2Q9{}, float{});
    case LB_D3Q19: return f64 ? w(lb::D3Q19{}, double{}) : w(lb::D3Q19{}, float{});
    case LB_D3Q27: return f64 ? w(lb::D3Q27{}, double{}) : w(lb::D3Q27{}, float{});
  }
  return -1;
}

int lb_plan_resident(const lb_desc* desc, int n_steps, int sms, size_t smem_max, int out[8]) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!out || sms < 1) return fail(LB_EINVAL, "bad argument");
  for (int k = 0; k < 8; ++k) out[k] = 0;
  if (g.d != 2) return LB_OK;   // the tile-resident kernel is for the 2-D stencil
  const lb::ResidentPlan pl = lb::plan_resident(g.ny, g.nz, g.q, desc->dtype == LB_F64 ? 8 : 4, n_steps, sms, smem_max, 0);
  if (!pl.ok) return LB_OK;
  out[0] = 1; out[1] = int(pl.ty); out[2] = int(pl.tz); out[3] = pl.rounds; out[4] = pl.steps_base; out[5] = pl.steps_extra;
  out[6] = int(pl.plane_max); out[7] = int(pl.smem);
  return LB_OK;
}

int lb_step(const lb_desc* desc, const lb_fields* fields, int n_steps, void* stream) {
  return step_any(desc, fields, n_steps, true, stream);
}

int lb_collide(const lb_desc* desc, const void* f_in, void* f_out, const uint8_t* ncm,
               const lb_boundary* boundaries, int n_boundaries, void* stream) {
  lb_fields fl{};
  fl.f = const_cast<void*>(f_in);
  fl.f_next = f_out;
  fl.no_collision_mask = ncm;
  fl.boundaries = boundaries;
  fl.n_boundaries = n_boundaries;
  return step_any(desc, &fl, 1, false, stream);
}

int lb_stream(const lb_desc* desc, const void* f_in, void* f_out, const uint32_t* sbits, void* stream) {
  if (!desc) return fail(LB_EINVAL, "null descriptor");
  if (sbits) return fail(LB_EINVAL, "masked lb_stream: pass the masks through lb_step with LB_COLL_NONE");
  lb_desc d = *desc;
  d.collision = LB_COLL_NONE;
  lb_fields fl{};
  fl.f = const_cast<void*>(f_in);
  fl.f_next = f_out;
  return step_any(&d, &fl, 1, true, stream);
}

int lb_pack_streaming_mask(const lb_desc* desc, const uint8_t* nsm, uint32_t* bits, void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!nsm || !bits) return fail(LB_EINVAL, "null mask pointer");
  const unsigned block = 256, grid = (g.n_nodes + block - 1) / block;
  pack_sbits_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(nsm, bits, g.n_nodes, g.q);
  LB_CUDA(cudaGetLastError());
  mark_sbits_neighbourhood_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(bits, g.nx, g.ny, g.nz);
  LB_CUDA(cudaGetLastError());
  return LB_OK;
}

size_t lb_obs_scratch_bytes(const lb_desc* desc) {
  Geometry g;
  if (check_desc(desc, &g)) return 0;
  return lb::obs_scratch_bytes(max_blocks(g));
}

}  // extern "C"

namespace {

template <class S, typename T>
int moments_typed(const Geometry& g, const void* f, void* rho_out, void* u_out, void* scratch,
                  double* obs_out, cudaStream_t st) {
  lb::MomentsParams<T> p{};
  p.f = static_cast<const T*>(f);
  p.rho_out = static_cast<T*>(rho_out);
  p.u_out = static_cast<T*>(u_out);
  p.n_nodes = g.n_nodes;
  p.obs_out = obs_out;
  if (obs_out) {
    const int rc = obs_prepare(g, scratch, &p.obs, st);
    if (rc) return rc;
  }
  unsigned obs_blocks = 0;
  cudaError_t e = lb::launch_moments<S, T>(p, st, &obs_blocks);
  if (e != cudaSuccess) return fail(LB_ECUDA, "moments kernel launch failed: %s", cudaGetErrorString(e));
  if (obs_out) {   // second stage of the reduction: block partials -> obs_out
    e = launch_obs_finish(p.obs, obs_blocks, obs_out, lb::N_OBS_PART, st);
    if (e != cudaSuccess) return fail(LB_ECUDA, "observable reduction launch failed: %s", cudaGetErrorString(e));
  }
  return LB_OK;
}

template <typename T>
int moments_stencil(const lb_desc* d, const Geometry& g, const void* f, void* r, void* u, void* s,
                    double* o, cudaStream_t st) {
  switch (d->stencil) {
    case LB_D2Q9: return moments_typed<lb::D2Q9, T>(g, f, r, u, s, o, st);
    case LB_D3Q19: return moments_typed<lb::D3Q19, T>(g, f, r, u, s, o, st);
    case LB_D3Q27: return moments_typed<lb::D3Q27, T>(g, f, r, u, s, o, st);
  }
  return fail(LB_EINVAL, "bad stencil");
}

// The same with V consecutive nodes of the contiguous axis per thread and 16-byte loads: 30 vector loads per V nodes
// instead of 36 scalar loads per node (the stencil taps come from L1 / L2: the kernel is bound by the number of
// load requests, not by HBM), one partial sum per block and obs_finish_kernel behind it (no ticket on the blocks).
// The whole line and the ghost offset must be multiples of V (host-checked).
template <typename T, int V>
__device__ __forceinline__ void diff6_vec(const T* __restrict__ u, size_t first, long long stride, long long i, long long n,
                                          bool periodic, T (&out)[V]) {
  // derivative along a NON-contiguous axis: six whole vectors at i-3 .. i+3 of that axis; `first` = this thread's
  // vector with the axis coordinate set to 0
  auto vec_at = [&](int s) {
    long long j = i + s;
    if (periodic) { if (j < 0) j += n; if (j >= n) j -= n; }
    return lb::ld_vec<T, V>(u + first + size_t(j) * size_t(stride));
  };
  const lb::Vec<T, V> m3 = vec_at(-3), m2 = vec_at(-2), m1 = vec_at(-1), p1 = vec_at(1), p2 = vec_at(2), p3 = vec_at(3);
#pragma unroll
  for (int k = 0; k < V; ++k)
    out[k] = T(-1.0 / 60) * m3.v[k] + T(3.0 / 20) * m2.v[k] + T(-3.0 / 4) * m1.v[k] + T(3.0 / 4) * p1.v[k] +
             T(-3.0 / 20) * p2.v[k] + T(1.0 / 60) * p3.v[k];
}

template <typename T, int V>
__device__ __forceinline__ void diff6_line(const T* __restrict__ line, unsigned zv, unsigned nzv, bool periodic, T (&out)[V]) {
  // derivative along the contiguous axis: a window of the line around this thread's vector
  constexpr int W = (3 + V - 1) / V;      // vectors each side that cover the radius
  T a[(2 * W + 1) * V];
#pragma unroll
  for (int w = -W; w <= W; ++w) {
    long long j = (long long)zv + w;
    if (periodic) { if (j < 0) j += nzv; if (j >= (long long)nzv) j -= nzv; }
    const lb::Vec<T, V> v = lb::ld_vec<T, V>(line + j * V);   // non-periodic: ghost planes make j valid (may be < 0)
#pragma unroll
    for (int k = 0; k < V; ++k) a[(w + W) * V + k] = v.v[k];
  }
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int c = W * V + k;
    out[k] = T(-1.0 / 60) * a[c - 3] + T(3.0 / 20) * a[c - 2] + T(-3.0 / 4) * a[c - 1] + T(3.0 / 4) * a[c + 1] +
             T(-3.0 / 20) * a[c + 2] + T(1.0 / 60) * a[c + 3];
  }
}

template <typename T, int D, int V>
__global__ void __launch_bounds__(lb::STEP_BLOCK)
enstrophy_vec_kernel(const T* __restrict__ u, unsigned n0, unsigned n1, unsigned n2, unsigned n_threads, unsigned ghost,
                     lb::ObsScratch scratch) {
  const unsigned tid = blockIdx.x * lb::STEP_BLOCK + threadIdx.x;
  double acc = 0.0;
  if (tid < n_threads) {
    const bool per = ghost == 0;
    if constexpr (D == 2) {
      const unsigned e1 = n1 + 2 * ghost, nzv = n1 / V;
      const size_t n_ext = size_t(n0) * e1;
      const unsigned i0 = tid / nzv, zv = tid - i0 * nzv;
      const T* u0 = u;
      const T* u1 = u + n_ext;
      T d1u0[V], d0u1[V];
      diff6_line<T, V>(u0 + size_t(i0) * e1 + ghost, zv, nzv, per, d1u0);                  // d u0 / d y
      diff6_vec<T, V>(u1, size_t(ghost) + size_t(zv) * V, e1, i0, n0, true, d0u1);         // d u1 / d x
#pragma unroll
      for (int k = 0; k < V; ++k) { const T w = d1u0[k] - d0u1[k]; acc += double(w * w); }
    } else {
      const unsigned e2 = n2 + 2 * ghost, nzv = n2 / V;
      const size_t n_ext = size_t(n0) * n1 * e2;
      const unsigned i0 = tid / (n1 * nzv);
      const unsigned r = tid - i0 * n1 * nzv;
      const unsigned i1 = r / nzv, zv = r - i1 * nzv;
      const T* u0 = u;
      const T* u1 = u + n_ext;
      const T* u2 = u + 2 * n_ext;
      const size_t zoff = size_t(ghost) + size_t(zv) * V;
      const size_t b0 = size_t(i1) * e2 + zoff;                // this vector with i0 = 0
      const size_t b1 = size_t(i0) * n1 * e2 + zoff;           // ... with i1 = 0
      const size_t ln = (size_t(i0) * n1 + i1) * e2 + ghost;   // this line's first interior element
      const long long s0 = (long long)n1 * e2, s1 = e2;
      T a[V], b[V];
      diff6_vec<T, V>(u0, b1, s1, i1, n1, true, a);            // wz = d1 u0 - d0 u1
      diff6_vec<T, V>(u1, b0, s0, i0, n0, true, b);
      T wz[V];
#pragma unroll
      for (int k = 0; k < V; ++k) wz[k] = a[k] - b[k];
      diff6_vec<T, V>(u2, b1, s1, i1, n1, true, a);            // wx = d1 u2 - d2 u1
      diff6_line<T, V>(u1 + ln, zv, nzv, per, b);
      T wx[V];
#pragma unroll
      for (int k = 0; k < V; ++k) wx[k] = a[k] - b[k];
      diff6_line<T, V>(u0 + ln, zv, nzv, per, a);              // wy = d2 u0 - d0 u2
      diff6_vec<T, V>(u2, b0, s0, i0, n0, true, b);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const T wy = a[k] - b[k];
        acc += double(wz[k] * wz[k]) + double(wx[k] * wx[k] + wy * wy);
      }
    }
  }
  double z1 = 0.0, z2 = 0.0;
  lb::block_reduce3<lb::STEP_BLOCK>(acc, z1, z2);
  if (threadIdx.x == 0) {
    double* bp = scratch.block_partials + size_t(blockIdx.x) * lb::N_OBS_PART;
    bp[0] = acc; bp[1] = 0.0; bp[2] = 0.0;
  }
}

template <typename T>
int enstrophy_typed(const lb_desc* d, const Geometry& g, const void* u, int ghost, void* scratch, double* obs_out,
                    cudaStream_t st) {
  // torch.roll wraps any size; the kernel wraps once per tap: every periodic axis needs >= 3 nodes
  for (int a = 0; a < g.d; ++a)
    if (d->dims[a] < 3 && !(a == g.d - 1 && ghost > 0))
      return fail(LB_EUNSUPPORTED, "the 6th-order gradient of the enstrophy needs at least 3 nodes per axis (dims[%d] = %lld)",
                  a, (long long)d->dims[a]);
  if (ghost != 0 && ghost < 3) return fail(LB_EINVAL, "ghost = %d: must be 0 (periodic) or >= 3 (stencil radius)", ghost);
  lb::ObsScratch sc;
  const int rc = obs_prepare(g, scratch, &sc, st);
  if (rc) return rc;
  constexpr int V = 16 / int(sizeof(T));
  const unsigned nlast = unsigned(d->dims[g.d - 1]);
  static const bool scalar_only = getenv("LB_ENSTROPHY_SCALAR") != nullptr;   // A/B timing only
  if (!scalar_only && nlast % V == 0 && unsigned(ghost) % V == 0 && aligned_to(u, 16) && nlast / V >= 2u * ((3 + V - 1) / V) + 1u) {
    const unsigned n_threads = g.n_nodes / V;
    const unsigned blocks = (n_threads + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK;
    if (g.d == 2)
      enstrophy_vec_kernel<T, 2, V><<<blocks, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(u), (unsigned)d->dims[0],
                                                                      (unsigned)d->dims[1], 1u, n_threads, unsigned(ghost), sc);
    else
      enstrophy_vec_kernel<T, 3, V><<<blocks, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(u), (unsigned)d->dims[0],
                                                                      (unsigned)d->dims[1], (unsigned)d->dims[2], n_threads,
                                                                      unsigned(ghost), sc);
    LB_CUDA(cudaGetLastError());
    LB_CUDA(launch_obs_finish(sc, blocks, obs_out + LB_OBS_ENSTROPHY, 1, st));
    return LB_OK;
  }
  const unsigned grid = (g.n_nodes + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK;
  if (g.d == 2)
    enstrophy_kernel<T, 2><<<grid, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(u), (unsigned)d->dims[0],
                                                           (unsigned)d->dims[1], 1u, g.n_nodes, unsigned(ghost), sc, obs_out);
  else
    enstrophy_kernel<T, 3><<<grid, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(u), (unsigned)d->dims[0],
                                                           (unsigned)d->dims[1], (unsigned)d->dims[2],
                                                           g.n_nodes, unsigned(ghost), sc, obs_out);
  LB_CUDA(cudaGetLastError());
  return LB_OK;
}

}  // namespace

namespace {

// 6th-order central difference along one axis (utility.py:37-99).  `i` is the coordinate along
// the axis, `n` its length, `stride` its element stride, `base` the element with that coordinate
// set to 0; PERIODIC wraps the index, otherwise the caller guarantees 3 valid planes either side.
template <typename T, bool PERIODIC>
__device__ __forceinline__ T diff6_at(const T* __restrict__ u, size_t base, long long i, long long n, size_t stride) {
  auto at = [&](int s) {
    long long j = i + s;
    if (PERIODIC) { if (j < 0) j += n; if (j >= n) j -= n; }
    return u[base + size_t(j) * stride];
  };
  return T(-1.0 / 60) * at(-3) + T(3.0 / 20) * at(-2) + T(-3.0 / 4) * at(-1) + T(3.0 / 4) * at(1) +
         T(-3.0 / 20) * at(2) + T(1.0 / 60) * at(3);
}

// one thread per interior node; kernel axes (X, Y, Z), Z contiguous and the only axis with ghosts
template <class S, typename T>
__global__ void __launch_bounds__(lb::STEP_BLOCK)
initialize_kernel(const T* __restrict__ rho, const T* __restrict__ u, T* __restrict__ f, unsigned nx, unsigned ny,
                  unsigned nz, unsigned ghost, T tau, int with_fneq) {
  const size_t tid = size_t(blockIdx.x) * lb::STEP_BLOCK + threadIdx.x;
  const size_t n = size_t(nx) * ny * nz;
  if (tid >= n) return;
  const unsigned z = unsigned(tid % nz);
  const unsigned y = unsigned((tid / nz) % ny);
  const unsigned x = unsigned(tid / (size_t(nz) * ny));
  const unsigned nze = nz + 2 * ghost;
  const size_t n_ext = size_t(nx) * ny * nze;
  const size_t e = (size_t(x) * ny + y) * nze + z + ghost;   // this node in the extended arrays
  const T r = rho[e];
  T uu[3] = {T(0), T(0), T(0)};
#pragma unroll
  for (int c = 0; c < S::D; ++c) uu[S::axis(c)] = u[c * n_ext + e];
  T feq[S::Q];
  lb::equilibrium<S, T>(r, uu, feq);
  if (with_fneq) {
    // S[a][b] = d u_a / d x_b in kernel axes (_flow.py:314)
    T g[3][3] = {};
#pragma unroll
    for (int c = 0; c < S::D; ++c) {
      const int a = S::axis(c);
      const T* ua = u + c * n_ext;
      if (S::D == 3) g[a][0] = diff6_at<T, true>(ua, size_t(y) * nze + z + ghost, x, nx, size_t(ny) * nze);
      g[a][1] = diff6_at<T, true>(ua, size_t(x) * ny * nze + z + ghost, y, ny, nze);
      g[a][2] = ghost ? diff6_at<T, false>(ua, (size_t(x) * ny + y) * nze, (long long)z + ghost, nze, 1)
                      : diff6_at<T, true>(ua, (size_t(x) * ny + y) * nze, z, nz, 1);
    }
    // Pi_1 = 1.0 * tau * rho * S / cs^2 (:315); fneq_q = w_q sum_ab Pi_1[a][b] (e_qa e_qb - cs^2 delta_ab) (:316-321)
    const T cs2 = T(LB_CS2);
    // the reference builds delta_ab cs^2 as `torch.eye(d) * cs ** 2` (_flow.py:328-330): torch.eye defaults
    // to float32, so that term carries cs^2 rounded to float32 whatever the lattice dtype
    const T cs2_eye = T(float(LB_CS2));
    T pi[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) pi[a][b] = tau * r * g[a][b] / cs2;
    lb::static_for<S::Q>([&](auto qc) {
      constexpr int q = decltype(qc)::value;
      T acc = T(0);
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          if (S::D == 2 && (a == 0 || b == 0)) continue;
          const T qab = T(S::e(q, a) * S::e(q, b)) - (a == b ? cs2_eye : T(0));
          acc += pi[a][b] * qab;
        }
      feq[q] -= T(S::w(q)) * acc;
    });
  }
  lb::static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    f[q * n + tid] = feq[q];
  });
}

// The same with V nodes of a z line per thread: 16-byte loads of rho and u, the 6 x 3 taps of a derivative as whole
// vectors (diff6_vec / diff6_line, shared with the enstrophy kernel), 16-byte stores of the populations.  The scalar
// kernel above issues 54 + 4 scalar loads and 19 - 27 scalar stores per node and ran at 0.23 of the HBM roofline (it
// writes Q values per node once: 1.0 ms at 256^3 D3Q19 fp32).  Element by element the arithmetic is the scalar kernel's.
template <class S, typename T, int V>
__global__ void __launch_bounds__(lb::STEP_BLOCK)
initialize_vec_kernel(const T* __restrict__ rho, const T* __restrict__ u, T* __restrict__ f, unsigned nx, unsigned ny,
                      unsigned nz, unsigned ghost, T tau, int with_fneq) {
  const unsigned nzv = nz / V;
  const size_t tid = size_t(blockIdx.x) * lb::STEP_BLOCK + threadIdx.x;
  const size_t n = size_t(nx) * ny * nz;
  if (tid >= size_t(nx) * ny * nzv) return;
  const unsigned zv = unsigned(tid % nzv);
  const unsigned y = unsigned((tid / nzv) % ny);
  const unsigned x = unsigned(tid / (size_t(nzv) * ny));
  const unsigned nze = nz + 2 * ghost;
  const size_t n_ext = size_t(nx) * ny * nze;
  const size_t zoff = size_t(ghost) + size_t(zv) * V;
  const size_t e = (size_t(x) * ny + y) * nze + zoff;        // this vector in the extended arrays
  const lb::Vec<T, V> rv = lb::ld_vec<T, V>(rho + e);
  T uu[3][V] = {};
#pragma unroll
  for (int c = 0; c < S::D; ++c) {
    const lb::Vec<T, V> v = lb::ld_vec<T, V>(u + c * n_ext + e);
#pragma unroll
    for (int k = 0; k < V; ++k) uu[S::axis(c)][k] = v.v[k];
  }
  T gd[3][3][V] = {};                                        // gd[a][b][k] = d u_a / d x_b of node k (kernel axes)
  if (with_fneq) {
#pragma unroll
    for (int c = 0; c < S::D; ++c) {
      const int a = S::axis(c);
      const T* ua = u + c * n_ext;
      if (S::D == 3) diff6_vec<T, V>(ua, size_t(y) * nze + zoff, (long long)ny * nze, x, nx, true, gd[a][0]);
      diff6_vec<T, V>(ua, size_t(x) * ny * nze + zoff, nze, y, ny, true, gd[a][1]);
      diff6_line<T, V>(ua + (size_t(x) * ny + y) * nze + ghost, zv, nzv, ghost == 0, gd[a][2]);
    }
  }
  T out[S::Q][V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const T r = rv.v[k];
    const T un[3] = {uu[0][k], uu[1][k], uu[2][k]};
    T feq[S::Q];
    lb::equilibrium<S, T>(r, un, feq);
    if (with_fneq) {
      const T cs2 = T(LB_CS2);
      const T cs2_eye = T(float(LB_CS2));                    // torch.eye(d) * cs ** 2 is float32 in the reference (see above)
      T pi[3][3];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) pi[a][b] = tau * r * gd[a][b][k] / cs2;
      lb::static_for<S::Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        T acc = T(0);
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) {
            if (S::D == 2 && (a == 0 || b == 0)) continue;
            const T qab = T(S::e(q, a) * S::e(q, b)) - (a == b ? cs2_eye : T(0));
            acc += pi[a][b] * qab;
          }
        feq[q] -= T(S::w(q)) * acc;
      });
    }
    lb::static_for<S::Q>([&](auto qc) { out[decltype(qc)::value][k] = feq[decltype(qc)::value]; });
  }
  const size_t node = tid * V;                               // interior index of this thread's first node
  lb::static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    lb::Vec<T, V> v;
#pragma unroll
    for (int k = 0; k < V; ++k) v.v[k] = out[q][k];
    lb::st_vec<T, V>(f + q * n + node, v);
  });
}

template <class S, typename T>
int initialize_typed(const lb_desc* d, const Geometry& g, const void* rho, const void* u, int ghost, int with_fneq,
                     void* f, cudaStream_t st) {
  const size_t n = size_t(g.n_nodes);
  constexpr int V = 16 / int(sizeof(T));
  static const bool scalar_only = getenv("LB_INIT_SCALAR") != nullptr;   // A/B timing only
  if (!scalar_only && g.nz % V == 0 && unsigned(ghost) % V == 0 && g.nz / V >= 3u && aligned_to(rho, 16) && aligned_to(u, 16) &&
      aligned_to(f, 16) && (n * sizeof(T)) % 16 == 0 && (size_t(g.nx) * g.ny * (g.nz + 2 * ghost) * sizeof(T)) % 16 == 0) {
    const size_t threads = n / V;
    const unsigned vgrid = unsigned((threads + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK);
    initialize_vec_kernel<S, T, V><<<vgrid, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(rho), static_cast<const T*>(u),
                                                                   static_cast<T*>(f), g.nx, g.ny, g.nz, unsigned(ghost),
                                                                   T(d->tau), with_fneq);
    LB_CUDA(cudaGetLastError());
    return LB_OK;
  }
  const unsigned grid = unsigned((n + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK);
  initialize_kernel<S, T><<<grid, lb::STEP_BLOCK, 0, st>>>(static_cast<const T*>(rho), static_cast<const T*>(u),
                                                        static_cast<T*>(f), g.nx, g.ny, g.nz, unsigned(ghost),
                                                        T(d->tau), with_fneq);
  LB_CUDA(cudaGetLastError());
  return LB_OK;
}

template <typename T>
int initialize_stencil(const lb_desc* d, const Geometry& g, const void* rho, const void* u, int ghost, int with_fneq,
                       void* f, cudaStream_t st) {
  switch (d->stencil) {
    case LB_D2Q9: return initialize_typed<lb::D2Q9, T>(d, g, rho, u, ghost, with_fneq, f, st);
    case LB_D3Q19: return initialize_typed<lb::D3Q19, T>(d, g, rho, u, ghost, with_fneq, f, st);
    case LB_D3Q27: return initialize_typed<lb::D3Q27, T>(d, g, rho, u, ghost, with_fneq, f, st);
  }
  return fail(LB_EINVAL, "bad stencil");
}

}  // namespace

extern "C" {

int lb_initialize(const lb_desc* desc, const void* rho, const void* u, int ghost, int with_fneq, void* f,
                  void* stream) {
  Geometry g;
  lb_desc dd;
  if (!desc) return fail(LB_EINVAL, "null descriptor");
  dd = *desc;
  dd.collision = LB_COLL_NONE;   // tau is only read when with_fneq; do not require it otherwise
  int rc = check_desc(&dd, &g);
  if (rc) return rc;
  if (!rho || !u || !f) return fail(LB_EINVAL, "null pointer");
  if (ghost != 0 && ghost < 3) return fail(LB_EINVAL, "ghost = %d: must be 0 (periodic) or >= 3 (stencil radius)", ghost);
  if (with_fneq && !(desc->tau > 0.0)) return fail(LB_EINVAL, "tau must be > 0 for the non-equilibrium part");
  if (with_fneq && (g.nz < 3 || g.ny < 3 || (g.d == 3 && g.nx < 3)))
    return fail(LB_EUNSUPPORTED, "the 6th-order gradient needs at least 3 nodes per axis");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return desc->dtype == LB_F32 ? initialize_stencil<float>(desc, g, rho, u, ghost, with_fneq, f, st)
                               : initialize_stencil<double>(desc, g, rho, u, ghost, with_fneq, f, st);
}

}  // extern "C"

namespace {
template <class S, typename T>
int output_typed(const Geometry& g, const void* f, double rho0, double p_scale, double u_scale, void* out,
                 cudaStream_t st) {
  lb::OutputParams<T> p{};
  p.f = static_cast<const T*>(f);
  p.out = static_cast<T*>(out);
  p.nz = g.nz;
  if (g.d == 2) { p.na = g.ny; p.nm = 1; }          // 2-D runs as X = 1: lattice x = kernel Y
  else { p.na = g.nx; p.nm = g.ny; }
  p.n_nodes = g.n_nodes;
  p.rho0 = T(rho0);
  p.p_scale = T(p_scale);
  p.u_scale = T(u_scale);
  cudaError_t e = lb::launch_output_fields<S, T>(p, st);
  if (e != cudaSuccess) return fail(LB_ECUDA, "output kernel launch failed: %s", cudaGetErrorString(e));
  return LB_OK;
}

template <typename T>
int output_stencil(const lb_desc* d, const Geometry& g, const void* f, double rho0, double ps, double us, void* out,
                   cudaStream_t st) {
  switch (d->stencil) {
    case LB_D2Q9: return output_typed<lb::D2Q9, T>(g, f, rho0, ps, us, out, st);
    case LB_D3Q19: return output_typed<lb::D3Q19, T>(g, f, rho0, ps, us, out, st);
    case LB_D3Q27: return output_typed<lb::D3Q27, T>(g, f, rho0, ps, us, out, st);
  }
  return fail(LB_EINVAL, "bad stencil");
}
}  // namespace

extern "C" {

int lb_moments(const lb_desc* desc, const void* f, void* rho_out, void* u_out, void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!f) return fail(LB_EINVAL, "null population buffer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return desc->dtype == LB_F32 ? moments_stencil<float>(desc, g, f, rho_out, u_out, nullptr, nullptr, st)
                               : moments_stencil<double>(desc, g, f, rho_out, u_out, nullptr, nullptr, st);
}

int lb_output_fields(const lb_desc* desc, const void* f, double rho0, double p_scale, double u_scale, void* out,
                     void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!f || !out) return fail(LB_EINVAL, "null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return desc->dtype == LB_F32 ? output_stencil<float>(desc, g, f, rho0, p_scale, u_scale, out, st)
                               : output_stencil<double>(desc, g, f, rho0, p_scale, u_scale, out, st);
}

int lb_observables(const lb_desc* desc, const void* f, int want_enstrophy, void* u_scratch,
                   void* obs_scratch, double* obs_out, void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!f || !obs_scratch || !obs_out) return fail(LB_EINVAL, "null pointer");
  if (want_enstrophy && !u_scratch) return fail(LB_EINVAL, "enstrophy needs u_scratch");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  rc = desc->dtype == LB_F32
           ? moments_stencil<float>(desc, g, f, nullptr, want_enstrophy ? u_scratch : nullptr, obs_scratch, obs_out, st)
           : moments_stencil<double>(desc, g, f, nullptr, want_enstrophy ? u_scratch : nullptr, obs_scratch, obs_out, st);
  if (rc || want_enstrophy != 1) return rc;   // 2: the caller only wanted u_scratch filled
  return desc->dtype == LB_F32 ? enstrophy_typed<float>(desc, g, u_scratch, 0, obs_scratch, obs_out, st)
                               : enstrophy_typed<double>(desc, g, u_scratch, 0, obs_scratch, obs_out, st);
}

int lb_enstrophy_slab(const lb_desc* desc, const void* u, int ghost, void* obs_scratch, double* obs_out, void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!u || !obs_scratch || !obs_out) return fail(LB_EINVAL, "null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return desc->dtype == LB_F32 ? enstrophy_typed<float>(desc, g, u, ghost, obs_scratch, obs_out, st)
                               : enstrophy_typed<double>(desc, g, u, ghost, obs_scratch, obs_out, st);
}

int lb_enstrophy(const lb_desc* desc, const void* u, void* obs_scratch, double* obs_out, void* stream) {
  return lb_enstrophy_slab(desc, u, 0, obs_scratch, obs_out, stream);
}

}  // extern "C"

// ---- lb_observables_static: lb_observables without a caller-provided (or any) allocation ----------------------
// The reduction scratch (ticket counters + per-block and per-group partials, lb::obs_scratch_bytes) lives in a
// file-scope device array of a fixed maximum size: LB_STATIC_OBS_MAX_NODES nodes.  Every GPU that loads the library
// holds these bytes until the process exits.  All calls on a device share the array.
namespace {
constexpr size_t STATIC_OBS_MAX_BLOCKS = (size_t(LB_STATIC_OBS_MAX_NODES) + lb::STEP_BLOCK - 1) / lb::STEP_BLOCK;
constexpr size_t STATIC_OBS_GROUPS = (STATIC_OBS_MAX_BLOCKS + lb::OBS_GROUP - 1) / lb::OBS_GROUP;
constexpr size_t STATIC_OBS_BYTES =   // == lb::obs_scratch_bytes(STATIC_OBS_MAX_BLOCKS), checked at set-up
    64 + ((STATIC_OBS_GROUPS * sizeof(unsigned) + 63) / 64) * 64 +
    (STATIC_OBS_MAX_BLOCKS + STATIC_OBS_GROUPS) * lb::N_OBS_PART * sizeof(double);
static __device__ __align__(256) unsigned char g_static_obs_scratch[STATIC_OBS_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
void* g_static_obs_ptr[STATIC_MAX_DEVICES] = {};   // device address of the array, per device ordinal
std::mutex g_static_obs_mutex;

// Device address of g_static_obs_scratch on the CURRENT device.  In host code the array's name is only a stand-in
// for the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that
// device current.
int static_obs_scratch(int dev, void** out) {
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return fail(LB_EUNSUPPORTED, "device ordinal %d beyond %d", dev, STATIC_MAX_DEVICES);
  std::lock_guard<std::mutex> lock(g_static_obs_mutex);
  if (!g_static_obs_ptr[dev]) {
    if (lb::obs_scratch_bytes(STATIC_OBS_MAX_BLOCKS) != STATIC_OBS_BYTES)
      return fail(LB_ECUDA, "static scratch size out of step with lb::obs_scratch_bytes");
    void* ptr = nullptr;
    LB_CUDA(cudaGetSymbolAddress(&ptr, g_static_obs_scratch));
    if (!ptr) return fail(LB_ECUDA, "cudaGetSymbolAddress returned a null address");
    g_static_obs_ptr[dev] = ptr;
  }
  *out = g_static_obs_ptr[dev];
  return LB_OK;
}
}  // namespace

extern "C" {

size_t lb_static_obs_max_nodes(void) { return size_t(LB_STATIC_OBS_MAX_NODES); }

int lb_observables_static(const lb_desc* desc, const void* f, double* obs_out, void* stream) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!f || !obs_out) return fail(LB_EINVAL, "null pointer");
  if (size_t(g.n_nodes) > size_t(LB_STATIC_OBS_MAX_NODES))
    return fail(LB_EUNSUPPORTED,
                "lb_observables_static: %u nodes exceed the built-in scratch of LB_STATIC_OBS_MAX_NODES = %llu nodes "
                "(%llu bytes); use lb_observables with caller-provided scratch",
                g.n_nodes, (unsigned long long)LB_STATIC_OBS_MAX_NODES, (unsigned long long)STATIC_OBS_BYTES);
  int dev = -1;
  LB_CUDA(cudaGetDevice(&dev));
  cudaPointerAttributes attr{};
  LB_CUDA(cudaPointerGetAttributes(&attr, f));
  if (attr.type == cudaMemoryTypeDevice && attr.device != dev)
    return fail(LB_EINVAL, "f lives on device %d but device %d is current: the built-in scratch is per device", attr.device, dev);
  void* scratch = nullptr;
  rc = static_obs_scratch(dev, &scratch);
  if (rc) return rc;
  // moments_typed clears the ticket counters (cudaMemsetAsync on `stream`, the only part that must start at zero;
  // the partials are written before they are read) and launches the same moments_kernel<OBS> as lb_observables
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return desc->dtype == LB_F32 ? moments_stencil<float>(desc, g, f, nullptr, nullptr, scratch, obs_out, st)
                               : moments_stencil<double>(desc, g, f, nullptr, nullptr, scratch, obs_out, st);
}

}  // extern "C"

// ---- host-buffer entry: H2D -> n steps -> D2H ------------------------------------------
namespace {
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
};
size_t extent(const Geometry& g, const lb_desc* d, const int64_t* stride) {
  size_t e = 1;
  for (int a = 0; a < g.d; ++a) e += size_t(d->dims[a] - 1) * size_t(stride[a]);
  return e;
}
}  // namespace

extern "C" int lb_run_host(const lb_desc* desc, void* f_host, const uint8_t* ncm_host,
                           const uint8_t* nsm_host, const lb_boundary* bnd_host, int n_boundaries,
                           int n_steps, double* obs_host) {
  Geometry g;
  int rc = check_desc(desc, &g);
  if (rc) return rc;
  if (!f_host) return fail(LB_EINVAL, "null host buffer");
  if (n_boundaries < 0 || n_boundaries > LB_MAX_BOUNDARIES) return fail(LB_EINVAL, "n_boundaries = %d", n_boundaries);
  const size_t esz = desc->dtype == LB_F64 ? 8 : 4;
  const size_t fbytes = size_t(g.q) * g.n_nodes * esz;
  DevBuf fa, fb, ncm, nsm, bits, scratch, obs, brho[LB_MAX_BOUNDARIES], bu[LB_MAX_BOUNDARIES];
  cudaStream_t st = nullptr;
  LB_CUDA(cudaMalloc(&fa.p, fbytes));
  LB_CUDA(cudaMalloc(&fb.p, fbytes));
  LB_CUDA(cudaMemcpyAsync(fa.p, f_host, fbytes, cudaMemcpyHostToDevice, st));
  lb_fields fl{};
  fl.f = fa.p;
  fl.f_next = fb.p;
  if (ncm_host) {
    LB_CUDA(cudaMalloc(&ncm.p, g.n_nodes));
    LB_CUDA(cudaMemcpyAsync(ncm.p, ncm_host, g.n_nodes, cudaMemcpyHostToDevice, st));
    fl.no_collision_mask = static_cast<const uint8_t*>(ncm.p);
  }
  if (nsm_host) {
    bool any = false;
    for (size_t i = 0; i < size_t(g.q) * g.n_nodes && !any; ++i) any = nsm_host[i] == 1;
    if (any) {
      LB_CUDA(cudaMalloc(&nsm.p, size_t(g.q) * g.n_nodes));
      LB_CUDA(cudaMalloc(&bits.p, size_t(g.n_nodes) * 4));
      LB_CUDA(cudaMemcpyAsync(nsm.p, nsm_host, size_t(g.q) * g.n_nodes, cudaMemcpyHostToDevice, st));
      rc = lb_pack_streaming_mask(desc, static_cast<const uint8_t*>(nsm.p), static_cast<uint32_t*>(bits.p), st);
      if (rc) return rc;
      fl.streaming_bits = static_cast<const uint32_t*>(bits.p);
    }
  }
  std::vector<lb_boundary> bd(bnd_host, bnd_host + n_boundaries);
  for (int i = 0; i < n_boundaries; ++i) {
    if (bd[i].kind != LB_BND_EQUILIBRIUM) continue;
    if (!bd[i].rho || !bd[i].u) return fail(LB_EINVAL, "boundary %d: equilibrium boundary needs rho and u", i);
    const size_t nr = extent(g, desc, bd[i].rho_stride) * esz;
    const size_t nu = (extent(g, desc, bd[i].u_stride) + size_t(g.d - 1) * size_t(bd[i].u_comp_stride)) * esz;
    LB_CUDA(cudaMalloc(&brho[i].p, nr));
    LB_CUDA(cudaMalloc(&bu[i].p, nu));
    LB_CUDA(cudaMemcpyAsync(brho[i].p, bd[i].rho, nr, cudaMemcpyHostToDevice, st));
    LB_CUDA(cudaMemcpyAsync(bu[i].p, bd[i].u, nu, cudaMemcpyHostToDevice, st));
    bd[i].rho = brho[i].p;
    bd[i].u = bu[i].p;
  }
  fl.boundaries = bd.data();
  fl.n_boundaries = n_boundaries;
  rc = lb_step(desc, &fl, n_steps, st);
  if (rc) return rc;
  void* result = (n_steps & 1) ? fb.p : fa.p;
  if (obs_host) {
    LB_CUDA(cudaMalloc(&scratch.p, lb_obs_scratch_bytes(desc)));
    LB_CUDA(cudaMalloc(&obs.p, LB_OBS_COUNT * sizeof(double)));
    LB_CUDA(cudaMemsetAsync(obs.p, 0, LB_OBS_COUNT * sizeof(double), st));
    rc = lb_observables(desc, result, 0, nullptr, scratch.p, static_cast<double*>(obs.p), st);
    if (rc) return rc;
    LB_CUDA(cudaMemcpyAsync(obs_host, obs.p, LB_OBS_COUNT * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  LB_CUDA(cudaMemcpyAsync(f_host, result, fbytes, cudaMemcpyDeviceToHost, st));
  LB_CUDA(cudaStreamSynchronize(st));
  return LB_OK;
}
