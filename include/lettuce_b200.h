/*
 * lettuce_b200.h -- C ABI of the B200-native stream-collide engine.
 *
 * This is the drop-in boundary for lettuce's per-timestep hot path.  Every entry point
 * takes plain pointers and sizes (device pointers unless the name says `_host`), returns
 * 0 on success or a negative LB_E* code, and leaves a message retrievable with
 * lb_last_error().  Nothing here depends on torch / pybind.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   - the per-step seam `Simulation._collide_and_stream(simulation)`
 *       lettuce/_simulation.py:92-96,148,202              -> lb_step()
 *   - the generated pybind11 extension entry
 *       `collide_and_stream_<hash>(f, f_next[, no_collision_mask, no_streaming_mask]
 *        [, tau_inv][, velocityK, densityK]...)`
 *       lettuce/cuda_native/_template.py:59-86,297-367     -> lb_step()
 *   - `Collision.__call__(flow)` (bgk_collision.py:17-22, trt_collision.py:16-27,
 *       kbc_collision.py:96-160, no_collision.py:9-11) and `Boundary.__call__(flow)`
 *       (bounce_back_boundary.py:17-18, equilibrium_boundary_pu.py:27-32)
 *                                                          -> lb_collide()
 *   - `Flow.rho/j/u` (lettuce/_flow.py:136-172)            -> lb_moments()
 *   - the field extraction of `VTKReporter.__call__` (ext/_reporter/vtk_reporter.py:29-47)
 *                                                          -> lb_output_fields()
 *   - the reporter reductions IncompressibleKineticEnergy / MaximumVelocity / Enstrophy
 *       (lettuce/ext/_reporter/observable_reporter.py:27-68, util/utility.py:37-99)
 *                                                          -> lb_observables(), and the
 *       fused per-step variant through lb_fields.obs_out
 *
 * Layout: populations are f[q][x][y][z] (2-D: f[q][x][y]), C-contiguous, last axis
 * fastest, exactly the reference's `flow.f` tensor (lettuce/_flow.py:89-91).
 * Stencil index order is the reference's (d2q9.py:8-10, d3q19.py:8-13, d3q27.py:8-12).
 */
#ifndef LETTUCE_B200_H
#define LETTUCE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define LB_API __attribute__((visibility("default")))
#else
#define LB_API
#endif

#define LB_ABI_VERSION 6
#define LB_MAX_BOUNDARIES 8

enum { LB_D2Q9 = 0, LB_D3Q19 = 1, LB_D3Q27 = 2 };                     /* lb_desc.stencil   */
enum { LB_F32 = 0, LB_F64 = 1 };                                      /* lb_desc.dtype     */
enum { LB_FORCE_NONE = 0, LB_FORCE_GUO = 1, LB_FORCE_SHAN_CHEN = 2 };  /* lb_desc.force     */
enum { LB_COLL_NONE = 0, LB_COLL_BGK = 1, LB_COLL_TRT = 2, LB_COLL_KBC = 3,
       LB_COLL_REGULARIZED = 4, /* regularized_collision.py:17-44; tau as the reference reads it from the units */
       LB_COLL_SMAGORINSKY = 5  /* smagorinsky_collision.py:19-36; lb_desc.tau_minus carries the Smagorinsky
                                   constant (default 0.17), two fixed-point iterations for tau_eff */
}; /* .collision */
enum { LB_BND_BOUNCE_BACK = 1, LB_BND_EQUILIBRIUM = 2,                 /* lb_boundary.kind  */
       LB_BND_ABB_OUTLET = 3, LB_BND_EQ_OUTLET_P = 4 };

enum {
  LB_OK = 0,
  LB_EINVAL = -1,      /* bad descriptor / null pointer / unsupported combination */
  LB_ECUDA = -2,       /* a CUDA runtime call failed; see lb_last_error()         */
  LB_EUNSUPPORTED = -3 /* e.g. KBC on D3Q19 (kbc_collision.py:101,115)            */
};

/* observables written per step into lb_fields.obs_out / by lb_observables():
 * all in LATTICE units, accumulated in double precision.                          */
enum {
  LB_OBS_ENERGY = 0,   /* sum_nodes 0.5 |u|^2           (_flow.py:178-181)          */
  LB_OBS_MAXU2 = 1,    /* max_nodes |u|^2               (observable_reporter.py:30) */
  LB_OBS_MASS = 2,     /* sum_nodes rho                                           */
  LB_OBS_ENSTROPHY = 3,/* sum |curl u|^2, 6th-order FD  (observable_reporter.py:53-68) */
  LB_OBS_COUNT = 4
};

typedef struct lb_desc {
  int32_t stencil;     /* LB_D2Q9 | LB_D3Q19 | LB_D3Q27 */
  int32_t dtype;       /* LB_F32 | LB_F64               */
  int32_t collision;   /* LB_COLL_*                     */
  int32_t ndim;        /* 2 or 3 (must match stencil)   */
  int64_t dims[3];     /* resolution, dims[ndim-1] is the contiguous axis */
  double tau;          /* BGK: tau; TRT: tau_plus; KBC: tau (beta = 1/(2 tau)) */
  double tau_minus;    /* TRT: tau_minus (trt_collision.py:12-14); Smagorinsky: the constant (:13) */
  /* Forcing scheme of BGKCollision(tau, force) (bgk_collision.py:17-22), LB_COLL_BGK only:
   *   u = j / rho + scale * acceleration / rho,   f - (f - feq(rho, u)) / tau + S
   *   LB_FORCE_GUO       scale = 1/2,       S_q = (1 - 1/(2 force_tau)) w_q ((e_q - u)/cs^2 + (e_q.u) e_q/cs^4) . a
   *                                         (ext/_force/guo.py:16-37)
   *   LB_FORCE_SHAN_CHEN scale = force_tau, S = 0                   (ext/_force/shan_chen.py:13-23)
   * acceleration: constant vector in lattice units, lattice axes; force_tau: the tau given to the Force. */
  double acceleration[3];
  double force_tau;
  int32_t force;       /* LB_FORCE_NONE | LB_FORCE_GUO | LB_FORCE_SHAN_CHEN */
  int32_t kernel_flags; /* 0 = the library chooses; LB_KF_* select a kernel variant explicitly (tests, A/B timing) */
} lb_desc;
/* lb_desc.kernel_flags: every variant produces bit-identical results */
enum {
  LB_KF_NO_TMA = 1,    /* never use the persistent TMA-prefetch step kernel                                */
  LB_KF_FORCE_TMA = 2, /* use it wherever it applies (plain lattices of whole, 16-byte aligned tiles)      */
  LB_KF_NO_LOOP = 4,   /* one launch per time step also where the cooperative multi-step kernel would fit */
  LB_KF_FORCE_PREFETCH = 8, /* persistent per-thread prefetch (cp.async) step kernel wherever it is compiled and applies */
  LB_KF_NO_PREFETCH = 16,   /* never use it                                                                    */
  LB_KF_NO_RESIDENT = 32    /* small 2-D lattices: not the tile-resident multi-step kernel (state in shared memory) */
};

/* One entry per boundary object, in the order lettuce sorted them
 * (_simulation.py:57-58): entry i applies where no_collision_mask == i + 1.
 * rho / u are in lattice units, element type = lb_desc.dtype, addressed as
 *   rho[x*rho_stride[0] + y*rho_stride[1] + z*rho_stride[2]]
 *   u  [c*u_comp_stride + x*u_stride[0] + y*u_stride[1] + z*u_stride[2]]
 * with strides in ELEMENTS; a stride of 0 broadcasts that axis
 * (equilibrium_boundary_pu.py:21-32; tests/boundary/test_equilibrium_bc_pu.py:125-148).
 *
 * The two outlet kinds act on a whole boundary PLANE of the domain (`direction`), in place and
 * whatever no_collision_mask says there, exactly like the reference objects:
 *   LB_BND_ABB_OUTLET   AntiBounceBackOutlet (anti_bounce_back_outlet.py:71-86): for every q with
 *       e_q . direction = 1:  f_opp(q) = -f_q + w_q rho (2 + (e_q.u_w)^2 / cs^4 - |u_w|^2 / cs^2),
 *       u_w = u + (u - u_neighbour) / 2, with rho, u of the plane node and u of the node next to it.
 *   LB_BND_EQ_OUTLET_P  EquilibriumOutletP (equilibrium_outlet_p.py:63-71): f = feq(value, u_neighbour).
 * Their streaming masks (which populations of the plane keep their local value) come through
 * lb_fields.streaming_bits like any other no_streaming_mask.  The neighbour's velocity is taken
 * from its state at the start of the step, corrected for a bounce-back or equilibrium boundary of
 * LOWER index on that node (collisions conserve rho and u); a lower-index OUTLET plane crossing the
 * neighbour plane is not supported (the host mirror refuses it). */
typedef struct lb_boundary {
  int32_t kind;
  int32_t direction;   /* outlet kinds: +-(lattice axis + 1), e.g. +1 = the x = nx-1 side, -2 = the y = 0 side */
  const void* rho;
  const void* u;
  int64_t rho_stride[3];
  int64_t u_stride[3];
  int64_t u_comp_stride;
  double value;        /* LB_BND_EQ_OUTLET_P: rho_outlet (lattice units) */
} lb_boundary;

/* z-slab decomposition across GPUs (BASELINE.json: slabs along the contiguous z axis; the
 * reference itself is single-device).  With lb_fields.halo set, the
 * populations that leave the local slab through z = nz-1 (ez = +1, "hi" face) or z = 0
 * (ez = -1, "lo" face) are not wrapped around locally.  Four modes:
 *
 * LB_HALO_PACKED  the step kernel stores them into send_hi[0] / send_lo[0], packed as
 *     [zface][lines] (zface = 3 / 5 / 9 directions for D2Q9 / D3Q19 / D3Q27, ascending q;
 *     lines = product of the non-contiguous axes, indexed by the DESTINATION line).  The caller
 *     moves them (NCCL send/recv) and scatters what it received with lb_unpack_halo() -- or
 *     hands it to the next call as recv_lo[0] / recv_hi[0], which the step kernel then consumes
 *     at load time instead of the boundary planes of f.  n_steps must be 1.
 * LB_HALO_PUSH    (default multi-GPU path) as PACKED, but the library itself moves the halos:
 *     after every step one small kernel copies send_hi -> the +z neighbour's recv_lo and
 *     send_lo -> the -z neighbour's recv_hi through peer-mapped pointers (CUDA IPC, NVLink) with
 *     128-bit stores, then performs the neighbour handshake.  The next step kernel consumes the
 *     received halo straight from recv_* at load time; only at the end of the call is it
 *     scattered into the population buffer so that f is canonical again.  All buffers are
 *     double-buffered by step parity ([s & 1]).
 * LB_HALO_DIRECT  the step kernel stores the crossing populations straight into the neighbours'
 *     population buffers (peer_up_f / peer_down_f, the buffers matching this rank's f_next [0]
 *     and f [1]) followed by the handshake.  No staging at all, but 4-byte remote stores.
 * LB_HALO_REMOTE  as PUSH without the copy: the step kernel writes its packed z-face elements,
 *     staged per thread block into contiguous runs, straight into the neighbours' recv_lo /
 *     recv_hi (peer_up_recv_lo / peer_down_recv_hi) and a one-block kernel does the handshake;
 *     received halos are consumed at load time and scattered once at the end of the call.
 *     No send buffers, no isolated remote stores: the transport that scales to 1024^2 faces.
 * LB_HALO_LL      (default multi-GPU path) as REMOTE, but WITHOUT any handshake: every packed halo element
 *     travels as a (payload, tag) pair written with one 8-byte store (fp64: two pairs in one 16-byte store),
 *     tag = epoch + s + 1 for step s of the call -- the global index of the producing time step.  The
 *     receiving step kernel loads its incoming elements together with its populations and re-polls only the
 *     ones whose tag is not there yet, so there is no kernel between two steps, no flag, no fence, and no wait
 *     on the critical path in the steady state.  Receive buffers are TRIPLE-buffered ([g % 3], g = the global
 *     step index): a block that writes slot g % 3 at step g has seen a tag g-1 from that neighbour, i.e. the
 *     neighbour has started step g-1 and therefore finished step g-2, the last reader of that slot's previous
 *     contents.  recv_* / peer_*_recv_* hold LLWord elements: 8 bytes per fp32 value, 16 per fp64 value
 *     (lb_halo_ll_bytes()).  The buffers must be zero before the first call (tag 0 is never sent), epoch = the
 *     number of time steps taken so far; the caller advances epoch by n_steps.  wait_slots[2] is the error flag.
 *
 * Masked streaming (lb_fields.streaming_bits) works across slabs with every mode except LB_HALO_DIRECT: the
 * sender always sends the element that crosses a z face, the RECEIVER drops it where its own node keeps the
 * population (the destination mask of _simulation.py:171-174 is the receiver's), at load time and in
 * lb_unpack_halo_ex().
 *
 * Handshake (PUSH, DIRECT, REMOTE): the rank stores the step counter into both neighbours' wait slots
 * (signal_up / signal_down, peer-mapped) and spins until both of its own slots reached it, so
 * that step t+1 never reads a halo its neighbour has not delivered, nor overwrites one its
 * neighbour still reads.  A call performs n_steps + 1 handshakes with the values
 * epoch+1 .. epoch+n_steps+1; the caller advances epoch by n_steps + 1.  A wait that exceeds
 * ~10 s sets wait_slots[2] = 1 and gives up instead of hanging the GPU. */
enum { LB_HALO_PACKED = 1, LB_HALO_PUSH = 2, LB_HALO_DIRECT = 3, LB_HALO_REMOTE = 4, LB_HALO_LL = 5 };

typedef struct lb_halo {
  int32_t mode;
  uint32_t epoch;
  void* send_hi[2];            /* local packed send buffers                                 */
  void* send_lo[2];
  void* recv_lo[3];            /* local: written by the -z neighbour (its hi face); [2] LL only */
  void* recv_hi[3];            /* local: written by the +z neighbour (its lo face)          */
  void* peer_up_recv_lo[3];    /* peer-mapped: the +z neighbour's recv_lo                   */
  void* peer_down_recv_hi[3];  /* peer-mapped: the -z neighbour's recv_hi                   */
  void* peer_up_f[2];          /* DIRECT: +z neighbour's buffers matching f_next [0], f [1] */
  void* peer_down_f[2];
  uint32_t* signal_up;         /* peer-mapped: +z neighbour's wait_slots[0]                 */
  uint32_t* signal_down;       /* peer-mapped: -z neighbour's wait_slots[1]                 */
  uint32_t* wait_slots;        /* local [8]: [0] written by the -z neighbour, [1] by the +z
                                  neighbour, [2] timeout flag, [3] block counter (scratch)  */
  uint32_t flags;              /* LB_HALO_LL only: LB_HALO_F_* (drivers that interleave the steps of
                                  several slabs on ONE device, e.g. the single-GPU tests)   */
  uint32_t _pad;
} lb_halo;
enum {
  LB_HALO_F_NO_UNPACK = 1,     /* do not scatter the last received halo into f at the end of the call
                                  (the caller does: lb_unpack_halo_ex with ll_tag = epoch + n_steps) */
  LB_HALO_F_CONSUME_FIRST = 2  /* f is not canonical: the first step of the call consumes the halo of
                                  tag `epoch` at load time, like every later step             */
};

typedef struct lb_fields {
  void* f;                          /* populations, state in / (n even) state out        */
  void* f_next;                     /* second buffer (flow.f_next); (n odd) state out    */
  const uint8_t* no_collision_mask; /* [*res] uint8, value = boundary index, or NULL     */
  const uint32_t* streaming_bits;   /* [*res] packed no_streaming_mask (bit q = nsm[q,x]),
                                       from lb_pack_streaming_mask(), or NULL            */
  const lb_boundary* boundaries;    /* n_boundaries entries (host memory)                */
  int32_t n_boundaries;
  int32_t _pad;
  double* obs_out;                  /* device, [n_steps][LB_OBS_COUNT] or NULL: energy, max|u|^2
                                       and mass of the state each step CONSUMES (f_t before
                                       collide), fused into the step kernel               */
  void* obs_scratch;                /* device scratch, lb_obs_scratch_bytes() bytes, if obs_out */
  void* u_out;                      /* device [d][*res] or NULL: velocity of the consumed
                                       state of the LAST step of the call (for enstrophy) */
  const lb_halo* halo;              /* NULL: the slab is periodic in z by itself             */
} lb_fields;

LB_API int lb_abi_version(void);
LB_API const char* lb_last_error(void);

/* Advance n_steps time steps (collide -> boundaries -> stream each).  Buffers ping-pong:
 * after the call the new state is in `f_next` if n_steps is odd, in `f` if even.
 * All launches go to `stream` (a cudaStream_t; NULL = legacy default stream); no host sync. */
LB_API int lb_step(const lb_desc* desc, const lb_fields* fields, int n_steps, void* stream);

/* Collision (where mask == 0) and boundaries (where mask == i) only, no streaming:
 * out may alias in.  Implements Collision.__call__ / Boundary.__call__. */
LB_API int lb_collide(const lb_desc* desc, const void* f_in, void* f_out,
               const uint8_t* no_collision_mask, const lb_boundary* boundaries,
               int n_boundaries, void* stream);

/* Pure streaming (_simulation.py:160-175), f_out != f_in.  streaming_bits must be NULL: destination-masked
 * streaming goes through lb_step with LB_COLL_NONE, which takes the masks (LB_EINVAL otherwise). */
LB_API int lb_stream(const lb_desc* desc, const void* f_in, void* f_out,
              const uint32_t* streaming_bits, void* stream);

/* bits[x] = OR_q (nsm[q][x] == 1) << q   (nsm: uint8 [q][*res]); bit 31 of a word is a hint written by
 * this function ("this node or one of its 26 periodic neighbours keeps a population") from which the
 * step kernel decides per warp whether streaming needs masked stores: lb_fields.streaming_bits must
 * come from this function. */
LB_API int lb_pack_streaming_mask(const lb_desc* desc, const uint8_t* nsm, uint32_t* bits, void* stream);

/* rho_out [*res] and/or u_out [d][*res] (either may be NULL) */
LB_API int lb_moments(const lb_desc* desc, const void* f, void* rho_out, void* u_out, void* stream);

/* The fields an output reporter writes, in ONE pass over f: out [1 + d][*reversed(res)] =
 *   out[0] = (rho - rho0) * p_scale     pressure in physical units   (lettuce/_flow.py:183-186, _unit.py:82-83)
 *   out[1 + c] = u_c * u_scale          velocity in physical units   (lettuce/_flow.py:188-190, _unit.py:52-53)
 * stored in the point order of a VTK file, lattice x FASTEST (pyevtk packs every array in Fortran order),
 * i.e. out[k][z][y][x] for a 3-D lattice f[q][x][y][z].  Replaces the field extraction of
 * `VTKReporter.__call__` (lettuce/ext/_reporter/vtk_reporter.py:29-47): flow.u_pu, flow.p_pu, the numpy
 * conversions and the reordering.  p_scale = cs^2 * p_char_pu / p_char_lu, u_scale = u_char_pu / u_char_lu. */
LB_API int lb_output_fields(const lb_desc* desc, const void* f, double rho0, double p_scale, double u_scale,
                     void* out, void* stream);

/* Initial populations from (rho, u) in lattice units -- replaces `Flow.initialize`
 * (lettuce/_flow.py:106-122: f = equilibrium(rho, u)) and, with `with_fneq` != 0, its first-order
 * non-equilibrium correction `initialize_f_neq` (lettuce/_flow.py:309-336):
 *   f_q = feq_q(rho, u) - w_q * sum_ab (desc->tau * rho * d_b u_a / cs^2) * (e_qa e_qb - cs^2 delta_ab)
 * with the 6th-order central difference of `torch_gradient` (lettuce/util/utility.py:37-99, dx = 1).
 * rho [*res_ext] and u [d][*res_ext] carry `ghost` extra planes on both sides of the LAST axis
 * (res_ext = res with res[d-1] + 2 ghost; ghost must be 0 or >= 3): ghost = 0 differentiates
 * periodically, ghost >= 3 reads the neighbouring slab's values from the ghost planes (z-slabs).
 * f [q][*res] receives the interior.  Nothing larger than (1 + d) / q of f is ever materialised,
 * which is what lets a 1024^3 lattice be initialised slab by slab. */
LB_API int lb_initialize(const lb_desc* desc, const void* rho, const void* u, int ghost, int with_fneq,
                  void* f, void* stream);

/* Standalone reduction of LB_OBS_* over state f into obs_out[LB_OBS_COUNT] (device, double).
 * `want_enstrophy` != 0 additionally needs u_scratch [d][*res] of dtype, which receives the velocity field;
 * 1 also runs the enstrophy pass over it, 2 leaves that to the caller (lb_enstrophy / lb_enstrophy_slab). */
LB_API size_t lb_obs_scratch_bytes(const lb_desc* desc);
LB_API int lb_observables(const lb_desc* desc, const void* f, int want_enstrophy, void* u_scratch,
                   void* obs_scratch, double* obs_out, void* stream);

/* Enstrophy (lattice units, dx = 1) of a velocity field u [d][*res] (e.g. lb_fields.u_out)
 * into obs_out[LB_OBS_ENSTROPHY]. */
LB_API int lb_enstrophy(const lb_desc* desc, const void* u, void* obs_scratch, double* obs_out, void* stream);
/* The same over one z-slab: u is [d][*res_ext] with `ghost` (0 or >= 3) planes of the neighbouring slabs on
 * both sides of the last axis (res_ext = res with res[d-1] + 2 ghost); the sum runs over the interior nodes.
 * The slab driver fills the ghost planes (3-plane velocity halo) and adds the ranks' sums. */
LB_API int lb_enstrophy_slab(const lb_desc* desc, const void* u_ext, int ghost, void* obs_scratch, double* obs_out,
                      void* stream);

/* lb_observables without any allocation, the caller's included (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernel and same result as lb_observables(desc, f, 0, NULL, scratch,
 * obs_out, stream): obs_out[LB_OBS_ENERGY / MAXU2 / MASS] (device, double; LB_OBS_ENSTROPHY is left untouched).  The
 * reduction scratch is a static __device__ array inside the library, sized for lattices of at most
 * LB_STATIC_OBS_MAX_NODES nodes (about 3.0 MiB; every GPU that loads the library holds it until the process exits);
 * a larger lattice returns LB_EUNSUPPORTED.  The call makes no cudaMalloc / cudaFree and calls no allocating
 * library routine; it launches on `stream` of the CURRENT device (f must live there) and does not synchronise.
 * All calls on a device share the one array: calls that may overlap on the GPU (different streams, no ordering
 * between them) race on it -- serialise them, or use lb_observables with a scratch buffer per stream. */
#define LB_STATIC_OBS_MAX_NODES 16777216ull /* 256^3 */
LB_API size_t lb_static_obs_max_nodes(void);
LB_API int lb_observables_static(const lb_desc* desc, const void* f, double* obs_out, void* stream);

/* Host-buffer convenience entry (end-to-end path): copies f_host -> device, runs n_steps,
 * copies the result back into f_host, synchronises.  masks / boundary arrays are HOST
 * pointers here as well (nsm is the raw uint8 [q][*res] mask or NULL). */
LB_API int lb_run_host(const lb_desc* desc, void* f_host, const uint8_t* ncm_host,
                const uint8_t* nsm_host, const lb_boundary* boundaries_host_ptrs,
                int n_boundaries, int n_steps, double* obs_host /* [LB_OBS_COUNT] or NULL */);

/* ---- multi-GPU plumbing (one process per GPU) --------------------------------------------- */
/* scatter packed halos received from the -z neighbour (its hi face -> our z = 0 plane) and the
 * +z neighbour (its lo face -> our z = nz-1 plane) into f; either may be NULL */
LB_API int lb_unpack_halo(const lb_desc* desc, void* f, const void* from_down, const void* from_up, void* stream);
/* the same with the receiver's packed no-streaming bits (elements whose population the node keeps are
 * dropped; NULL = none) and, when ll_tag != 0, LLWord buffers: every element is polled until its tag equals
 * ll_tag (bounded; *ll_err = 1 on timeout) */
LB_API int lb_unpack_halo_ex(const lb_desc* desc, void* f, const void* from_down, const void* from_up,
                      const uint32_t* streaming_bits, uint32_t ll_tag, uint32_t* ll_err, void* stream);
LB_API size_t lb_halo_ll_bytes(const lb_desc* desc);   /* bytes of ONE LL receive buffer: zface * lines * word */
LB_API int lb_zface_count(int stencil);  /* directions per z face: 3 / 5 / 9 */
/* device memory that other processes of the node can map (cudaMalloc + CUDA IPC) */
LB_API int lb_ipc_alloc(size_t bytes, void** dptr);
LB_API int lb_ipc_free(void* dptr);
LB_API int lb_ipc_export(void* dptr, unsigned char handle[64]);
LB_API int lb_ipc_open(const unsigned char handle[64], void** dptr);
LB_API int lb_ipc_close(void* dptr);

/* Introspection for tests / roofline arithmetic */
LB_API int lb_stencil_q(int stencil);
LB_API size_t lb_bytes_per_update(const lb_desc* desc); /* 2 * Q * sizeof(dtype) */
LB_API int lb_vector_width(const lb_desc* desc, int has_streaming_bits); /* nodes per thread chosen for aligned fields */
/* How a call of n_steps time steps on a small 2-D lattice would be run by the tile-resident kernel (state in shared
 * memory, csrc/lb_resident.cuh) on a device with `sms` multiprocessors and `smem_max` bytes of shared memory per block:
 * out = {applies, tile_y, tile_z, rounds, steps_base, steps_extra, region elements per population, shared bytes};
 * round r advances steps_base + (r < steps_extra) time steps.  Pure host arithmetic (tests). */
LB_API int lb_plan_resident(const lb_desc* desc, int n_steps, int sms, size_t smem_max, int out[8]);

#ifdef __cplusplus
}
#endif
#endif /* LETTUCE_B200_H */
