"""z-slab decomposition across the GPUs of one node: one process per GPU (torchrun), each rank
owns a contiguous range of the last (contiguous) axis, periodic ring of neighbours.  New work
defined by BASELINE.json -- the reference is single-device.

Transports behind one class:
  'p2p-push'  (LB_HALO_PUSH) the step kernel packs the populations that cross a z
          face; one small kernel per step pushes them into the neighbours' receive buffers through
          CUDA-IPC peer pointers over NVLink (128-bit stores) and does the neighbour handshake;
          the next step kernel consumes the received halo at load time.  No host involvement
          between steps: simulation(n) is one C call.
  'p2p'   (LB_HALO_DIRECT) the step kernel itself stores the crossing populations
          straight into the neighbours' boundary planes through CUDA-IPC peer pointers over NVLink:
          compute and exchange in ONE kernel, no pack / copy / unpack; a one-block handshake kernel
          follows each step.  Cheapest on small faces, but isolated 4-byte remote stores into planes
          that the receiver is writing too: up to 20 % of a step on 1024^2 faces.
  'p2p-packed'  (LB_HALO_REMOTE) the step kernel stages its packed z-face elements per thread block
          and writes them as contiguous runs straight into the neighbours' receive buffers (peer
          pointers); a one-block handshake kernel follows each step; received halos are consumed at
          load time.  No isolated remote stores: the transport for large faces (1024^2).
  'p2p-ll'  (LB_HALO_LL, the default on CUDA) as 'p2p-packed', but every halo element travels as a (payload, tag)
          pair written with one 8-byte store, tag = the global index of the producing time step, into
          TRIPLE-buffered receive buffers.  The data is its own flag: the receiving step kernel loads its incoming
          elements with its populations and re-polls only those whose tag is not there yet.  No handshake kernel,
          no flags, no fences between steps: simulation(n) is n step kernels back to back on every rank.
  'auto'  (default on CUDA) = 'p2p-ll'.
  'nccl'  (LB_HALO_PACKED) packed send buffers, torch.distributed send/recv, lb_unpack_halo:
          the baseline design north_star names.
  'host'  same exchange logic over any torch.distributed backend with an injected local stepper;
          this is what the world_size-2 gloo tests on CPU drive (tests inject the oracle).
"""
import ctypes as C
from timeit import default_timer as timer
from typing import List, Optional

import torch
import torch.distributed as dist

from . import _native as nat
from ._flows import Obstacle, TaylorGreenVortex
from ._reporter import ObservableReporter
from ._simulation import Simulation

PEER_TRANSPORTS = ('p2p-ll', 'p2p-packed', 'p2p-push', 'p2p')   # halos move through CUDA-IPC peer pointers


def auto_transport(resolution) -> str:
    """The handshake-free 'p2p-ll' transport for every face size (round 1 picked 'p2p' below
    512 x 512 lines and 'p2p-packed' above, both with a handshake kernel per step)."""
    return 'p2p-ll'

__all__ = ['slab_range', 'zface_populations', 'SlabTaylorGreenVortex', 'SlabObstacle', 'SlabSimulation']


def slab_range(n: int, rank: int, world: int):
    """[z0, z1) of rank's slab; equal slabs are required by the direct transport"""
    if n % world != 0:
        raise ValueError(f"the slab axis ({n}) must be divisible by the number of ranks ({world})")
    nl = n // world
    return rank * nl, (rank + 1) * nl


def zface_populations(stencil):
    """(hi, lo): directions leaving through the +z / -z face of a slab, in ascending q --
    the order of the packed halo buffers (lb::zface_index)"""
    hi = [q for q, e in enumerate(stencil.e) if e[-1] == 1]
    lo = [q for q, e in enumerate(stencil.e) if e[-1] == -1]
    return hi, lo


class SlabTaylorGreenVortex(TaylorGreenVortex):
    """The local slab of a global Taylor-Green vortex.  Initialisation evaluates the analytic
    (p, u) on the slab extended by the radius of the 6th-order gradient stencil (3 planes each
    side, periodic indices into the GLOBAL grid) and hands (rho, u) with those ghost planes to the
    engine's `lb_initialize`, which writes the interior populations directly: the populations equal
    the corresponding slice of the global initial state, and nothing larger than (1 + d) / q of the
    slab is materialised (a 1024^3 lattice is initialised slab by slab).  Without CUDA (the gloo
    tests) the same is done with the torch initialisation on the extended slab."""
    GHOST = 3

    def __init__(self, context, global_resolution, reynolds_number, mach_number, stencil,
                 rank: int, world: int, initialize_fneq: bool = True):
        stencil = stencil() if callable(stencil) else stencil
        self.global_resolution = list(global_resolution)
        self.rank, self.world = rank, world
        self.z0, self.z1 = slab_range(self.global_resolution[-1], rank, world)
        self._ghost = self.GHOST if world > 1 else 0
        self._interior = self.global_resolution[:-1] + [self.z1 - self.z0]
        self._extended = self.global_resolution[:-1] + [self.z1 - self.z0 + 2 * self._ghost]
        TaylorGreenVortex.__init__(self, context, list(self._extended), reynolds_number, mach_number, stencil,
                                   initialize_fneq=initialize_fneq)
        self.resolution = list(self._interior)

    def initialize(self):
        g = self._ghost
        p, u = self.initial_pu()                       # on the extended slab (self.grid)
        rho = self.context.convert_to_tensor(self.units.convert_pressure_pu_to_density_lu(p))
        u = self.context.convert_to_tensor(self.units.convert_velocity_to_lu(u))
        self.resolution = list(self._interior)
        if self._native_initialize_ok(rho, u, ghost=g):
            self.f = self._native_initialize(rho, u, ghost=g)
            return
        self.resolution = list(self._extended)         # torch path: initialise the extended slab, crop
        TaylorGreenVortex.initialize(self)
        if g:
            self.f = self.f[..., g:-g].contiguous()
        self.resolution = list(self._interior)

    def make_units(self, reynolds_number, mach_number, resolution):
        return TaylorGreenVortex.make_units(self, reynolds_number, mach_number, self.global_resolution)

    @property
    def grid(self):
        axes = [torch.linspace(0, 2 * torch.pi * (1 - 1 / n), steps=n, device=self.context.device,
                               dtype=self.context.dtype) for n in self.global_resolution]
        g = self._ghost
        idx = torch.arange(self.z0 - g, self.z1 + g, device=self.context.device) % self.global_resolution[-1]
        axes[-1] = axes[-1][idx]
        return torch.meshgrid(*axes, indexing='ij')


class SlabObstacle(Obstacle):
    """The local z-slab of `lt.Obstacle` (obstacle.py:16-125; SURVEY section 8e: "masks are sharded the same
    way"): same units as the undivided flow (they depend on resolution[0] only), the grid of the slab's own
    z range, and the stock boundary list -- equilibrium inlet on x = 0, AntiBounceBackOutlet on the +x side,
    bounce-back on the mask -- built from the LOCAL mask.  All three are node-local or read a neighbour along x,
    so no boundary needs a halo; the no-streaming mask of the outlet is the destination's own and is applied by
    the receiving rank (include/lettuce_b200.h).  `mask` accepts the local slab or the whole lattice."""

    def __init__(self, context, global_resolution, reynolds_number, mach_number, domain_length_x,
                 rank: int, world: int, char_length=1, char_velocity=1, stencil=None, equilibrium=None):
        self.global_resolution = list(global_resolution)
        self.rank, self.world = rank, world
        self.z0, self.z1 = slab_range(self.global_resolution[-1], rank, world)
        local = self.global_resolution[:-1] + [self.z1 - self.z0]
        Obstacle.__init__(self, context, local, reynolds_number, mach_number, domain_length_x,
                          char_length=char_length, char_velocity=char_velocity, stencil=stencil,
                          equilibrium=equilibrium)

    @property
    def mask(self):
        return self._mask

    @mask.setter
    def mask(self, m):
        if list(m.shape) == self.global_resolution and self.world > 1:
            m = m[..., self.z0:self.z1]
        assert list(m.shape) == list(self.resolution), (m.shape, self.resolution)
        self._mask = self.context.convert_to_tensor(m, dtype=torch.bool)

    @property
    def grid(self):
        """coordinates of the slab's nodes in the global lattice"""
        ranges = [torch.arange(n) for n in self.resolution[:-1]] + [torch.arange(self.z0, self.z1)]
        return torch.meshgrid(*(self.units.convert_length_to_pu(r) for r in ranges), indexing='ij')

    @property
    def global_grid(self):
        xyz = tuple(self.units.convert_length_to_pu(torch.arange(n)) for n in self.global_resolution)
        return torch.meshgrid(*xyz, indexing='ij')


class _DeviceMemory:
    """cudaMalloc'ed, IPC-exportable memory exposed to torch through __cuda_array_interface__"""

    def __init__(self, shape, dtype: torch.dtype):
        self.shape, self.dtype = tuple(shape), dtype
        n = 1
        for s in shape:
            n *= s
        self.nbytes = n * torch.empty((), dtype=dtype).element_size()
        p = C.c_void_p()
        nat.check(nat.lib().lb_ipc_alloc(self.nbytes, C.byref(p)))
        self.ptr = p.value
        typestr = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": typestr,
                                         "data": (self.ptr, False), "version": 2}

    def tensor(self, device):
        return torch.as_tensor(self, device=device)

    def handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        nat.check(nat.lib().lb_ipc_export(self.ptr, buf))
        return buf.raw

    def free(self):
        if self.ptr:
            nat.lib().lb_ipc_free(self.ptr)
            self.ptr = None


def _open_handle(handle: bytes) -> int:
    p = C.c_void_p()
    nat.check(nat.lib().lb_ipc_open(handle, C.byref(p)))
    return p.value


class SlabSimulation(Simulation):
    """`Simulation` over one z-slab per rank.  Same call protocol: simulation(num_steps) -> global
    MLUPS; fused reporters receive globally reduced observables."""

    def __init__(self, flow, collision, reporter, transport: Optional[str] = None, group=None,
                 local_stepper=None, host_no_streaming_mask=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.up = (self.rank + 1) % self.world
        self.down = (self.rank - 1) % self.world
        self.local_stepper = local_stepper
        self.host_no_streaming_mask = host_no_streaming_mask
        if transport is None:
            transport = 'host' if local_stepper is not None else 'auto'
        if transport == 'auto':
            transport = auto_transport(flow.resolution)
        if transport not in PEER_TRANSPORTS + ('nccl', 'host'):
            raise ValueError(transport)
        self.transport = transport
        self._epoch = 0
        self._mem = []
        if transport == 'host':
            # host-logic path (tests): no native engine, the injected stepper does the local update
            self.flow, self.collision, self.reporter = flow, collision, reporter
            self.flow.collision = collision
            self.context = flow.context
            self.boundaries = [None]
            self.no_collision_mask = self.no_streaming_mask = None
        else:
            Simulation.__init__(self, flow, collision, reporter)
            if transport == 'p2p' and self.no_streaming_mask is not None and bool((self.no_streaming_mask == 1).any()):
                raise NotImplementedError("no_streaming_mask across slabs needs a packed transport ('p2p-ll', "
                                          "'p2p-packed', 'p2p-push', 'nccl'): the direct peer stores of 'p2p' go "
                                          "into planes whose mask the sender does not have")
            self._check_slab_outlets()
        flow._obs_allreduce = self._reduce_obs
        if self.world > 1 and transport != 'host':
            flow._obs_enstrophy = self._enstrophy
        self.hi_q, self.lo_q = zface_populations(flow.stencil)
        self.lines = 1
        for n in flow.resolution[:-1]:
            self.lines *= n
        if self.world > 1:
            if transport == 'p2p-ll':
                self._setup_ll()
            elif transport in ('p2p-push', 'p2p-packed'):
                self._setup_push()
            elif transport == 'p2p':
                self._setup_direct()
            else:
                self._setup_packed()

    # ---- packed transport (nccl / host) --------------------------------------------------------
    def _setup_packed(self):
        f = self.flow.f
        shape = [len(self.hi_q), self.lines]
        self.send_hi, self.send_lo = f.new_zeros(shape), f.new_zeros(shape)
        self.recv_down, self.recv_up = f.new_zeros(shape), f.new_zeros(shape)

    def _exchange_packed(self):
        """send_hi -> up neighbour's recv_down, send_lo -> down neighbour's recv_up (ring)"""
        ops = [dist.P2POp(dist.isend, self.send_hi, self.up, self.group),
               dist.P2POp(dist.isend, self.send_lo, self.down, self.group),
               dist.P2POp(dist.irecv, self.recv_down, self.down, self.group),
               dist.P2POp(dist.irecv, self.recv_up, self.up, self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()

    # ---- peer-memory transports ------------------------------------------------------------------
    def _share(self, mems):
        """export our IPC allocations, open the neighbours': {rank: [device pointers]}"""
        handles = [None] * self.world
        dist.all_gather_object(handles, [m.handle() for m in mems], group=self.group)
        peer = {}
        for r in {self.up, self.down}:
            peer[r] = [_open_handle(h) for h in handles[r]]
        dist.barrier(group=self.group)
        return peer

    def _setup_push(self):
        flow, dev = self.flow, self.flow.f.device
        shape = [len(self.hi_q), self.lines]
        with torch.cuda.device(dev):
            # local send buffers; IPC-visible receive buffers [lo0, lo1, hi0, hi1] + flags
            self._send = [flow.f.new_zeros(shape) for _ in range(4)]        # hi0 hi1 lo0 lo1
            recv = [_DeviceMemory(shape, flow.f.dtype) for _ in range(4)]   # lo0 lo1 hi0 hi1
            flags = _DeviceMemory([8], torch.int32)
            self._mem = recv + [flags]
            self._recv = [m.tensor(dev) for m in recv]
            self._flags = flags.tensor(dev)
            torch.cuda.synchronize(dev)
            self._peer = self._share(self._mem)

    def _setup_ll(self):
        """three IPC-visible receive buffers per face (tagged words, zero = "nothing sent yet") + the flag word"""
        flow, dev = self.flow, self.flow.f.device
        nbytes = nat.lib().lb_halo_ll_bytes(self._desc())
        with torch.cuda.device(dev):
            recv = [_DeviceMemory([nbytes], torch.uint8) for _ in range(6)]    # lo0 lo1 lo2 hi0 hi1 hi2
            flags = _DeviceMemory([8], torch.int32)
            self._mem = recv + [flags]
            self._recv = [m.tensor(dev) for m in recv]
            self._flags = flags.tensor(dev)
            torch.cuda.synchronize(dev)
            self._peer = self._share(self._mem)

    def _check_slab_outlets(self):
        """an outlet plane normal to the slab axis would read the neighbour rank's plane: not supported"""
        from ._boundary import AntiBounceBackOutlet
        if self.world > 1:
            for b in self.boundaries[1:]:
                if isinstance(b, AntiBounceBackOutlet) and b.direction[-1] != 0:
                    raise NotImplementedError("outlet boundaries normal to the slab axis (the last lattice axis) are "
                                              "not supported across slabs")

    def _setup_direct(self):
        flow, dev = self.flow, self.flow.f.device
        with torch.cuda.device(dev):
            bufs = [_DeviceMemory(flow.f.shape, flow.f.dtype) for _ in range(2)]
            flags = _DeviceMemory([8], torch.int32)
            self._mem = bufs + [flags]
            self._buf = [b.tensor(dev) for b in bufs]
            self._buf[0].copy_(flow.f)
            flow.f, flow.f_next = self._buf[0], self._buf[1]
            self._flags = flags.tensor(dev)
            torch.cuda.synchronize(dev)
            self._peer = self._share(self._mem)

    def _halo_struct(self, epoch):
        flow = self.flow
        h = nat.LbHalo()
        h.epoch = epoch
        if self.transport == 'p2p-ll':
            up, down = self._peer[self.up], self._peer[self.down]
            h.mode = nat.LB_HALO_LL
            for k in range(3):
                h.recv_lo[k], h.recv_hi[k] = self._recv[k].data_ptr(), self._recv[3 + k].data_ptr()
                h.peer_up_recv_lo[k] = up[k]            # our hi face lands in the +z neighbour's recv_lo
                h.peer_down_recv_hi[k] = down[3 + k]    # our lo face lands in the -z neighbour's recv_hi
            h.wait_slots = self._flags.data_ptr()
            return h
        if self.transport in ('p2p-push', 'p2p-packed'):
            up, down = self._peer[self.up], self._peer[self.down]
            h.mode = nat.LB_HALO_PUSH if self.transport == 'p2p-push' else nat.LB_HALO_REMOTE
            for k in range(2):
                h.send_hi[k], h.send_lo[k] = self._send[k].data_ptr(), self._send[2 + k].data_ptr()
                h.recv_lo[k], h.recv_hi[k] = self._recv[k].data_ptr(), self._recv[2 + k].data_ptr()
                h.peer_up_recv_lo[k] = up[k]            # our hi face lands in the +z neighbour's recv_lo
                h.peer_down_recv_hi[k] = down[2 + k]    # our lo face lands in the -z neighbour's recv_hi
            flag_idx = 4
        elif self.transport == 'p2p':
            up, down = self._peer[self.up], self._peer[self.down]
            h.mode = nat.LB_HALO_DIRECT
            # the neighbours store into the two IPC buffers: if flow.f was rebound (Flow.load, a reporter, a
            # .contiguous() copy) move the state back into them instead of silently losing the halos
            ptrs = (self._buf[0].data_ptr(), self._buf[1].data_ptr())
            if flow.f.data_ptr() not in ptrs:
                self._buf[0].copy_(flow.f)
                flow.f, flow.f_next = self._buf[0], self._buf[1]
            elif flow.f_next.data_ptr() not in ptrs or flow.f_next.data_ptr() == flow.f.data_ptr():
                flow.f_next = self._buf[1] if flow.f.data_ptr() == ptrs[0] else self._buf[0]
            cur = 0 if flow.f.data_ptr() == self._buf[0].data_ptr() else 1
            # step s writes into f_next (s even) / f (s odd): the neighbours' matching buffers
            h.peer_up_f[0], h.peer_up_f[1] = up[1 - cur], up[cur]
            h.peer_down_f[0], h.peer_down_f[1] = down[1 - cur], down[cur]
            flag_idx = 2
        else:
            h.mode = nat.LB_HALO_PACKED
            h.send_hi[0], h.send_lo[0] = self.send_hi.data_ptr(), self.send_lo.data_ptr()
            return h
        h.signal_up = up[flag_idx] + 0        # +z neighbour's wait_slots[0] (we are its -z neighbour)
        h.signal_down = down[flag_idx] + 4    # -z neighbour's wait_slots[1]
        h.wait_slots = self._flags.data_ptr()
        return h

    def _fields(self, desc, obs_out=None, u_out=None):
        fl = Simulation._fields(self, desc, obs_out, u_out)
        if self.world > 1:
            self._halo = self._halo_struct(self._epoch)
            fl.halo = C.pointer(self._halo)
        return fl

    def _advance(self, n, obs_out=None, u_out=None):
        if n <= 0:
            return
        if self.transport == 'host':
            for _ in range(n):
                self._host_step()
            return
        if self.world == 1:
            return Simulation._advance(self, n, obs_out, u_out)
        if self.transport == 'p2p-ll':
            Simulation._advance(self, n, obs_out, u_out)
            self._epoch += n          # = time steps taken: the tag of the next step is epoch + 1
            return
        if self.transport in ('p2p-push', 'p2p-packed', 'p2p'):
            Simulation._advance(self, n, obs_out, u_out)
            self._epoch += n + 1
            return
        # nccl: one step per call, exchange + unpack in between
        dev = self.context.device
        for s in range(n):
            Simulation._advance(self, 1, None if obs_out is None else obs_out[s:s + 1],
                                u_out if s == n - 1 else None)
            self._exchange_packed()
            desc = self._desc()
            with torch.cuda.device(dev):
                nat.check(nat.lib().lb_unpack_halo_ex(desc, self.flow.f.data_ptr(), self.recv_down.data_ptr(),
                                                      self.recv_up.data_ptr(), nat.ptr(self._streaming_bits), 0, None,
                                                      nat.current_stream_ptr(dev)))

    def _host_step(self):
        """local update by the injected stepper, which follows the kernel's contract: it returns
        the locally streamed slab (periodic in z by itself) -- the planes that wrapped around are
        exactly the populations owed to the neighbours.  With a no-streaming mask the stepper is a pair
        (collide, stream-without-mask) and the mask is applied HERE, after the exchange: the sender always
        sends, the receiving slab keeps its own post-collision value where its bit is set
        (_simulation.py:171-174), exactly what the kernels do."""
        flow = self.flow
        fstar = None
        if isinstance(self.local_stepper, tuple):
            collide, stream = self.local_stepper
            fstar = collide(flow.f)
            f = stream(fstar)
        else:
            f = self.local_stepper(flow.f)
        if self.world > 1:
            for k, q in enumerate(self.hi_q):
                self.send_hi[k] = f[q][..., 0].reshape(-1)       # wrapped to z=0: belongs to `up`
            for k, q in enumerate(self.lo_q):
                self.send_lo[k] = f[q][..., -1].reshape(-1)      # wrapped to z=nz-1: belongs to `down`
            self._exchange_packed()
            plane = f[0][..., 0].shape
            for k, q in enumerate(self.hi_q):
                f[q][..., 0] = self.recv_down[k].reshape(plane)
            for k, q in enumerate(self.lo_q):
                f[q][..., -1] = self.recv_up[k].reshape(plane)
        if fstar is not None and self.host_no_streaming_mask is not None:
            f = torch.where(self.host_no_streaming_mask == 1, fstar, f)
        flow.f = f

    # ---- reporters: global reductions ------------------------------------------------------------
    def _reduce_obs(self, obs: torch.Tensor):
        """[n, LB_OBS_COUNT] local partials (device) -> global: sums except the max|u|^2 column"""
        if self.world == 1:
            return obs
        t = obs.clone()
        mx = t[:, nat.LB_OBS_MAXU2].clone()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
        t[:, nat.LB_OBS_MAXU2] = mx
        return t

    def _finalize_obs(self, obs):
        return self._reduce_obs(obs).cpu()

    GRADIENT_RADIUS = 4   # planes of velocity halo: the 6th-order central difference (util.torch_gradient,
    #                       utility.py:37-99) needs 3; 4 keeps the interior 16-byte aligned for the vector kernel

    def _exchange_velocity_halo(self, u):
        """u [d, *res] -> u_ext [d, *res[:-1], nz + 2 g]: the g planes next to each z face come from the neighbours
        (a real exchange step, on enstrophy-reporting steps only: torch.distributed send / recv)"""
        g = self.GRADIENT_RADIUS
        if u.shape[-1] < g:
            raise NotImplementedError(f"Enstrophy across slabs needs slabs of at least {g} planes")
        lo, hi = u[..., :g].contiguous(), u[..., -g:].contiguous()
        from_down, from_up = torch.empty_like(hi), torch.empty_like(lo)
        ops = [dist.P2POp(dist.isend, hi, self.up, self.group), dist.P2POp(dist.isend, lo, self.down, self.group),
               dist.P2POp(dist.irecv, from_down, self.down, self.group),
               dist.P2POp(dist.irecv, from_up, self.up, self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        return torch.cat([from_down, u, from_up], dim=-1).contiguous()

    def _enstrophy(self, desc, u_buf, obs_row, scratch=None):
        if self.world == 1:
            return Simulation._enstrophy(self, desc, u_buf, obs_row)
        if min(self.flow.resolution[:-1]) < 3:
            raise NotImplementedError("Enstrophy needs at least 3 nodes per axis (6th-order periodic gradient)")
        dev = self.context.device
        scratch = self._obs_scratch if scratch is None else scratch
        u_ext = self._exchange_velocity_halo(u_buf)
        with torch.cuda.device(dev):
            nat.check(nat.lib().lb_enstrophy_slab(desc, u_ext.data_ptr(), self.GRADIENT_RADIUS,
                                                  scratch.data_ptr(), obs_row.data_ptr(),
                                                  nat.current_stream_ptr(dev)))

    def _run_segment(self, n, fused):
        if self.transport == 'host':
            self._advance(n)
            self.flow.i += n
            return
        return Simulation._run_segment(self, n, fused)

    def __call__(self, num_steps):
        if self.transport == 'host':
            beg = timer()
            self._advance(num_steps)
            self.flow.i += num_steps
            return num_steps * self.flow.f[0].numel() * self.world / 1e6 / (timer() - beg)
        i_before = self.flow.i
        mlups_local = Simulation.__call__(self, num_steps)
        if self.world > 1 and self.transport in PEER_TRANSPORTS:
            if int(self._flags[2].item()) != 0:
                self.flow.i = i_before
                raise RuntimeError("halo exchange timed out: a neighbour rank did not deliver its halo / reach the "
                                   "same step within ~10 s.  The populations of this call are INVALID (steps after "
                                   "the timeout ran on stale halos); flow.i was left at its value before the call.")
        return mlups_local * self.world

    def close(self):
        if self.world > 1 and self.transport in PEER_TRANSPORTS:
            torch.cuda.synchronize(self.context.device)
            dist.barrier(group=self.group)
            for r, ptrs in self._peer.items():
                for p in ptrs:
                    nat.lib().lb_ipc_close(p)
            self._peer = {}
