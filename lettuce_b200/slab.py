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
  'auto'  (default on CUDA) 'p2p' for faces below 512 x 512 lines, 'p2p-packed' from there on.
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
from ._flows import TaylorGreenVortex
from ._reporter import ObservableReporter
from ._simulation import Simulation

AUTO_PACKED_MIN_LINES = 512 * 512   # faces (lines per z plane) from which 'auto' picks 'p2p-packed'


def auto_transport(resolution) -> str:
    """'p2p' (direct peer stores) for small z faces, 'p2p-packed' for large ones.  Measured (DESIGN.md
    section 6): 256^2 faces 95.6 % ('p2p') vs 93.8 % ('p2p-packed') of the single-GPU rate at 8 GPUs;
    1024^2 faces 83 % vs 98.5 %."""
    lines = 1
    for n in resolution[:-1]:
        lines *= int(n)
    return 'p2p' if lines < AUTO_PACKED_MIN_LINES else 'p2p-packed'

__all__ = ['slab_range', 'zface_populations', 'SlabTaylorGreenVortex', 'SlabSimulation']


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
                 local_stepper=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.up = (self.rank + 1) % self.world
        self.down = (self.rank - 1) % self.world
        self.local_stepper = local_stepper
        if transport is None:
            transport = 'host' if local_stepper is not None else 'auto'
        if transport == 'auto':
            transport = auto_transport(flow.resolution)
        if transport not in ('p2p-push', 'p2p-packed', 'p2p', 'nccl', 'host'):
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
            if self.no_streaming_mask is not None and bool(self.no_streaming_mask.any()):
                raise NotImplementedError("no_streaming_mask is not supported across slabs")
        flow._obs_allreduce = self._reduce_obs
        self.hi_q, self.lo_q = zface_populations(flow.stencil)
        self.lines = 1
        for n in flow.resolution[:-1]:
            self.lines *= n
        if self.world > 1:
            if transport in ('p2p-push', 'p2p-packed'):
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
                nat.check(nat.lib().lb_unpack_halo(desc, self.flow.f.data_ptr(), self.recv_down.data_ptr(),
                                                   self.recv_up.data_ptr(), nat.current_stream_ptr(dev)))

    def _host_step(self):
        """local update by the injected stepper, which follows the kernel's contract: it returns
        the locally streamed slab (periodic in z by itself) -- the planes that wrapped around are
        exactly the populations owed to the neighbours."""
        flow = self.flow
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

    def _run_segment(self, n, fused):
        if any(r.native_slot == nat.LB_OBS_ENSTROPHY for r in fused) and self.world > 1:
            raise NotImplementedError("Enstrophy across slabs needs a 3-plane velocity halo, which the slab "
                                      "transports do not carry")
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
        mlups_local = Simulation.__call__(self, num_steps)
        if self.world > 1 and self.transport in ('p2p-push', 'p2p-packed', 'p2p'):
            if int(self._flags[2].item()) != 0:
                raise RuntimeError("halo handshake timed out: a neighbour rank did not reach the same step")
        return mlups_local * self.world

    def close(self):
        if self.world > 1 and self.transport in ('p2p-push', 'p2p-packed', 'p2p'):
            torch.cuda.synchronize(self.context.device)
            dist.barrier(group=self.group)
            for r, ptrs in self._peer.items():
                for p in ptrs:
                    nat.lib().lb_ipc_close(p)
            self._peer = {}
