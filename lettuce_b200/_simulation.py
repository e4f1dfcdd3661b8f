"""Simulation driver.  Mirrors `lettuce.Simulation` (reference/lettuce/_simulation.py:41-207):
same constructor, same mask construction, same `simulation(num_steps) -> MLUPS` call and the
same per-step seam `simulation._collide_and_stream(simulation)` -- but the seam is always the
native B200 engine (lb_step through the C ABI).  There is no torch / CPU stepping path.

Differences from the reference, all behind the same API:
  * `__call__` batches time steps into single C calls.  Reporters that are fed by fused
    reductions (ObservableReporter over IncompressibleKineticEnergy / MaximumVelocity /
    Enstrophy) receive their values from the step kernel; any other reporter forces a batch
    boundary at every multiple of its `interval` and is called there with `flow.f` valid.
  * no device-wide sync per step (the reference's generated extension syncs every step,
    cuda_native/_template.py:366); one sync at the end of the call for the MLUPS timer.
"""
import ctypes as C
import warnings
from timeit import default_timer as timer
from typing import List, Optional

import torch

from . import _native as nat
from ._collision import Collision
from ._reporter import ObservableReporter, Reporter

__all__ = ['Simulation', 'build_masks']


def build_masks(flow, boundaries, context):
    """no_collision_mask [*res] uint8 (value = boundary index, later index wins) and
    no_streaming_mask [q, *res] uint8 (OR of the boundaries' masks); (None, None) without
    boundaries.  `boundaries` is the [None] + sorted list (_simulation.py:57-86)."""
    if len(boundaries) <= 1:
        return None, None
    ncm_all = context.zero_tensor(flow.resolution, dtype=torch.uint8)
    nsm_all = context.zero_tensor([flow.stencil.q, *flow.resolution], dtype=torch.uint8)
    for i, boundary in enumerate(boundaries[1:], start=1):
        ncm = boundary.make_no_collision_mask(list(flow.f.shape[1:]), context=context)
        if ncm is not None:
            ncm_all[ncm.to(context.device)] = i
        nsm = boundary.make_no_streaming_mask(list(flow.f.shape), context=context)
        if nsm is not None:
            nsm_all |= nsm.to(context.device)
    return ncm_all, nsm_all


class Simulation:
    flow: 'Flow'
    context: 'Context'
    collision: 'Collision'
    boundaries: List['Boundary']
    no_collision_mask: Optional[torch.Tensor]
    no_streaming_mask: Optional[torch.Tensor]
    reporter: List['Reporter']

    def __init__(self, flow: 'Flow', collision: 'Collision', reporter: List['Reporter']):
        self.flow = flow
        self.flow.collision = collision
        self.context = flow.context
        self.collision = collision
        self.reporter = reporter
        # index = position in the list sorted by str(), as the reference does (:57-58)
        self.boundaries = [None] + sorted(flow.boundaries, key=lambda b: str(b))

        if self.context.device.type != 'cuda' or not self.context.use_native:
            raise RuntimeError(
                "lettuce_b200.Simulation runs on the native CUDA engine only: construct the flow "
                "with Context(device='cuda', use_native=True). There is no torch / CPU fallback.")
        nat.lib()  # fail loudly, now, when the shared library is not built

        self.no_collision_mask, self.no_streaming_mask = build_masks(flow, self.boundaries, self.context)

        # ---- native descriptors -----------------------------------------------------------------
        if type(flow.stencil).__name__ not in nat.STENCIL_IDS:
            raise NotImplementedError(f"stencil '{type(flow.stencil).__name__}' has no native kernel "
                                      f"({', '.join(nat.STENCIL_IDS)} have)")
        if not isinstance(collision, Collision) or not collision.native_available():
            raise NotImplementedError(f"collision '{type(collision).__name__}' has no native kernel")
        if flow.equilibrium is not None and not flow.equilibrium.native_available():
            raise NotImplementedError(
                f"equilibrium '{type(flow.equilibrium).__name__}' has no native kernel")
        for b in self.boundaries[1:]:
            if not (b.native_available() and hasattr(b, 'native_descriptor')):
                raise NotImplementedError(f"boundary '{type(b).__name__}' has no native kernel")
        self._check_outlets()
        if len(self.boundaries) - 1 > nat.LB_MAX_BOUNDARIES:
            raise NotImplementedError(f"at most {nat.LB_MAX_BOUNDARIES} boundary objects are supported")

        self._streaming_bits = None
        self._mask_version = None
        self._obs_scratch = None
        self._keepalive = []
        self._collide_and_stream = self._native_seam

    def _check_outlets(self):
        """The kernel takes the velocity next to an outlet plane from that node's state at the start
        of the step, corrected for a lower-index bounce-back / equilibrium boundary on it
        (include/lettuce_b200.h).  What it cannot reproduce is an outlet of LOWER index whose plane
        crosses the neighbour plane of another outlet: refuse that combination."""
        from ._boundary import AntiBounceBackOutlet
        outlets = [(i, b) for i, b in enumerate(self.boundaries[1:], start=1) if isinstance(b, AntiBounceBackOutlet)]
        if outlets and getattr(self.collision, 'force', None) is not None:
            # the reference's outlets read flow.u() AFTER the forced collision (momentum j + a); the kernel takes
            # the neighbour's velocity from its incoming populations, which is only right without a force
            raise NotImplementedError("outlet boundaries together with a forcing scheme are not supported by the "
                                      "native kernel")
        for i, b in outlets:
            if min(self.flow.resolution) < 2:
                raise ValueError("outlet boundaries need at least 2 nodes along every axis")
            axis = next(a for a, v in enumerate(b.direction) if v != 0)
            for j, other in outlets:
                if j < i and next(a for a, v in enumerate(other.direction) if v != 0) != axis:
                    raise NotImplementedError(
                        f"outlet boundary {i} ({type(b).__name__}, direction {b.direction}): the lower-index "
                        f"outlet {j} (direction {other.direction}) crosses its neighbour plane; this ordering "
                        f"is not supported by the native kernel")

    # ------------------------------------------------------------------------------------------
    def step(self, num_steps: int):
        warnings.warn("lt.Simulation.step() is deprecated and will be removed in a future version. "
                      "Instead, call simulation directly: simulation(num_steps)", DeprecationWarning)
        return self(num_steps)

    @property
    def units(self):
        return self.flow.units

    # ---- native plumbing -----------------------------------------------------------------------
    def _desc(self):
        d = self.collision.native_descriptor(self.flow)
        return nat.make_desc(type(self.flow.stencil).__name__, self.flow.f.dtype, self.flow.resolution,
                             d['collision'], d.get('tau', 1.0), d.get('tau_minus', 1.0),
                             d.get('force', nat.LB_FORCE_NONE), d.get('acceleration', ()), d.get('force_tau', 1.0),
                             kernel_flags=getattr(self, 'kernel_flags', 0))

    def _pack_streaming_bits(self, desc):
        """no_streaming_mask [q, *res] uint8 -> one uint32 per node (bit q), once per mask version;
        None when the mask is all zero (then streaming is unmasked)."""
        nsm = self.no_streaming_mask
        if nsm is None:
            return None
        version = (nsm.data_ptr(), nsm._version)
        if self._mask_version != version:
            self._mask_version = version
            if bool((nsm == 1).any().item()):
                bits = torch.empty(self.flow.resolution, dtype=torch.int32, device=nsm.device)
                nat.check(nat.lib().lb_pack_streaming_mask(
                    desc, nsm.contiguous().data_ptr(), bits.data_ptr(),
                    nat.current_stream_ptr(nsm.device)))
                self._streaming_bits = bits
            else:
                self._streaming_bits = None
        return self._streaming_bits

    def _boundary_array(self):
        nb = len(self.boundaries) - 1
        arr = (nat.LbBoundary * max(nb, 1))()
        keep = []
        res = self.flow.resolution
        for k, b in enumerate(self.boundaries[1:]):
            d = b.native_descriptor(self.flow)
            arr[k].kind = d['kind']
            arr[k].direction = int(d.get('direction', 0))
            arr[k].value = float(d.get('value', 0.0))
            if d['kind'] == nat.LB_BND_EQUILIBRIUM:
                rho, u = d['rho'], d['u']
                if u.shape[0] != self.flow.stencil.d:
                    raise ValueError(f"boundary velocity must have {self.flow.stencil.d} components, "
                                     f"got shape {tuple(u.shape)}")
                rho_lead = max(0, rho.dim() - len(res))
                if any(n != 1 for n in rho.shape[:rho_lead]):
                    raise ValueError(f"boundary density of shape {tuple(rho.shape)} does not broadcast")
                rs = nat.broadcast_strides(rho, rho_lead, res)
                us = nat.broadcast_strides(u, 1, res)
                arr[k].rho, arr[k].u = rho.data_ptr(), u.data_ptr()
                for a in range(3):
                    arr[k].rho_stride[a] = rs[a]
                    arr[k].u_stride[a] = us[a]
                arr[k].u_comp_stride = u.stride(0)
                keep += [rho, u]
        return arr, nb, keep

    def _fields(self, desc, obs_out=None, u_out=None):
        flow = self.flow
        if not flow.f.is_contiguous():
            flow.f = flow.f.contiguous()
        cdev = self.context.device
        if flow.f.device.type != cdev.type or (cdev.index is not None and flow.f.device != cdev):
            raise ValueError("flow.f is not on the context device")
        fl = nat.LbFields()
        fl.f, fl.f_next = flow.f.data_ptr(), flow.f_next.data_ptr()
        if self.no_collision_mask is not None:
            fl.no_collision_mask = self.no_collision_mask.data_ptr()
            bits = self._pack_streaming_bits(desc)
            fl.streaming_bits = nat.ptr(bits)
        arr, nb, keep = self._boundary_array()
        fl.boundaries = C.cast(arr, C.POINTER(nat.LbBoundary))
        fl.n_boundaries = nb
        if obs_out is not None:
            if self._obs_scratch is None:
                self._obs_scratch = torch.empty(nat.lib().lb_obs_scratch_bytes(desc), dtype=torch.uint8,
                                                device=self.context.device)
            fl.obs_out = obs_out.data_ptr()
            fl.obs_scratch = self._obs_scratch.data_ptr()
            fl.u_out = nat.ptr(u_out)
        self._keepalive = [arr, keep]
        return fl

    def _advance(self, n, obs_out=None, u_out=None):
        """n native time steps; flow.f / flow.f_next are swapped so that flow.f is the new state"""
        if n <= 0:
            return
        desc = self._desc()
        with torch.cuda.device(self.context.device):
            fl = self._fields(desc, obs_out, u_out)
            nat.check(nat.lib().lb_step(desc, fl, n, nat.current_stream_ptr(self.context.device)))
        if n % 2 == 1:
            self.flow.f, self.flow.f_next = self.flow.f_next, self.flow.f

    def _native_seam(self, *_, **__):
        """the reference's `_collide_and_stream(simulation)`: exactly one time step"""
        self._advance(1)

    # ---- reference-named pieces, each through the engine (kept for API parity) -------------------
    def _collide(self):
        desc = self._desc()
        flow = self.flow
        from ._boundary import AntiBounceBackOutlet
        # an outlet node reads every population of the node next to it: out of place then (lb_collide refuses aliasing)
        out_of_place = any(isinstance(b, AntiBounceBackOutlet) for b in self.boundaries[1:])
        with torch.cuda.device(self.context.device):
            fl = self._fields(desc)
            out = flow.f_next if out_of_place else flow.f
            nat.check(nat.lib().lb_collide(desc, flow.f.data_ptr(), out.data_ptr(),
                                           fl.no_collision_mask, fl.boundaries, fl.n_boundaries,
                                           nat.current_stream_ptr(self.context.device)))
        if out_of_place:
            flow.f, flow.f_next = flow.f_next, flow.f
        return flow.f

    def _stream(self):
        saved = self.collision, self.no_collision_mask, self.boundaries
        # pure streaming = a NoCollision step with only the streaming mask active
        from ._collision import NoCollision
        try:
            self.collision = NoCollision()
            if self.no_collision_mask is not None:
                self.no_collision_mask = torch.zeros_like(self.no_collision_mask)
            self._advance(1)
        finally:
            self.collision, self.no_collision_mask, self.boundaries = saved
        return self.flow.f

    def _report(self):
        for reporter in self.reporter:
            reporter(self)

    # ---- the hot loop -----------------------------------------------------------------------------
    def __call__(self, num_steps):
        beg = timer()
        flow = self.flow
        fused = [r for r in self.reporter if isinstance(r, ObservableReporter) and r.native_slot is not None]
        generic = [r for r in self.reporter if r not in fused]

        if flow.i == 0:
            self._report()          # initial state: every reporter through its own __call__ (:198-199)

        end_i = flow.i + num_steps
        while flow.i < end_i:
            # next index at which some non-fused reporter must see flow.f
            stop = end_i
            for r in generic:
                k = max(1, int(getattr(r, 'interval', 1) or 1))
                stop = min(stop, (flow.i // k + 1) * k)
            self._run_segment(stop - flow.i, fused)
            # at the batch boundary: fused entries were recorded by _run_segment; call the rest
            for r in generic:
                r(self)

        for r in self.reporter:                    # reporters that write files in the background: files of this call
            flush = getattr(r, 'flush', None)      # exist when the call returns
            if callable(flush):
                flush()
        torch.cuda.synchronize(self.context.device)
        end = timer()
        n_nodes = 1
        for n in flow.resolution:     # == flow.rho().numel() (_simulation.py:207) without a moments pass
            n_nodes *= int(n)
        return num_steps * n_nodes / 1e6 / (end - beg)

    def _enstrophy(self, desc, u_buf, obs_row):
        """sum |curl u|^2 of the velocity field u_buf [d, *res] (lattice units) into obs_row[LB_OBS_ENSTROPHY];
        the slab driver overrides this with the 3-plane velocity halo"""
        dev = self.context.device
        if min(self.flow.resolution) < 3:
            raise NotImplementedError("Enstrophy needs at least 3 nodes per axis (6th-order periodic gradient)")
        with torch.cuda.device(dev):
            nat.check(nat.lib().lb_enstrophy(desc, u_buf.data_ptr(), self._obs_scratch.data_ptr(),
                                             obs_row.data_ptr(), nat.current_stream_ptr(dev)))

    def _finalize_obs(self, obs: torch.Tensor) -> torch.Tensor:
        """device [n, LB_OBS_COUNT] -> host; the slab driver all-reduces across ranks here"""
        return obs.cpu()

    def _run_segment(self, n, fused):
        """Advance n steps from flow.i, recording every fused reporter at each index
        i in (flow.i, flow.i + n] with i % interval == 0."""
        flow = self.flow
        i0, i1 = flow.i, flow.i + n
        wanted = {}                      # index -> list of reporters
        for r in fused:
            k = max(1, int(r.interval))
            first = (i0 // k + 1) * k
            for i in range(first, i1 + 1, k):
                wanted.setdefault(i, []).append(r)
        if not wanted:
            self._advance(n)
            flow.i = i1
            return

        dev = self.context.device
        need_u = {i for i, rs in wanted.items() if any(r.native_slot == nat.LB_OBS_ENSTROPHY for r in rs)}
        # the step that CONSUMES f_i yields obs(f_i): indices < i1 are fused, i1 is a standalone pass
        idx = sorted(i for i in wanted if i < i1)
        obs = torch.zeros([len(idx) + 1, nat.LB_OBS_COUNT], dtype=torch.float64, device=dev)
        row = {i: k for k, i in enumerate(idx)}
        u_buf = None
        if need_u:
            u_buf = torch.empty([flow.stencil.d, *flow.resolution], dtype=flow.f.dtype, device=dev)
        desc = self._desc()
        L = nat.lib()
        cur = i0
        while cur < i1:
            if cur in row:
                # run of consecutive fused steps (a step needing u_out ends its run)
                m = 1
                while (cur + m - 1) not in need_u and (cur + m) in row:
                    m += 1
                last = cur + m - 1
                self._advance(m, obs_out=obs[row[cur]:row[cur] + m],
                              u_out=u_buf if last in need_u else None)
                if last in need_u:
                    self._enstrophy(desc, u_buf, obs[row[last]])
                cur += m
            else:
                nxt = min([i for i in idx if i > cur] + [i1])
                self._advance(nxt - cur)
                cur = nxt
        if i1 in wanted:
            if self._obs_scratch is None:
                self._obs_scratch = torch.empty(L.lb_obs_scratch_bytes(desc), dtype=torch.uint8, device=dev)
            with torch.cuda.device(dev):
                # want_enstrophy = 2: only fill u_buf; the enstrophy pass itself goes through the hook below
                nat.check(L.lb_observables(desc, flow.f.data_ptr(), 2 if i1 in need_u else 0, nat.ptr(u_buf),
                                           self._obs_scratch.data_ptr(), obs[len(idx)].data_ptr(),
                                           nat.current_stream_ptr(dev)))
            if i1 in need_u:
                self._enstrophy(desc, u_buf, obs[len(idx)])
            row[i1] = len(idx)
        host = self._finalize_obs(obs).tolist()   # the one device->host sync of the segment; plain floats from here
        flow.i = i1
        for i in sorted(wanted):
            values = host[row[i]]
            for r in wanted[i]:
                r.record_lattice(self, i, values[r.native_slot])
