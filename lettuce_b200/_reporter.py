"""Reporters and the observables on the hot path.  Mirrors reference/lettuce/_simulation.py:30-38
(Reporter) and ext/_reporter/observable_reporter.py:17-68,161-199 (Observable, the three
reductions, ObservableReporter).

Each of IncompressibleKineticEnergy / MaximumVelocity / Enstrophy is a reduction the native
engine produces in LATTICE units (lb_observables, or fused into the step kernel through
lb_fields.obs_out); `from_lattice` applies the reference's PU scaling on the host.
`Simulation.__call__` feeds fused values straight into `ObservableReporter.record`, so a run
with these reporters never materialises u or rho."""
import sys
from abc import ABC, abstractmethod
from typing import Optional

import numpy as np
import torch

from . import _native as nat

__all__ = ['Reporter', 'Observable', 'ObservableReporter', 'MaximumVelocity',
           'IncompressibleKineticEnergy', 'Enstrophy', 'Mass']


class Reporter(ABC):
    interval: int

    def __init__(self, interval: int):
        self.interval = interval

    @abstractmethod
    def __call__(self, simulation: 'Simulation'):
        ...


class Observable(ABC):
    #: LB_OBS_* slot when the native engine computes this observable, else None
    native_slot: Optional[int] = None

    def __init__(self, flow: 'Flow'):
        self.context = flow.context
        self.flow = flow

    @abstractmethod
    def __call__(self, f: Optional[torch.Tensor] = None):
        ...

    def from_lattice(self, value_lu: float) -> float:
        raise NotImplementedError

    def _native(self, want_enstrophy=False) -> float:
        """standalone native reduction over flow.f -> lattice-unit value of this observable"""
        flow = self.flow
        f = flow.f
        if not f.is_cuda:
            raise nat.NativeError("lettuce_b200 observables run on CUDA tensors only (no CPU fallback)")
        f = f.contiguous()
        desc = nat.make_desc(type(flow.stencil).__name__, f.dtype, flow.resolution)
        L = nat.lib()
        scratch = torch.empty(L.lb_obs_scratch_bytes(desc), dtype=torch.uint8, device=f.device)
        out = torch.zeros(nat.LB_OBS_COUNT, dtype=torch.float64, device=f.device)
        u = torch.empty([flow.stencil.d, *flow.resolution], dtype=f.dtype, device=f.device) \
            if want_enstrophy else None
        # the slab driver replaces the enstrophy pass (3-plane velocity halo from the neighbouring ranks)
        enstrophy_ = getattr(flow, '_obs_enstrophy', None) if want_enstrophy else None
        with torch.cuda.device(f.device):
            nat.check(L.lb_observables(desc, f.data_ptr(), (2 if enstrophy_ else 1) if want_enstrophy else 0,
                                       nat.ptr(u), scratch.data_ptr(), out.data_ptr(),
                                       nat.current_stream_ptr(f.device)))
        if enstrophy_ is not None:
            enstrophy_(desc, u, out, scratch)
        reduce_ = getattr(flow, '_obs_allreduce', None)   # set by the slab driver: global reduction
        if reduce_ is not None:
            out = reduce_(out[None, :])[0]
        return float(out[self.native_slot].item())

    def _as_tensor(self, value: float) -> torch.Tensor:
        return torch.tensor(value, dtype=self.context.dtype, device=self.context.device)


class MaximumVelocity(Observable):
    """max_nodes |u_pu|  (observable_reporter.py:27-31)"""
    native_slot = nat.LB_OBS_MAXU2

    def from_lattice(self, value_lu):
        return self.flow.units.convert_velocity_to_pu(value_lu ** 0.5)

    def __call__(self, f=None):
        return self._as_tensor(self.from_lattice(self._native()))


class IncompressibleKineticEnergy(Observable):
    """dx^d * E_pu(sum_nodes 0.5 |u|^2)  (observable_reporter.py:34-42)"""
    native_slot = nat.LB_OBS_ENERGY

    def from_lattice(self, value_lu):
        units = self.flow.units
        dx = units.convert_length_to_pu(1.0)
        return units.convert_incompressible_energy_to_pu(value_lu) * dx ** self.flow.stencil.d

    def __call__(self, f=None):
        return self._as_tensor(self.from_lattice(self._native()))


class Enstrophy(Observable):
    """dx^d * sum |curl u_pu|^2 with 6th-order periodic differences (observable_reporter.py:45-68).
    The kernel works in lattice units with dx = 1: curl_pu = curl_lu * (u_pu/u_lu) / dx."""
    native_slot = nat.LB_OBS_ENSTROPHY

    def from_lattice(self, value_lu):
        units = self.flow.units
        dx = units.convert_length_to_pu(1.0)
        c = units.convert_velocity_to_pu(1.0)
        return value_lu * (c / dx) ** 2 * dx ** self.flow.stencil.d

    def __call__(self, f=None):
        return self._as_tensor(self.from_lattice(self._native(want_enstrophy=True)))


class Mass(Observable):
    """Total mass in lattice units, as the reference defines it: the sum of f over the interior
    `[..., 1:-1, 1:-1]` minus the masked part (observable_reporter.py:140-158).  Not a fused
    reduction (it slices f); evaluated with torch on the device tensor."""

    def __init__(self, flow, no_mass_mask=None):
        super().__init__(flow)
        self.mask = no_mass_mask

    def __call__(self, f=None):
        f = self.flow.f if f is None else f
        mass = f[..., 1:-1, 1:-1].sum()
        if self.mask is not None:
            mass -= (f * self.mask.to(dtype=torch.float)).sum()
        return mass


class ObservableReporter(Reporter):
    """Appends `[i, t_pu, value...]` to `out` (a list) or prints it, every `interval` steps
    (observable_reporter.py:161-199)."""

    def __init__(self, observable, interval=1, out=sys.stdout):
        super().__init__(interval)
        self.observable = observable
        self.out = [] if out is None else out
        self._parameter_name = observable.__class__.__name__
        print('steps    ', 'time    ', self._parameter_name)

    @property
    def native_slot(self):
        """slot of the fused reduction that feeds this reporter, or None"""
        return getattr(self.observable, 'native_slot', None)

    def record(self, simulation, i, observed):
        entry = [i, simulation.units.convert_time_to_pu(i)] + list(observed)
        if isinstance(self.out, list):
            self.out.append(entry)
        else:
            print(*entry, file=self.out)

    def record_lattice(self, simulation, i, value_lu):
        """entry point of the fused path: value in lattice units from the step kernel"""
        v = self.observable.from_lattice(value_lu)
        # round through the context dtype as the reference's tensor -> numpy conversion does.  (Not through a torch
        # tensor per value: with an every-step reporter that was ~5 us of host time per step and reporter, as much
        # as half a time step of a 2048 x 1024 lattice.)
        dtype = self.observable.context.dtype
        if dtype == torch.float32:
            v = float(np.float32(v))
        elif dtype != torch.float64:
            v = torch.tensor(v, dtype=dtype).item()
        self.record(simulation, i, [v])

    def __call__(self, simulation: 'Simulation'):
        if simulation.flow.i % self.interval == 0:
            observed = self.observable.context.convert_to_ndarray(
                self.observable(simulation.flow.f))
            assert len(observed.shape) < 2
            observed = [observed.item()] if len(observed.shape) == 0 else observed.tolist()
            self.record(simulation, simulation.flow.i, observed)
