"""Boundaries on the hot path: BounceBackBoundary (reference/lettuce/ext/_boundary/
bounce_back_boundary.py:10-32) and EquilibriumBoundaryPU (equilibrium_boundary_pu.py:13-46).
Same constructors and mask methods as the reference; `native_descriptor` hands the step kernel
a POD record (kind + broadcastable rho/u fields in LATTICE units) instead of generated code."""
from typing import List, Optional

import torch

from . import _native as nat
from ._flow import Boundary

__all__ = ['BounceBackBoundary', 'EquilibriumBoundaryPU']


class BounceBackBoundary(Boundary):
    def __init__(self, mask: torch.Tensor):
        self._mask = mask

    def __call__(self, flow: 'Flow'):
        return flow.f[flow.stencil.opposite]

    def make_no_streaming_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        return None

    def make_no_collision_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        return self._mask

    def native_available(self) -> bool:
        return True

    def native_descriptor(self, flow: 'Flow') -> dict:
        return dict(kind=nat.LB_BND_BOUNCE_BACK)


class EquilibriumBoundaryPU(Boundary):
    """Sets the populations on the masked nodes to the equilibrium of a prescribed velocity and
    pressure, both given in physical units and converted every time they are used
    (equilibrium_boundary_pu.py:27-29), so user code may change them between calls."""

    def __init__(self, context: 'Context', mask, velocity, pressure=0):
        velocity = [velocity] if not hasattr(velocity, '__len__') else velocity
        self.velocity = context.convert_to_tensor(velocity)
        self.pressure = context.convert_to_tensor(pressure)
        self._mask = mask

    def _lattice_units(self, flow):
        rho = flow.units.convert_pressure_pu_to_density_lu(self.pressure)
        u = flow.units.convert_velocity_to_lu(self.velocity)
        return rho, u

    def __call__(self, flow: 'Flow'):
        rho, u = self._lattice_units(flow)
        feq = flow.equilibrium(flow, rho, u)
        return flow.einsum("q,q->q", [feq, torch.ones_like(flow.f)])

    def make_no_collision_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        return self._mask

    def make_no_streaming_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        return None

    def native_available(self) -> bool:
        return True

    def native_descriptor(self, flow: 'Flow') -> dict:
        rho, u = self._lattice_units(flow)
        dev, dt = flow.f.device, flow.f.dtype
        rho = rho.to(device=dev, dtype=dt)
        u = u.to(device=dev, dtype=dt)
        return dict(kind=nat.LB_BND_EQUILIBRIUM, rho=rho.contiguous(), u=u.contiguous())
