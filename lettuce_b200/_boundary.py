"""Boundaries on the hot path: BounceBackBoundary (reference/lettuce/ext/_boundary/
bounce_back_boundary.py:10-32) and EquilibriumBoundaryPU (equilibrium_boundary_pu.py:13-46), and the
two outlets of the stock `lt.Obstacle` family, AntiBounceBackOutlet (anti_bounce_back_outlet.py:13-104)
and EquilibriumOutletP (equilibrium_outlet_p.py:12-85).
Same constructors and mask methods as the reference; `native_descriptor` hands the step kernel
a POD record (kind + broadcastable rho/u fields in LATTICE units) instead of generated code."""
from typing import List, Optional

import torch

from . import _native as nat
from ._flow import Boundary

__all__ = ['BounceBackBoundary', 'EquilibriumBoundaryPU', 'AntiBounceBackOutlet', 'EquilibriumOutletP']


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


class AntiBounceBackOutlet(Boundary):
    """Lets distributions leave the domain through one side (Krueger et al. 2016, p. 195).
    `direction` names the side: [1, 0, 0] = the last x plane, [0, -1] = the first y plane, ...
    Same constructor, masks and in-place semantics as the reference
    (anti_bounce_back_outlet.py:22-104); the update itself runs inside the step kernel
    (LB_BND_ABB_OUTLET)."""
    _kind = nat.LB_BND_ABB_OUTLET

    def __init__(self, direction: List[int], flow: 'Flow', collision: 'Collision' = None):
        direction = [int(round(float(v))) for v in direction]
        assert len(direction) in [1, 2, 3], \
            f"Invalid direction parameter. Expected direction of of length 1, 2 or 3 but got {len(direction)}."
        assert (direction.count(0) == (len(direction) - 1)) and ((1 in direction) ^ (-1 in direction)), \
            f"Invalid direction parameter. Expected direction with all entries 0 except one 1 or -1 but got {direction}."
        assert len(direction) == flow.stencil.d
        self.direction = direction
        self.stencil = flow.stencil
        # the populations that leave through this side: e_q . direction == 1 (:39-43)
        self.velocities = [q for q, e in enumerate(flow.stencil.e)
                           if sum(int(a) * b for a, b in zip(e, direction)) == 1]
        self.index = []      # the boundary plane ...
        self.neighbor = []   # ... and the plane next to it (:46-57)
        for i in direction:
            if i == 0:
                self.index.append(slice(None)); self.neighbor.append(slice(None))
            elif i == 1:
                self.index.append(-1); self.neighbor.append(-2)
            else:
                self.index.append(0); self.neighbor.append(1)

    def _kept_populations(self) -> List[int]:
        """populations of the plane that streaming must not overwrite (:88-93)"""
        return [int(self.stencil.opposite[q]) for q in self.velocities]

    def make_no_streaming_mask(self, f_shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        m = torch.zeros(size=list(f_shape), dtype=torch.bool, device=context.device)
        m[(self._kept_populations(), *self.index)] = True
        return m

    def make_no_collision_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        m = torch.zeros(size=list(shape), dtype=torch.bool, device=context.device)
        m[tuple(self.index)] = True
        return m

    def native_available(self) -> bool:
        return True

    def native_descriptor(self, flow: 'Flow') -> dict:
        axis = next(a for a, v in enumerate(self.direction) if v != 0)
        return dict(kind=self._kind, direction=(axis + 1) * self.direction[axis], value=0.0)

    def __call__(self, flow: 'Flow'):
        raise NotImplementedError("outlet boundaries are applied inside the native step kernel "
                                  "(lettuce_b200 has no torch stepping path)")


class EquilibriumOutletP(AntiBounceBackOutlet):
    """Equilibrium outlet with constant pressure: the plane is set to the equilibrium of
    `rho_outlet` and the velocity of the neighbouring plane (equilibrium_outlet_p.py:12-85)."""
    _kind = nat.LB_BND_EQ_OUTLET_P

    def __init__(self, direction: List[int], flow: 'Flow', rho_outlet: float = 1.0):
        AntiBounceBackOutlet.__init__(self, direction, flow)
        self.context = flow.context
        self.rho_outlet = self.context.convert_to_tensor(rho_outlet)

    def _kept_populations(self) -> List[int]:
        # everything except the populations that leave (:68-73)
        return [q for q in range(self.stencil.q) if q not in self.velocities]

    def native_descriptor(self, flow: 'Flow') -> dict:
        d = AntiBounceBackOutlet.native_descriptor(self, flow)
        d['value'] = float(self.rho_outlet)
        return d
