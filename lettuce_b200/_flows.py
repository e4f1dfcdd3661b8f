"""Flows exercised by the BASELINE configurations: ExtFlow (reference/lettuce/ext/_flows/
_ext_flow.py:8-42), TaylorGreenVortex (taylorgreen.py:16-98) and Obstacle (obstacle.py:16-125).
Host-side set-up code (runs once); same constructor signatures as the reference."""
import warnings
from abc import ABC, abstractmethod
from typing import List, Optional, Union

import numpy as np
import torch

from ._boundary import AntiBounceBackOutlet, BounceBackBoundary, EquilibriumBoundaryPU
from ._equilibrium import QuadraticEquilibrium
from ._flow import Flow
from ._stencil import D1Q3, D2Q9, D3Q19
from ._unit import UnitConversion
from .util import append_axes

__all__ = ['ExtFlow', 'TaylorGreenVortex', 'TaylorGreenVortex2D', 'TaylorGreenVortex3D', 'Obstacle']


class ExtFlow(Flow, ABC):
    """Common constructor: subclasses provide make_resolution / make_units."""

    def __init__(self, context: 'Context', resolution: Union[int, List[int]], reynolds_number,
                 mach_number, stencil: Optional['Stencil'] = None,
                 equilibrium: Optional['Equilibrium'] = None):
        resolution = self.make_resolution(resolution, stencil)
        assert len(resolution) in [1, 2, 3], \
            f"flow supports dimensions 1, 2 and 3 but {len(resolution)} dimensions where requested."
        stencil = stencil or [D1Q3(), D2Q9(), D3Q19()][len(resolution) - 1]
        stencil = stencil() if callable(stencil) else stencil
        equilibrium = equilibrium or QuadraticEquilibrium()
        Flow.__init__(self, context, resolution,
                      self.make_units(reynolds_number, mach_number, resolution), stencil, equilibrium)

    @abstractmethod
    def make_resolution(self, resolution, stencil=None) -> List[int]:
        ...

    @abstractmethod
    def make_units(self, reynolds_number, mach_number, resolution: List[int]) -> 'UnitConversion':
        ...


class TaylorGreenVortex(ExtFlow):
    def __init__(self, context, resolution, reynolds_number, mach_number, stencil=None,
                 equilibrium=None, initialize_fneq: bool = True):
        self.initialize_fneq = initialize_fneq
        if stencil is None and not isinstance(resolution, list):
            warnings.warn("Requiring information about dimensionality! Either via stencil or "
                          "resolution. Setting dimension to 2.", UserWarning)
            self.stencil = D2Q9()
        else:
            self.stencil = stencil() if callable(stencil) else stencil
        ExtFlow.__init__(self, context, resolution, reynolds_number, mach_number, stencil, equilibrium)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        if isinstance(resolution, int):
            return [resolution] * self.stencil.d
        assert len(resolution) in [2, 3], \
            'the resolution of a taylor-green-vortex must be 2- or 3-dimensional!'
        return resolution

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0],
                              characteristic_length_pu=2 * torch.pi,
                              characteristic_velocity_pu=1)

    @property
    def grid(self):
        # like np.linspace(0, 2 pi, n, endpoint=False)  (taylorgreen.py:52-61)
        axes = tuple(torch.linspace(0, 2 * torch.pi * (1 - 1 / n), steps=n,
                                    device=self.context.device, dtype=self.context.dtype)
                     for n in self.resolution)
        return torch.meshgrid(*axes, indexing='ij')

    def initial_pu(self):
        return self.analytic_solution(t=0)

    def analytic_solution(self, t: float):
        """taylorgreen.py:66-94"""
        if t > 0 and self.stencil.d > 2:
            warnings.warn("The analytic solution is only true for the 2D TGV!")
        g = self.grid
        nu = self.context.convert_to_tensor(self.units.viscosity_pu)
        if len(self.resolution) == 2:
            u = torch.stack([torch.cos(g[0]) * torch.sin(g[1]) * torch.exp(-2 * nu * t),
                             -torch.sin(g[0]) * torch.cos(g[1]) * torch.exp(-2 * nu * t)])
            p = -torch.stack([0.25 * (torch.cos(2 * g[0]) + torch.cos(2 * g[1]))
                              * torch.exp(-4 * nu * t)])
        else:
            u = torch.stack([torch.sin(g[0]) * torch.cos(g[1]) * torch.cos(g[2]),
                             -torch.cos(g[0]) * torch.sin(g[1]) * torch.cos(g[2]),
                             torch.zeros_like(g[0])])
            p = torch.stack([1 / 16. * (torch.cos(2 * g[0]) + torch.cos(2 * g[1]))
                             * (torch.cos(2 * g[2]) + 2)])
        return p, u

    @property
    def boundaries(self) -> List['Boundary']:
        return []


def TaylorGreenVortex3D(context, resolution, reynolds_number, mach_number, stencil=None, equilibrium=None):
    warnings.warn("TaylorGreenVortex3D is deprecated. Use TaylorGreenVortex instead", DeprecationWarning)
    return TaylorGreenVortex(context=context, resolution=resolution, reynolds_number=reynolds_number,
                             mach_number=mach_number, stencil=stencil, equilibrium=equilibrium)


def TaylorGreenVortex2D(context, resolution, reynolds_number, mach_number, stencil=None, equilibrium=None):
    warnings.warn("TaylorGreenVortex2D is deprecated. Use TaylorGreenVortex instead", DeprecationWarning)
    return TaylorGreenVortex(context=context, resolution=resolution, reynolds_number=reynolds_number,
                             mach_number=mach_number, stencil=stencil, equilibrium=equilibrium)


class Obstacle(ExtFlow):
    """Flow around an obstacle given by a boolean mask (obstacle.py:16-125), with the reference's
    stock boundary list: equilibrium inlet on x = 0, AntiBounceBackOutlet on the +x side,
    bounce-back on the mask (obstacle.py:107-122)."""

    def __init__(self, context, resolution, reynolds_number, mach_number, domain_length_x,
                 char_length=1, char_velocity=1, stencil=None, equilibrium=None):
        self.char_length_lu = resolution[0] / domain_length_x * char_length
        self.char_length = char_length
        self.char_velocity = char_velocity
        self.resolution = self.make_resolution(resolution, stencil)
        self._mask = torch.zeros(self.resolution, dtype=torch.bool)
        ExtFlow.__init__(self, context, resolution, reynolds_number, mach_number, stencil, equilibrium)

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=self.char_length_lu,
                              characteristic_length_pu=self.char_length,
                              characteristic_velocity_pu=self.char_velocity)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        if isinstance(resolution, int):
            return [resolution] * (stencil.d or self.stencil.d)
        return resolution

    @property
    def mask(self):
        return self._mask

    @mask.setter
    def mask(self, m):
        assert (isinstance(m, (np.ndarray, torch.Tensor))
                and all(m.shape[dim] == self.resolution[dim] for dim in range(self.stencil.d)))
        self._mask = self.context.convert_to_tensor(m, dtype=torch.bool)

    def initial_pu(self):
        p = np.zeros_like(self.grid[0], dtype=float)[None, ...]
        u_char = append_axes(self.units.characteristic_velocity_pu * self._unit_vector(), self.stencil.d)
        u = ~self.mask.cpu() * u_char
        return p, u

    @property
    def grid(self):
        xyz = tuple(self.units.convert_length_to_pu(torch.arange(n)) for n in self.resolution)
        return torch.meshgrid(*xyz, indexing='ij')

    @property
    def boundaries(self):
        x = self.grid[0]
        u_in = self.units.characteristic_velocity_pu * self._unit_vector()
        return [EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                AntiBounceBackOutlet(self._unit_vector().tolist(), self),
                BounceBackBoundary(self.mask)]

    def _unit_vector(self, i=0):
        return torch.eye(self.stencil.d)[i]
