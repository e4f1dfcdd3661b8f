"""The reference's stock flows, i.e. the callers of the hot path: ExtFlow (reference/lettuce/ext/
_flows/_ext_flow.py:8-42), TaylorGreenVortex (taylorgreen.py:16-98), Obstacle (obstacle.py:16-125),
CouetteFlow2D (couette.py:17-82), PoiseuilleFlow2D (poiseuille.py:18-95), DoublyPeriodicShear2D
(doublyshear.py:20-86), DecayingTurbulence (decayingturbulence.py:23-199), Cavity2D
(liddrivencavity.py:14-71) and the `flow_by_name` table (_flow_by_name.py:10-16).
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

__all__ = ['ExtFlow', 'TaylorGreenVortex', 'TaylorGreenVortex2D', 'TaylorGreenVortex3D', 'Obstacle',
           'CouetteFlow2D', 'PoiseuilleFlow2D', 'DoublyPeriodicShear2D', 'DecayingTurbulence',
           'Cavity2D', 'flow_by_name']


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


def _unit_grid(flow, endpoint: bool):
    """meshgrid over [0, 1] per axis (endpoint=True) or [0, 1 - 1/n] (endpoint=False)"""
    axes = tuple(torch.linspace(0, 1 if endpoint else 1 - 1 / n, steps=n,
                                device=flow.context.device, dtype=flow.context.dtype)
                 for n in flow.resolution)
    return torch.meshgrid(*axes, indexing='ij')


class CouetteFlow2D(ExtFlow):
    """couette.py:17-82.  As in the reference the characteristic velocity in PU is the background
    velocity u0 = 0, which makes the PU -> LU velocity conversion singular (the reference's own
    energy-spectrum test skips this flow for that reason, tests/reporter/test_energy_spectrum.py:8-9)."""

    def __init__(self, context, resolution, reynolds_number, mach_number, stencil=None, equilibrium=None):
        self.u0 = 0  # background velocity
        super().__init__(context, resolution, reynolds_number, mach_number, stencil, equilibrium)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        return [resolution] * 2 if isinstance(resolution, int) else resolution

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0], characteristic_length_pu=1,
                              characteristic_velocity_pu=self.u0)

    def analytic_solution(self):
        dvdy = 1 / self.resolution[0]
        x, y = self.grid
        return (dvdy * y + self.u0)[None, ...]

    def initial_pu(self):
        zeros = self.context.zero_tensor(self.resolution)
        return zeros[None, ...], torch.stack([zeros, zeros], dim=0)

    @property
    def grid(self):
        return _unit_grid(self, endpoint=True)

    @property
    def boundaries(self):
        ktop = torch.zeros(self.resolution, dtype=torch.bool)
        ktop[:, 1] = True
        kbottom = torch.zeros(self.resolution, dtype=torch.bool)
        kbottom[:, -1] = True
        return [EquilibriumBoundaryPU(self.context, mask=ktop, velocity=np.array([1.0, 0.0])),
                BounceBackBoundary(kbottom)]


class PoiseuilleFlow2D(ExtFlow):
    """Channel between two bounce-back walls driven by `acceleration` (poiseuille.py:18-95)."""

    def __init__(self, context, resolution, reynolds_number, mach_number, stencil=None,
                 equilibrium=None, initialize_with_zeros=True):
        self.stencil = D2Q9() if stencil is None else (stencil() if callable(stencil) else stencil)
        self.initialize_with_zeros = initialize_with_zeros
        super().__init__(context, resolution, reynolds_number, mach_number, self.stencil, equilibrium)

    def analytic_solution(self, t=0):
        half_lattice_spacing = 0.5 / self.resolution[0]
        x, y = self.grid
        nu = self.units.viscosity_pu
        rho = 1
        ux = (self.acceleration[0] / (2 * rho * nu)
              * ((y - half_lattice_spacing) * (1 - half_lattice_spacing - y)))
        u = torch.stack([ux, self.context.zero_tensor(self.resolution)], dim=0)
        p = y * 0 + self.units.convert_density_lu_to_pressure_pu(rho)
        return p, u

    def initial_pu(self):
        if self.initialize_with_zeros:
            zeros = self.context.zero_tensor(self.resolution)
            return zeros[None, ...], torch.stack(2 * [zeros], dim=0)
        return self.analytic_solution()

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0] - 1, characteristic_length_pu=1,
                              characteristic_velocity_pu=1)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        if isinstance(resolution, list):
            assert len(resolution) == self.stencil.d
        if isinstance(resolution, int):
            resolution = [resolution] * self.stencil.d
        return resolution

    @property
    def grid(self):
        return _unit_grid(self, endpoint=True)

    @property
    def boundaries(self):
        mask = self.context.zero_tensor(self.resolution, dtype=bool)
        mask[:, [0, -1]] = True
        return [BounceBackBoundary(mask=mask)]

    @property
    def acceleration(self):
        return self.context.convert_to_tensor([0.001, 0])


class DoublyPeriodicShear2D(ExtFlow):
    """Two tanh shear layers with a sinusoidal perturbation (doublyshear.py:20-86)."""

    def __init__(self, context, resolution, reynolds_number, mach_number, stencil=None, equilibrium=None,
                 shear_layer_width=80, initial_perturbation_magnitude=0.05, initialize_fneq: bool = True):
        self.initialize_fneq = initialize_fneq
        self.initial_perturbation_magnitude = initial_perturbation_magnitude
        self.shear_layer_width = shear_layer_width
        self.stencil = D2Q9() if stencil is None else (stencil() if callable(stencil) else stencil)
        super().__init__(context, resolution, reynolds_number, mach_number, self.stencil, equilibrium)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        if isinstance(resolution, int):
            return [resolution] * self.stencil.d
        assert len(resolution) == 2, 'expected 2-dimensional resolution'
        return resolution

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0], characteristic_length_pu=1,
                              characteristic_velocity_pu=1)

    def analytic_solution(self, t=0):
        raise NotImplementedError

    def initial_pu(self):
        pert = self.initial_perturbation_magnitude
        w = self.shear_layer_width
        x, y = self.grid
        u1 = torch.where(y > 0.5, torch.tanh(w * (y - 0.25)), torch.tanh(w * (0.75 - y)))
        u2 = pert * torch.sin(2 * torch.pi * (x + 0.25))
        return torch.zeros_like(u1)[None, ...], torch.stack([u1, u2])

    @property
    def grid(self):
        return _unit_grid(self, endpoint=False)

    @property
    def boundaries(self):
        return []


class DecayingTurbulence(ExtFlow):
    """Random solenoidal velocity field with the energy spectrum k^4 exp(-2 (k/k0)^2) in 2-D or
    3-D (decayingturbulence.py:23-199).  The initial field is built on the host with numpy's FFT
    and `np.random.seed(randseed)`, exactly the reference's sequence of draws; `initial_pu` sets
    `units.characteristic_velocity_pu` to the largest speed of the field."""

    def __init__(self, context, resolution, reynolds_number, mach_number, k0=20, ic_energy=0.5,
                 stencil=None, equilibrium=None, initialize_pressure: bool = True,
                 initialize_fneq: bool = True, randseed: Optional[int] = None):
        self.initialize_pressure = initialize_pressure
        self.initialize_fneq = initialize_fneq
        self.randseed = randseed
        self.k0 = k0
        self.ic_energy = ic_energy
        self.wavenumbers = []
        self.spectrum = []
        stencil = stencil or [D1Q3(), D2Q9(), D3Q19()][len(resolution) - 1]
        stencil = stencil() if callable(stencil) else stencil
        if stencil.d != 2:
            self.initialize_pressure = False
        super().__init__(context, resolution, reynolds_number, mach_number, stencil, equilibrium)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        return [resolution] * stencil.d if isinstance(resolution, int) else resolution

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0],
                              characteristic_length_pu=2 * np.pi, characteristic_velocity_pu=None)

    def analytic_solution(self, x, t=0):
        return

    def _generate_wavenumbers(self):
        self.dimensions = tuple(self.resolution)
        frequencies = [np.fft.fftfreq(dim, d=1 / dim) for dim in self.dimensions]
        wavenumber = np.meshgrid(*frequencies)        # numpy's default 'xy' indexing, as the reference
        wavenorms = np.linalg.norm(wavenumber, axis=0)
        self.wavenumbers = np.arange(int(np.max(wavenorms)))
        wavemask = ((wavenorms[..., None] > self.wavenumbers - 0.5)
                    & (wavenorms[..., None] <= self.wavenumbers + 0.5))
        return wavenorms, wavenumber, wavemask

    def _generate_spectrum(self):
        wavenorms, wavenumber, wavemask = self._generate_wavenumbers()
        ek = wavenorms ** 4 * np.exp(-2 * (wavenorms / self.k0) ** 2)
        ek /= np.sum(ek)
        ek *= self.ic_energy
        self.spectrum = np.sum(ek[..., None] * wavemask, axis=tuple(range(self.stencil.d)))
        return ek, wavenumber

    def _generate_initial_velocity(self, ek, wavenumber):
        d = self.stencil.d
        axes = tuple(range(d))
        dx = self.units.convert_length_to_pu(1.0)
        np.random.seed(self.randseed)
        phase = np.random.random(np.array(wavenumber).shape) * 2 * np.pi + 0j
        uh = [np.fft.fftn(phase[c], axes=axes) for c in range(d)]
        re = [h.real.copy() for h in uh]
        im = [h.imag.copy() for h in uh]
        for c in range(d):
            re[c].ravel()[0] = 0
            im[c].ravel()[0] = 0
        # amplitudes from the target spectrum
        amp = [np.sqrt(2 / d * ek / (im[c] ** 2 + re[c] ** 2 + 1.e-15)) for c in range(d)]
        re_h = [amp[c] * re[c] for c in range(d)]
        im_h = [amp[c] * im[c] for c in range(d)]
        for c in range(d):
            re_h[c].ravel()[0] = 0
            im_h[c].ravel()[0] = 0
        # projection onto divergence-free modes with the modified wavenumber sin(k dx) / dx of
        # second-order central differences
        k_mod = [np.sin(wavenumber[c] * dx) / dx for c in range(d)]
        k_norm = np.linalg.norm(k_mod, axis=0) + 1e-16
        div_re = np.zeros(self.dimensions)
        div_im = np.zeros(self.dimensions)
        for c in range(d):
            div_re += k_mod[c] * re_h[c]
            div_im += k_mod[c] * im_h[c]
        re = [re_h[c] - div_re * k_mod[c] / k_norm ** 2 for c in range(d)]
        im = [im_h[c] - div_im * k_mod[c] / k_norm ** 2 for c in range(d)]
        for c in range(d):
            re[c].ravel()[0] = 0
            im[c].ravel()[0] = 0
        # scale to the requested initial energy
        e_kin = np.sum([np.sum(re[c] ** 2 + im[c] ** 2) for c in range(d)]) * .5
        factor = np.sqrt(self.ic_energy / e_kin)
        re = [r * factor for r in re]
        im = [i * factor for i in im]
        # back to physical space
        norm = ((self.resolution[0] * dx ** (1 - d) * np.sqrt(self.units.characteristic_length_pu))
                if d == 3 else (self.resolution[0] / dx))
        return np.asarray([(np.fft.ifftn(re[c] + im[c] * 1.0j, axes=axes) * norm).real for c in range(d)])

    def _compute_initial_pressure(self):
        return np.zeros(self.dimensions)[None, ...]

    def initial_pu(self):
        """(p, u) in PU; sets the characteristic velocity in PU (decayingturbulence.py:172-180)"""
        ek, wavenumber = self._generate_spectrum()
        u = self._generate_initial_velocity(ek, wavenumber)
        self.units.characteristic_velocity_pu = np.array(u).max()
        p = self._compute_initial_pressure()
        self.units.characteristic_velocity_pu = np.linalg.norm(u, axis=0).max()
        return p, u

    @property
    def energy_spectrum(self):
        return self.spectrum, self.wavenumbers

    @property
    def grid(self):
        axes = tuple(torch.linspace(0, 2 * torch.pi * (1 - 1 / n), steps=n,
                                    device=self.context.device, dtype=self.context.dtype)
                     for n in self.resolution)
        return torch.meshgrid(*axes, indexing='ij')

    @property
    def boundaries(self) -> List['Boundary']:
        return []


class Cavity2D(ExtFlow):
    """Lid-driven cavity: bounce-back walls left / right / bottom, moving equilibrium lid on top
    (liddrivencavity.py:14-71)."""

    def __init__(self, context, resolution, reynolds_number, mach_number):
        super().__init__(context, resolution, reynolds_number, mach_number)

    def make_resolution(self, resolution, stencil=None) -> List[int]:
        if isinstance(resolution, int):
            return [resolution] * 2
        assert len(resolution) == 2, 'expected 2-dimensional resolution'
        return resolution

    def make_units(self, reynolds_number, mach_number, resolution) -> 'UnitConversion':
        return UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number,
                              characteristic_length_lu=resolution[0], characteristic_length_pu=1,
                              characteristic_velocity_pu=1)

    def initial_pu(self):
        zeros = torch.zeros_like(self.grid[0])
        return torch.stack([zeros]), torch.stack([zeros] * 2)

    @property
    def grid(self):
        return _unit_grid(self, endpoint=False)

    @property
    def boundaries(self):
        x = self.grid[0]
        boundary = self.context.zero_tensor(x.shape, dtype=bool)
        top = self.context.zero_tensor(x.shape, dtype=bool)
        boundary[[0, -1], 1:] = True   # left and right
        boundary[:, 0] = True          # bottom
        top[:, -1] = True              # lid
        return [BounceBackBoundary(boundary),
                EquilibriumBoundaryPU(self.context, top, [float(self.units.characteristic_velocity_pu), 0.0])]


#: name -> (flow class, default stencil), as the reference's command line uses it (_flow_by_name.py:10-16)
flow_by_name = {
    'taylor2d': (TaylorGreenVortex, D2Q9),
    'taylor3d': (TaylorGreenVortex, D3Q19),
    'poiseuille2d': (PoiseuilleFlow2D, D2Q9),
    'shear2d': (DoublyPeriodicShear2D, D2Q9),
    'couette2d': (CouetteFlow2D, D2Q9),
    'decay2d': (DecayingTurbulence, D2Q9)}
