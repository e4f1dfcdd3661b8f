"""Flow state, moments and the Equilibrium / Boundary protocols.  Mirrors
reference/lettuce/_flow.py:16-236 (same attribute and method names).

`f` is the reference's `[q, *resolution]` C-contiguous tensor; the native engine borrows its
data pointer.  On a CUDA flow, rho() / j() / u() run the native moments kernel
(lb_moments); the torch expressions below them serve set-up on host tensors only."""
import pickle
from abc import ABC, abstractmethod
from typing import List, Optional, Union

import numpy as np
import torch

from . import _native as nat
from ._stencil import TorchStencil
from .util import torch_gradient, torch_jacobi

__all__ = ['Equilibrium', 'Flow', 'Boundary', 'initialize_f_neq', 'pressure_poisson',
           'initialize_pressure_poisson']


class Equilibrium(ABC):
    @abstractmethod
    def __call__(self, flow: 'Flow', rho=None, u=None) -> torch.Tensor:
        ...

    @abstractmethod
    def native_available(self) -> bool:
        ...

    def native_generator(self):
        """The reference returns a code generator here (_flow.py:25-27); the B200 engine has
        fixed kernels, so this returns the kernel descriptor id instead."""
        return None


class Boundary(ABC):
    @abstractmethod
    def __call__(self, flow: 'Flow'):
        ...

    @abstractmethod
    def make_no_collision_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        ...

    @abstractmethod
    def make_no_streaming_mask(self, shape: List[int], context: 'Context') -> Optional[torch.Tensor]:
        ...

    @abstractmethod
    def native_available(self) -> bool:
        ...

    def native_generator(self, index: int):
        return None


class Flow(ABC):
    """Physical configuration and state of the simulation (_flow.py:54-122)."""
    initialize_pressure: bool = False
    initialize_fneq: bool = False

    def __init__(self, context: 'Context', resolution: List[int], units: 'UnitConversion',
                 stencil: 'Stencil', equilibrium: 'Equilibrium'):
        self.context = context
        self.resolution = resolution
        self.units = units
        self.stencil = stencil
        self.torch_stencil = TorchStencil(stencil, context)
        self.equilibrium = equilibrium
        self.i = 0
        self.f = None          # allocated by initialize() (the reference allocates an empty tensor first, _flow.py:89-91)
        self._f_next = None
        self.initialize()

    @property
    @abstractmethod
    def boundaries(self) -> List['Boundary']:
        return []

    @abstractmethod
    def initial_pu(self):
        """initial solution in physical units: (p, u)"""
        ...

    def initialize(self):
        """equilibrium of the initial (p, u), optionally with first-order non-equilibrium
        (_flow.py:106-122).  On a CUDA context the populations are written by the engine's
        `lb_initialize` kernel straight from (rho, u): no [q, *res]-sized temporaries."""
        initial_p, initial_u = self.initial_pu()
        initial_rho = self.context.convert_to_tensor(
            self.units.convert_pressure_pu_to_density_lu(initial_p))
        initial_u = self.context.convert_to_tensor(self.units.convert_velocity_to_lu(initial_u))
        if self.initialize_pressure:
            initial_rho = pressure_poisson(self.units, initial_u, initial_rho)
        if self._native_initialize_ok(initial_rho, initial_u):
            self.f = self._native_initialize(initial_rho, initial_u, ghost=0)
            return
        self.f = self.equilibrium(self, rho=initial_rho, u=initial_u)
        if self.initialize_fneq:
            self.f = initialize_f_neq(self)

    def _native_initialize_ok(self, rho, u, ghost: int = 0) -> bool:
        from ._equilibrium import QuadraticEquilibrium
        ext = list(self.resolution[:-1]) + [self.resolution[-1] + 2 * ghost]
        return (rho.is_cuda and u.is_cuda and rho.dtype in nat.DTYPE_IDS and u.dtype == rho.dtype
                and type(self.stencil).__name__ in nat.STENCIL_IDS
                and isinstance(self.equilibrium, QuadraticEquilibrium)
                and list(rho.shape) == [1, *ext] and list(u.shape) == [self.stencil.d, *ext]
                and (not self.initialize_fneq or min(self.resolution) >= 3))

    def _native_initialize(self, rho, u, ghost: int) -> torch.Tensor:
        """lb_initialize: rho [1, *res_ext], u [d, *res_ext] (lattice units; `ghost` extra planes on
        both sides of the last axis) -> f [q, *res]"""
        desc = nat.make_desc(type(self.stencil).__name__, rho.dtype, self.resolution,
                             tau=float(self.units.relaxation_parameter_lu))
        rho, u = rho.contiguous(), u.contiguous()
        f = torch.empty([self.stencil.q, *self.resolution], dtype=rho.dtype, device=rho.device)
        with torch.cuda.device(rho.device):
            nat.check(nat.lib().lb_initialize(desc, rho.data_ptr(), u.data_ptr(), int(ghost),
                                              int(bool(self.initialize_fneq)), f.data_ptr(),
                                              nat.current_stream_ptr(rho.device)))
        return f

    # ---- second buffer (lazy, _flow.py:124-134) ------------------------------------------
    @property
    def f_next(self) -> torch.Tensor:
        if self._f_next is None or self._f_next.shape != self.f.shape \
                or self._f_next.dtype != self.f.dtype or self._f_next.device != self.f.device:
            self._f_next = torch.empty_like(self.f, memory_format=torch.contiguous_format)
        return self._f_next

    @f_next.setter
    def f_next(self, f_next_: torch.Tensor):
        self._f_next = f_next_

    # ---- moments ---------------------------------------------------------------------------
    def _native_moments_ok(self, f):
        return (f.is_cuda and f.is_contiguous() and f.dtype in nat.DTYPE_IDS
                and type(self.stencil).__name__ in nat.STENCIL_IDS
                and list(f.shape) == [self.stencil.q, *self.resolution])

    def _native_moments(self, f, want_rho, want_u):
        desc = nat.make_desc(type(self.stencil).__name__, f.dtype, self.resolution)
        rho = torch.empty([1, *self.resolution], dtype=f.dtype, device=f.device) if want_rho else None
        u = torch.empty([self.stencil.d, *self.resolution], dtype=f.dtype, device=f.device) if want_u else None
        with torch.cuda.device(f.device):
            nat.check(nat.lib().lb_moments(desc, f.data_ptr(), nat.ptr(rho), nat.ptr(u),
                                           nat.current_stream_ptr(f.device)))
        return rho, u

    def output_fields_pu(self, f: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[1 + d, nodes]: p_pu then u_pu, each FLAT IN FILE ORDER (lattice x fastest), what a field-writing
        reporter needs (vtk_reporter.py:29-47).  On a CUDA flow one `lb_output_fields` launch: the
        populations are read once and the fields leave the kernel already converted and reordered.
        `out`: a [1 + d, nodes] device tensor to write into (reporters reuse theirs)."""
        f = self.f if f is None else f
        d = self.stencil.d
        if self._native_moments_ok(f):
            units = self.units
            desc = nat.make_desc(type(self.stencil).__name__, f.dtype, self.resolution)
            if out is None:
                out = torch.empty([1 + d, f[0].numel()], dtype=f.dtype, device=f.device)
            assert out.is_contiguous() and tuple(out.shape) == (1 + d, f[0].numel()) and out.dtype == f.dtype
            p_scale = float(units.cs ** 2 / units.characteristic_pressure_lu * units.characteristic_pressure_pu)
            u_scale = float(units.characteristic_velocity_pu / units.characteristic_velocity_lu)
            with torch.cuda.device(f.device):
                nat.check(nat.lib().lb_output_fields(desc, f.data_ptr(), float(units.characteristic_density_lu),
                                                     p_scale, u_scale, out.data_ptr(),
                                                     nat.current_stream_ptr(f.device)))
            return out
        rho, u = self.rho_u(f)               # host tensors (set-up, tests): the reference's expressions
        fields = torch.cat([self.units.convert_density_lu_to_pressure_pu(rho), self.units.convert_velocity_to_pu(u)])
        return fields.permute(0, *range(d, 0, -1)).reshape(1 + d, -1)

    def rho_u(self, f: Optional[torch.Tensor] = None):
        """(rho, u) in lattice units from ONE pass over the populations (the output-side reporters
        need both; the reference computes them separately, _flow.py:136-172)"""
        f = self.f if f is None else f
        if self._native_moments_ok(f):
            return self._native_moments(f, True, True)
        rho = self.rho(f)
        return rho, self.u(f, rho=rho)

    def rho(self, f: Optional[torch.Tensor] = None) -> torch.Tensor:
        """density, shape [1, *res] (_flow.py:136-138)"""
        f = self.f if f is None else f
        if self._native_moments_ok(f):
            return self._native_moments(f, True, False)[0]
        return torch.sum(f, dim=0)[None, ...]

    def j(self, f: Optional[torch.Tensor] = None) -> torch.Tensor:
        """momentum (_flow.py:152-155)"""
        return self.einsum("qd,q->d", [self.torch_stencil.e, self.f if f is None else f])

    def u(self, f: Optional[torch.Tensor] = None, rho=None, acceleration=None) -> torch.Tensor:
        """velocity j / rho (_flow.py:157-172)"""
        ff = self.f if f is None else f
        if rho is None and acceleration is None and self._native_moments_ok(ff):
            return self._native_moments(ff, False, True)[1]
        rho = self.rho(f=f) if rho is None else rho
        v = self.j(f=f) / rho
        if acceleration is None:
            return v + 0.0
        if len(acceleration.shape) == 1:
            acceleration = acceleration[(Ellipsis,) + (None,) * self.stencil.d]
        return v + acceleration / (2 * rho)

    @property
    def velocity(self):
        return self.j() / self.rho()

    @property
    def rho_pu(self) -> torch.Tensor:
        return self.units.convert_density_to_pu(self.rho())

    @property
    def p_pu(self) -> torch.Tensor:
        return self.units.convert_density_lu_to_pressure_pu(self.rho())

    @property
    def u_pu(self):
        return self.units.convert_velocity_to_pu(self.u())

    def incompressible_energy(self, f: Optional[torch.Tensor] = None) -> torch.Tensor:
        """0.5 |u|^2 per node (_flow.py:178-181)"""
        u = self.u(f)
        return 0.5 * self.einsum("d,d->", [u, u])

    def entropy(self) -> torch.Tensor:
        f_log = -torch.log(self.einsum("q,q->q", [self.f, 1 / self.torch_stencil.w]))
        return self.einsum("q,q->", [self.f, f_log])

    def pseudo_entropy_global(self) -> torch.Tensor:
        f_w = self.einsum("q,q->q", [self.f, 1 / self.torch_stencil.w])
        return self.rho() - self.einsum("q,q->", [self.f, f_w])

    def pseudo_entropy_local(self, f: Optional[torch.Tensor] = None) -> torch.Tensor:
        f = self.f if f is None else f
        f_feq = f / self.equilibrium(self)
        return self.rho(f) - self.einsum("q,q->", [f, f_feq])

    def shear_tensor(self, f: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Pi_ab = sum_q f_q e_qa e_qb (_flow.py:198-205)"""
        e = self.torch_stencil.e
        return self.einsum("q,qab->ab", [self.f if f is None else f, self.einsum("qa,qb->qab", [e, e])])

    def einsum(self, equation, fields, *args) -> torch.Tensor:
        """Einstein summation on local fields: operands that carry the grid axes get an
        ellipsis appended (_flow.py:210-224)."""
        inputs, output = equation.split("->")
        inputs = inputs.split(",")
        for k, inp in enumerate(inputs):
            extra = len(fields[k].shape) - len(inp)
            if extra == 0:
                continue
            assert extra == self.stencil.d, "Bad dimension."
            inputs[k] += "..."
            if not output.endswith("..."):
                output += "..."
        return torch.einsum(",".join(inputs) + "->" + output, fields, *args)

    # ---- checkpoint (_flow.py:226-236) -----------------------------------------------------
    def dump(self, filename):
        with open(filename, "wb") as file:
            pickle.dump(self.context.convert_to_ndarray(self.f), file)

    def load(self, filename):
        with open(filename, "rb") as file:
            self.f = self.context.convert_to_tensor(pickle.load(file), dtype=self.context.dtype)
        self._f_next = None


def initialize_f_neq(flow: 'Flow'):
    """f = feq - fneq with fneq from the 6th-order velocity gradients (Krueger et al. 2017;
    _flow.py:309-336).  Set-up only."""
    rho = flow.rho()
    u = flow.u()
    S = torch.cat([torch_gradient(u[c], dx=1, order=6)[None, ...] for c in range(flow.stencil.d)])
    Pi_1 = 1.0 * flow.units.relaxation_parameter_lu * rho * S / flow.torch_stencil.cs ** 2
    e = flow.torch_stencil.e
    Q = (torch.einsum('ia,ib->iab', [e, e])
         - torch.eye(flow.stencil.d, device=e.device) * flow.stencil.cs ** 2)
    Pi_1_Q = flow.einsum('ab,iab->i', [Pi_1, Q])
    fneq = flow.einsum('i,i->i', [flow.torch_stencil.w, Pi_1_Q])
    return flow.equilibrium(flow, rho, u) - fneq


def pressure_poisson(units: 'UnitConversion', u, rho0, tol_abs=1e-10, max_num_steps=100000):
    """Density whose pressure solves lap(p) = -d_i d_j (u_i u_j), by Jacobi iteration from rho0
    (_flow.py:239-287).  Like the reference it always iterates over the first TWO axes
    (`dim=2` is hard-coded at :280; the 3-D case is marked as not working there).  Set-up only."""
    dx = units.convert_length_to_pu(1.0)
    u = units.convert_velocity_to_pu(u)
    p = units.convert_density_lu_to_pressure_pu(rho0)
    with torch.no_grad():
        u_mod = torch.zeros_like(u[0])
        dim = u.shape[0]
        for i in range(dim):
            for j in range(dim):
                u_mod -= torch_gradient(torch_gradient(u[i] * u[j], dx)[i], dx)[j]
    p_mod = torch_jacobi(u_mod, p[0], dx, dim=2, tol_abs=tol_abs, max_num_steps=max_num_steps)[None, ...]
    return units.convert_pressure_pu_to_density_lu(p_mod)


def initialize_pressure_poisson(flow: 'Flow', max_num_steps=100000, tol_pressure=1e-6):
    """equilibrium populations of (rho from pressure_poisson, current u) (_flow.py:290-305)"""
    u = flow.u()
    rho = pressure_poisson(flow.units, u, flow.rho(), tol_abs=tol_pressure, max_num_steps=max_num_steps)
    return flow.equilibrium(flow=flow, rho=rho, u=u)
