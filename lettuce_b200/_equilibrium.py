"""Equilibria.  `QuadraticEquilibrium` mirrors reference/lettuce/ext/_equilibrium/
quadratic_equilibrium.py:10-31.  Inside the step kernel the same formula runs in registers
(csrc/lb_physics.cuh `equilibrium`); the torch expression here is the set-up-time helper that
`Flow.initialize` and `EquilibriumBoundaryPU.__call__` use.
`QuadraticEquilibriumLessMemory` (quadratic_equilibrium_less_memory.py:9-37) is the same function
evaluated with fewer temporaries; the kernel never materialises the equilibrium, so it maps to the
same native code.  `IncompressibleQuadraticEquilibrium` (incompressible_quadratic_equilibrium.py:
6-26) is a different function with no kernel: usable for set-up, refused by `Simulation`."""
import torch

from ._flow import Equilibrium

__all__ = ['QuadraticEquilibrium', 'QuadraticEquilibriumLessMemory', 'IncompressibleQuadraticEquilibrium']


class QuadraticEquilibrium(Equilibrium):
    def __call__(self, flow: 'Flow', rho=None, u=None):
        rho = flow.rho() if rho is None else rho
        u = flow.u() if u is None else u
        cs2 = flow.torch_stencil.cs ** 2
        exu = torch.tensordot(flow.torch_stencil.e, u, dims=1)
        uxu = flow.einsum("d,d->", [u, u])
        inner = rho * ((2 * exu - uxu) / (2 * cs2) + 0.5 * (exu / cs2) ** 2 + 1)
        return flow.einsum("q,q->q", [flow.torch_stencil.w, inner])

    def native_available(self) -> bool:
        return True

    def native_generator(self):
        return "quadratic"


class QuadraticEquilibriumLessMemory(QuadraticEquilibrium):
    def __call__(self, flow: 'Flow', rho=None, u=None):
        rho = flow.rho() if rho is None else rho
        u = flow.u() if u is None else u
        e, cs2 = flow.torch_stencil.e, flow.torch_stencil.cs ** 2
        return flow.einsum("q,q->q", [flow.torch_stencil.w,
                                      rho * ((2 * torch.tensordot(e, u, dims=1) - flow.einsum("d,d->", [u, u]))
                                             / (2 * cs2)
                                             + 0.5 * (torch.tensordot(e, u, dims=1) / cs2) ** 2 + 1)])


class IncompressibleQuadraticEquilibrium(Equilibrium):
    def __init__(self, rho0=1.0):
        self.rho0 = rho0

    def __call__(self, flow: 'Flow', rho=None, u=None):
        rho = flow.rho() if rho is None else rho
        u = flow.u() if u is None else u
        cs2 = flow.torch_stencil.cs ** 2
        exu = flow.einsum("qd,d->q", [flow.torch_stencil.e, u])
        uxu = flow.einsum("d,d->", [u, u])
        return flow.einsum("q,q->q", [flow.torch_stencil.w,
                                      rho + self.rho0 * ((2 * exu - uxu) / (2 * cs2) + 0.5 * (exu / cs2) ** 2)])

    def native_available(self) -> bool:
        return False
