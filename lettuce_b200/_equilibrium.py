"""Equilibria.  `QuadraticEquilibrium` mirrors reference/lettuce/ext/_equilibrium/
quadratic_equilibrium.py:10-31.  Inside the step kernel the same formula runs in registers
(csrc/lb_physics.cuh `equilibrium`); the torch expression here is the set-up-time helper that
`Flow.initialize` and `EquilibriumBoundaryPU.__call__` use."""
import torch

from ._flow import Equilibrium

__all__ = ['QuadraticEquilibrium']


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
