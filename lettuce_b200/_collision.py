"""Collision operators with the reference's constructor signatures.  Each carries a POD
kernel descriptor (`native_descriptor`) instead of the reference's code generator, and
`__call__(flow)` runs the native collide-only kernel (lb_collide) so the operator can still be
used on its own, e.g. by conservation tests (reference/tests/collision/*).

BGK  reference/lettuce/ext/_collision/bgk_collision.py:12-35
TRT  .../trt_collision.py:6-27        KBC  .../kbc_collision.py:11-167
None .../no_collision.py:9-17
Regularized .../regularized_collision.py:8-49   Smagorinsky .../smagorinsky_collision.py:7-41
(the reference has no native kernels for these two)"""
import warnings
from abc import ABC, abstractmethod

import torch

from . import _native as nat

__all__ = ['Collision', 'BGKCollision', 'TRTCollision', 'KBCCollision', 'KBCCollision2D',
           'KBCCollision3D', 'NoCollision', 'RegularizedCollision', 'SmagorinskyCollision', 'MRTCollision']


class Collision(ABC):
    """reference/lettuce/_simulation.py:16-27"""

    @abstractmethod
    def native_descriptor(self, flow: 'Flow') -> dict:
        """{'collision': LB_COLL_*, 'tau': float, 'tau_minus': float}"""
        ...

    def native_available(self) -> bool:
        return True

    def native_generator(self):
        return self

    def __call__(self, flow: 'Flow') -> torch.Tensor:
        """collision(flow) -> post-collision populations (new tensor), on the GPU."""
        f = flow.f
        if not f.is_cuda:
            raise nat.NativeError("lettuce_b200 collisions run on CUDA tensors only (no CPU fallback)")
        d = self.native_descriptor(flow)
        desc = nat.make_desc(type(flow.stencil).__name__, f.dtype, flow.resolution, d['collision'],
                             d.get('tau', 1.0), d.get('tau_minus', 1.0), d.get('force', nat.LB_FORCE_NONE),
                             d.get('acceleration', ()), d.get('force_tau', 1.0))
        f = f.contiguous()
        out = torch.empty_like(f)
        with torch.cuda.device(f.device):
            nat.check(nat.lib().lb_collide(desc, f.data_ptr(), out.data_ptr(), None, None, 0,
                                           nat.current_stream_ptr(f.device)))
        return out


class NoCollision(Collision):
    def native_descriptor(self, flow):
        return dict(collision=nat.LB_COLL_NONE)

    def __call__(self, flow):
        return flow.f


class BGKCollision(Collision):
    def __init__(self, tau, force=None):
        if force is not None and not hasattr(force, 'native_descriptor'):
            raise NotImplementedError(f"{type(force).__name__} has no native kernel (Guo and ShanChen do)")
        self.tau = tau
        self.force = force

    def name(self):
        if self.force is not None:
            return f"{self.__class__.__name__}_{self.force.__class__.__name__}"
        return self.__class__.__name__

    def native_descriptor(self, flow):
        d = dict(collision=nat.LB_COLL_BGK, tau=float(self.tau))
        if self.force is not None:
            d.update(self.force.native_descriptor())
        return d


class TRTCollision(Collision):
    def __init__(self, tau, tau_minus=1.0):
        self.tau_plus = tau
        self.tau_minus = tau_minus

    def native_descriptor(self, flow):
        return dict(collision=nat.LB_COLL_TRT, tau=float(self.tau_plus), tau_minus=float(self.tau_minus))


class KBCCollision(Collision):
    """Entropic multi-relaxation model (Karlin et al.).  As in the reference, the constructor's
    tau is ignored: tau is read from flow.units.relaxation_parameter_lu at first use
    (kbc_collision.py:97-99)."""

    def __init__(self, tau: float = None):
        self.tau = tau
        self.beta = None

    def native_descriptor(self, flow):
        if self.beta is None:
            name = type(flow.stencil).__name__
            if flow.stencil.d == 3:
                assert name == "D3Q27", "KBC Collision is only implemented for D3Q27!"
            elif flow.stencil.d == 2:
                assert name == "D2Q9", "KBC Collision is only implemented for D2Q9!"
            else:
                raise NotImplementedError("KBC Collision is only implemented for 2d and 3d!")
            self.tau = flow.units.relaxation_parameter_lu
            self.beta = 1. / (2 * self.tau)
        return dict(collision=nat.LB_COLL_KBC, tau=float(self.tau))


class KBCCollision2D(KBCCollision):
    def __init__(self, tau: float = None):
        warnings.warn("KBCCollision2D is is deprecated! Use KBCCollision instead!")
        super().__init__()


class KBCCollision3D(KBCCollision):
    def __init__(self, tau: float = None):
        warnings.warn("KBCCollision2D is is deprecated! Use KBCCollision instead!")
        super().__init__()


class RegularizedCollision(Collision):
    """Regularized LBM according to Jonas Latt and Bastien Chopard (2006).  As in the reference, the
    constructor's tau is ignored: tau is read from flow.units.relaxation_parameter_lu at first use
    (regularized_collision.py:18-19)."""

    def __init__(self, tau: float = None):
        self.tau = tau
        self._bound = False

    def native_descriptor(self, flow):
        if not self._bound:
            self.tau = flow.units.relaxation_parameter_lu
            self._bound = True
        return dict(collision=nat.LB_COLL_REGULARIZED, tau=float(self.tau))


class SmagorinskyCollision(Collision):
    """Smagorinsky large eddy simulation (LES) collision model with BGK operator
    (smagorinsky_collision.py:7-36): two fixed-point iterations for the effective relaxation time per
    node, inside the step kernel.  `tau_eff` of the reference is a per-node field that exists only in
    registers here."""

    def __init__(self, tau, smagorinsky_constant=0.17, force=None):
        if force is not None:
            raise NotImplementedError("forcing schemes are outside the hot-path scope (SURVEY.md "
                                      "section 2 row 7); the reference's native path has none either")
        self.force = force
        self.tau = tau
        self.iterations = 2
        self.constant = smagorinsky_constant

    def native_descriptor(self, flow):
        # lb_desc.tau_minus carries the Smagorinsky constant (include/lettuce_b200.h)
        return dict(collision=nat.LB_COLL_SMAGORINSKY, tau=float(self.tau), tau_minus=float(self.constant))


class MRTCollision(Collision):
    """Multiple-relaxation-time operator over a user-supplied moment transform
    (reference/lettuce/ext/_collision/mrt_collision.py:6-36).  The reference has no native kernel for it
    (`native_available() == False`, :32-36) and neither has this engine: `Simulation` refuses it.  The
    operator itself stays callable on a flow (torch expressions on the flow's device) for analysis."""

    def __init__(self, transform: 'Transform', relaxation_parameters: list, context: 'Context'):
        self.transform = transform
        self.relaxation_parameters = context.convert_to_tensor(relaxation_parameters)

    def __call__(self, flow: 'Flow'):
        m = self.transform.transform(flow.f)
        meq = self.transform.equilibrium(m, flow)
        m = m - flow.einsum("q,q->q", [1 / self.relaxation_parameters, m - meq])
        return self.transform.inverse_transform(m)

    def native_available(self) -> bool:
        return False

    def native_descriptor(self, flow):
        raise NotImplementedError("MRTCollision has no native kernel")
