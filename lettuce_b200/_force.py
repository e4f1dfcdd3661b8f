"""Forcing schemes for BGKCollision(tau, force): Guo (reference/lettuce/ext/_force/guo.py:9-43) and ShanChen
(shan_chen.py:7-29), SURVEY section 8f rank 3.  Same constructors as the reference; the arithmetic runs
inside the step kernel (lb_desc.force / .acceleration / .force_tau), these objects only carry the
parameters.  The acceleration is a constant vector in lattice units."""
from abc import ABC, abstractmethod

from . import _native as nat

__all__ = ['Force', 'Guo', 'ShanChen']


class Force(ABC):
    """reference/lettuce/ext/_force/_force.py:6-33"""

    def __init__(self, flow: 'Flow', tau, acceleration):
        self.flow = flow
        self.tau = tau
        self.acceleration = flow.context.convert_to_tensor(acceleration)
        if self.acceleration.dim() != 1 or self.acceleration.numel() != flow.stencil.d:
            raise NotImplementedError("the native forcing schemes take one constant acceleration vector "
                                      f"of {flow.stencil.d} components")

    @property
    @abstractmethod
    def ueq_scaling_factor(self):
        ...

    @property
    @abstractmethod
    def native_kind(self) -> int:
        ...

    def native_available(self) -> bool:
        return True

    def native_descriptor(self) -> dict:
        return dict(force=self.native_kind, force_tau=float(self.tau),
                    acceleration=[float(a) for a in self.acceleration.tolist()])

    def u_eq(self, flow: 'Flow' = None):
        """ueq_scaling_factor * acceleration / rho (guo.py:31-33, shan_chen.py:16-19)"""
        flow = self.flow if flow is None else flow
        acc = self.acceleration.to(flow.f.device)[(Ellipsis,) + (None,) * flow.stencil.d]
        return self.ueq_scaling_factor * acc / flow.rho()


class Guo(Force):
    @property
    def ueq_scaling_factor(self):
        return 0.5

    @property
    def native_kind(self):
        return nat.LB_FORCE_GUO


class ShanChen(Force):
    @property
    def ueq_scaling_factor(self):
        return self.tau * 1

    @property
    def native_kind(self):
        return nat.LB_FORCE_SHAN_CHEN
