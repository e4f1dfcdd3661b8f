"""Small host-side helpers with the reference's names (reference/lettuce/util/utility.py).
Used at set-up time only (initial conditions); nothing here is on the per-step path."""
import inspect

import torch

__all__ = ['get_subclasses', 'LettuceException', 'LettuceWarning', 'InefficientCodeWarning',
           'ExperimentalWarning', 'torch_gradient', 'grid_fine_to_coarse', 'torch_jacobi', 'append_axes']


def get_subclasses(cls, module):
    """direct subclasses of `cls` among the members of `module` (utility.py:15-18)"""
    for _, obj in inspect.getmembers(module):
        if hasattr(obj, "__bases__") and cls in obj.__bases__:
            yield obj


class LettuceException(Exception):
    pass


class LettuceWarning(UserWarning):
    pass


class InefficientCodeWarning(LettuceWarning):
    pass


class ExperimentalWarning(LettuceWarning):
    pass


# central-difference coefficients c_s on f[x + s] for s = 1..3 (antisymmetric), Fornberg 1988;
# utility.py:52-59 lists them as weights on roll(+3..-3), i.e. on f[x - s]
_COEFF = {2: (1 / 2,), 4: (2 / 3, -1 / 12), 6: (3 / 4, -3 / 20, 1 / 60)}


def torch_gradient(f: torch.Tensor, dx=1, order=2) -> torch.Tensor:
    """Periodic first derivative along every axis of a 2-D / 3-D field; returns [dim, *f.shape]
    (utility.py:37-99)."""
    if f.ndim not in (2, 3):
        raise LettuceException("Invalid dimension!")
    coeff = _COEFF[order]
    with torch.no_grad():
        out = torch.empty([f.ndim, *f.shape], dtype=f.dtype, device=f.device)
        inv_dx = torch.tensor(1.0 / dx, dtype=f.dtype, device=f.device)
        for axis in range(f.ndim):
            # same term order as the reference: shifts +S..+1 then -1..-S
            acc = None
            for s in range(len(coeff), 0, -1):
                term = -coeff[s - 1] * f.roll(shifts=s, dims=axis)
                acc = term if acc is None else acc + term
            for s in range(1, len(coeff) + 1):
                acc = acc + coeff[s - 1] * f.roll(shifts=-s, dims=axis)
            out[axis] = acc * inv_dx
    return out


def append_axes(array, n):
    """array[..., None * n] (utility.py append_axes)"""
    return array[(Ellipsis,) + (None,) * n]


def grid_fine_to_coarse(flow: 'Flow', f_fine, tau_fine, tau_coarse):
    """Every second node of a fine-grid field with the non-equilibrium part rescaled by
    2 tau_coarse / tau_fine (utility.py:102-118).  Set-up helper (restart on a coarser grid)."""
    if f_fine.dim() not in (3, 4):
        raise LettuceException("Invalid dimension!")
    sub = f_fine[(slice(None),) + (slice(None, None, 2),) * (f_fine.dim() - 1)]
    f_eq = flow.equilibrium(flow, rho=flow.rho(sub), u=flow.u(sub))
    return f_eq + 2 * tau_coarse / tau_fine * (sub - f_eq)


def torch_jacobi(f, p, dx, dim, tol_abs=1e-10, max_num_steps=100000):
    """Jacobi iteration for the periodic Poisson equation lap(p) = f with second-order
    differences; stops when mean(residuum^2) <= tol_abs (utility.py:121-166).  The residuum is
    taken with the updated p, as the reference does."""
    axes = range(dim)
    error, it = 1, 0
    while error > tol_abs and it < max_num_steps:
        it += 1
        nb = None
        for s in (1, -1):
            for a in axes:
                r = p.roll(shifts=s, dims=a)
                nb = r if nb is None else nb + r
        p = (f * (dx ** 2) - nb) * -1 / (2 * dim)
        nb = None
        for s in (1, -1):
            for a in axes:
                r = p.roll(shifts=s, dims=a)
                nb = r if nb is None else nb + r
        residuum = f - (nb - 2 * dim * p) / (dx ** 2)
        error = torch.mean(residuum ** 2)
    return p
