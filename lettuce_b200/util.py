"""Small host-side helpers with the reference's names (reference/lettuce/util/utility.py).
Used at set-up time only (initial conditions); nothing here is on the per-step path."""
import torch

__all__ = ['LettuceException', 'LettuceWarning', 'InefficientCodeWarning', 'ExperimentalWarning',
           'torch_gradient', 'append_axes']


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
