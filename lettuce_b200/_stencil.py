"""Stencils.  Same classes, attribute names and index order as the reference
(reference/lettuce/_stencil.py:12-46, ext/_stencil/d1q3.py, d2q9.py:8-10, d3q15.py,
d3q19.py:8-13, d3q27.py:8-12) -- the order of `e` is the order of `f[q]` and is API."""
from typing import List

import numpy as np
import torch

__all__ = ['Stencil', 'TorchStencil', 'D1Q3', 'D2Q9', 'D3Q15', 'D3Q19', 'D3Q27']


class Stencil:
    e: List[List[int]]
    w: List[float]
    opposite: List[int]
    cs: float = 1 / np.sqrt(3.0)

    @property
    def d(self):
        return len(self.e[0])

    @property
    def q(self):
        return len(self.e)


class TorchStencil:
    cs: float = 1 / np.sqrt(3.0)

    def __init__(self, stencil: 'Stencil', context: 'Context'):
        self.e = context.convert_to_tensor(stencil.e)
        self.w = context.convert_to_tensor(stencil.w)
        self.opposite = context.convert_to_tensor(stencil.opposite)

    @property
    def d(self):
        return self.e.shape[1]

    @property
    def q(self):
        return self.e.shape[0]


def _pairs(q):
    return [0] + [i + 1 if i % 2 else i - 1 for i in range(1, q)]


_AXIS = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
_EDGE = [[0, 1, 1], [0, -1, -1], [0, 1, -1], [0, -1, 1], [1, 0, 1], [-1, 0, -1],
         [1, 0, -1], [-1, 0, 1], [1, 1, 0], [-1, -1, 0], [1, -1, 0], [-1, 1, 0]]
_CORNER = [[1, 1, 1], [-1, -1, -1], [1, 1, -1], [-1, -1, 1],
           [1, -1, 1], [-1, 1, -1], [1, -1, -1], [-1, 1, 1]]


class D1Q3(Stencil):
    def __init__(self):
        self.e = [[0], [1], [-1]]
        self.w = [2.0 / 3.0, 1.0 / 6.0, 1.0 / 6.0]
        self.opposite = [0, 2, 1]


class D2Q9(Stencil):
    def __init__(self):
        self.e = [[0, 0], [1, 0], [0, 1], [-1, 0], [0, -1], [1, 1], [-1, 1], [-1, -1], [1, -1]]
        self.w = [4.0 / 9.0] + [1.0 / 9.0] * 4 + [1.0 / 36.0] * 4
        self.opposite = [0, 3, 4, 1, 2, 7, 8, 5, 6]


class D3Q15(Stencil):
    def __init__(self):
        self.e = [[0, 0, 0]] + _AXIS + _CORNER
        self.w = [2.0 / 9.0] + [1.0 / 9.0] * 6 + [1.0 / 72.0] * 8
        self.opposite = _pairs(15)


class D3Q19(Stencil):
    def __init__(self):
        self.e = [[0, 0, 0]] + _AXIS + _EDGE
        self.w = [1.0 / 3.0] + [1.0 / 18.0] * 6 + [1.0 / 36.0] * 12
        self.opposite = _pairs(19)


class D3Q27(Stencil):
    def __init__(self):
        self.e = [[0, 0, 0]] + _AXIS + _EDGE + _CORNER
        self.w = [8.0 / 27.0] + [2.0 / 27.0] * 6 + [1.0 / 54.0] * 12 + [1.0 / 216.0] * 8
        self.opposite = _pairs(27)
