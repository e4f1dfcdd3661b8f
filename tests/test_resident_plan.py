"""Host logic of the tile-resident kernel for small 2-D lattices (lettuce_b200/csrc/lb_resident.cuh), no GPU needed:
* `lb_plan_resident` (pure host arithmetic behind the C ABI): the plan's invariants for many lattices / step counts;
* the scheme itself restated in numpy -- tile + halo of h nodes, h pushes with a shrinking valid region, tile written
  back, rounds given the parity of the step count -- equals n plain periodic steps BIT FOR BIT for any node-local
  collision (the GPU tests check the CUDA kernel against one launch per step)."""
import ctypes as C

import numpy as np
import pytest
import torch

import lettuce_b200 as lt

SMS, SMEM = 148, 232448   # B200: multiprocessors, opt-in shared memory per block


def plan(res, dtype, n_steps, sms=SMS, smem=SMEM):
    nat = lt.native
    desc = nat.make_desc("D2Q9", dtype, res)
    out = (C.c_int * 8)()
    nat.check(nat.lib().lb_plan_resident(desc, n_steps, sms, smem, out))
    return list(out)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("res", [[256, 256], [128, 128], [64, 64], [96, 160], [100, 36], [97, 64], [16, 16], [384, 320], [7, 5]])
def test_plan_invariants(res, dtype):
    elem = 4 if dtype == torch.float32 else 8
    for n in (2, 3, 4, 5, 7, 8, 9, 16, 37, 50, 100, 999, 1000):
        ok, ty, tz, rounds, base, extra, plane_max, smem = plan(res, dtype, n)
        if not ok:
            continue
        ny, nz = res                      # 2-D: kernel axes (Y, Z) = lattice axes (x, y)
        assert ny % ty == 0 and nz % tz == 0 and (ny // ty) * (nz // tz) <= SMS       # whole tiles, all co-resident
        assert rounds >= 1 and rounds % 2 == n % 2                                     # state ends where n steps leave it
        assert base * rounds + extra == n and 0 <= extra < rounds and base >= 1        # every step is taken once
        hmax = base + (1 if extra else 0)
        assert plane_max == (ty + 2 * hmax + 2) * (tz + 2 * hmax + 2)   # region + one ring of padding
        assert smem == 2 * 9 * plane_max * elem and smem <= SMEM                       # two region buffers fit
        assert hmax <= ny and hmax <= nz                                               # one periodic wrap per halo node


def test_plan_refuses_what_does_not_fit():
    assert plan([256, 256], torch.float64, 1)[0] == 0            # a single step: nothing to keep on chip
    assert plan([2048, 1024], torch.float32, 100)[0] == 0        # 2 M nodes: no tile per SM fits shared memory
    assert plan([256, 256], torch.float64, 100, sms=4)[0] == 0   # too few multiprocessors for tiles that fit
    nat = lt.native
    out = (C.c_int * 8)()
    nat.check(nat.lib().lb_plan_resident(nat.make_desc("D3Q19", torch.float32, [32, 32, 32]), 10, SMS, SMEM, out))
    assert list(out)[0] == 0                                     # 2-D stencil only


E = [(0, 0), (1, 0), (0, 1), (-1, 0), (0, -1), (1, 1), (-1, 1), (-1, -1), (1, -1)]


def _collide(g):   # any node-local map will do
    s = g.sum(0)
    return g * 0.5 + 0.5 * s / 9 + 0.01 * g ** 2


def _reference_steps(f, n):
    for _ in range(n):
        g = _collide(f)
        f = np.stack([np.roll(g[q], E[q], (0, 1)) for q in range(9)])
    return f


def _resident(f, ty, tz, rounds, base, extra):
    ny, nz = f.shape[1:]
    bufs = [f.copy(), np.full_like(f, np.nan)]
    for r in range(rounds):
        h = base + (1 if r < extra else 0)
        src, dst = bufs[r & 1], bufs[(r + 1) & 1]
        for t in range((ny // ty) * (nz // tz)):
            ty0, tz0 = (t // (nz // tz)) * ty, (t % (nz // tz)) * tz
            RY, RZ = ty + 2 * h, tz + 2 * h
            A = src[:, (ty0 - h + np.arange(RY)) % ny][:, :, (tz0 - h + np.arange(RZ)) % nz].copy()
            for k in range(1, h + 1):
                o = k - 1
                B = np.full_like(A, np.nan)
                g = _collide(A[:, o:RY - o, o:RZ - o])
                for q, (ey, ez) in enumerate(E):   # push, clipped to the buffer
                    ys, zs = np.arange(o, RY - o) + ey, np.arange(o, RZ - o) + ez
                    my, mz = (ys >= 0) & (ys < RY), (zs >= 0) & (zs < RZ)
                    B[q][np.ix_(ys[my], zs[mz])] = g[q][np.ix_(my, mz)]
                A = B
            dst[:, ty0:ty0 + ty, tz0:tz0 + tz] = A[:, h:h + ty, h:h + tz]
    return bufs[rounds & 1]


@pytest.mark.parametrize("res,n", [([24, 32], 7), ([16, 16], 5), ([12, 20], 4), ([24, 32], 2), ([16, 24], 9), ([30, 18], 13)])
def test_scheme_equals_plain_steps_bit_for_bit(res, n):
    ok, ty, tz, rounds, base, extra, _, _ = plan(res, torch.float64, n, sms=12, smem=60000)   # force several tiles
    assert ok
    f = np.random.default_rng(res[0] + n).random((9, *res))
    assert np.array_equal(_resident(f, ty, tz, rounds, base, extra), _reference_steps(f, n))
