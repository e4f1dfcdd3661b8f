"""CPU tests of the host-side mirror of the lettuce API: stencils, units, flow initialisation
(against the reference's own initial fields stored in the golden files), mask construction."""
import numpy as np
import pytest
import torch

import lettuce_b200 as lt
from golden_util import DT, load_case


@pytest.mark.parametrize("S", [lt.D1Q3, lt.D2Q9, lt.D3Q15, lt.D3Q19, lt.D3Q27])
def test_stencil_invariants(S):
    """reference/tests/stencil/test_{first_zero,opposite,symmetry,weights}.py"""
    s = S()
    e, w = np.array(s.e), np.array(s.w)
    assert (e[0] == 0).all()
    assert (e[s.opposite] == -e).all()
    assert (e.sum(axis=0) == 0).all()
    assert w.sum() == pytest.approx(1.0)
    assert s.cs == 1 / np.sqrt(3.0)


def test_stencil_tables_match_oracle_restated_reference():
    from oracle.lbm_oracle import STENCILS
    for name, tab in STENCILS.items():
        s = getattr(lt, name)()
        assert s.e == tab["e"] and s.opposite == tab["opposite"] and s.w == tab["w"]


def test_context_policy():
    """reference/lettuce/_context.py:17-62"""
    c = lt.Context(device="cpu")
    assert c.device.type == "cpu" and c.use_native is False and c.dtype == torch.float32
    with pytest.raises(AssertionError):
        lt.Context(device="cpu", use_native=True)
    if not torch.cuda.is_available():
        assert lt.Context().device.type == "cpu"
        with pytest.raises(AssertionError):
            lt.Context(device="cuda")
    t = c.convert_to_tensor(np.zeros(3, dtype=bool))
    assert t.dtype == torch.bool
    assert c.convert_to_tensor(np.zeros(3, dtype=np.uint8)).dtype == torch.uint8
    assert c.convert_to_tensor([1, 2]).dtype == torch.float32


def test_units():
    """reference/lettuce/_unit.py:35-60 and the [probe] in SURVEY appendix A.13"""
    u = lt.UnitConversion(reynolds_number=100, mach_number=0.05, characteristic_length_lu=256,
                          characteristic_length_pu=2 * np.pi, characteristic_velocity_pu=1)
    cs = 1 / np.sqrt(3.0)
    assert u.characteristic_velocity_lu == cs * 0.05
    assert u.relaxation_parameter_lu == 256 * cs * 0.05 / 100 / cs ** 2 + 0.5
    assert u.convert_velocity_to_pu(u.convert_velocity_to_lu(0.3)) == pytest.approx(0.3)
    assert u.convert_pressure_pu_to_density_lu(0.0) == 1.0
    assert u.convert_density_lu_to_pressure_pu(u.convert_pressure_pu_to_density_lu(0.02)) == pytest.approx(0.02)
    assert u.convert_time_to_pu(u.convert_time_to_lu(2.0)) == pytest.approx(2.0)
    assert u.convert_length_to_pu(1.0) == 2 * np.pi / 256


@pytest.mark.parametrize("name", ["tgv_d2q9_bgk_f64", "tgv_d3q19_bgk_f64", "tgv_d3q27_bgk_f32",
                                  "tgv_d2q9_bgk_f32", "tgv_d3q19_bgk_f32"])
def test_tgv_initialisation_matches_reference(name):
    """our TaylorGreenVortex (analytic p,u -> feq -> f_neq correction) reproduces the initial
    populations the reference built (golden f0 of the tgv_* cases)."""
    case = load_case(name)
    meta = case["meta"]
    ctx = lt.Context(device="cpu", dtype=DT[meta["dtype"]], use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=meta["res"], reynolds_number=100.0, mach_number=0.05,
                                stencil=getattr(lt, meta["stencil"]))
    assert flow.units.relaxation_parameter_lu == meta["tau"]
    tol = 1e-6 if meta["dtype"] == "f32" else 1e-13
    assert np.abs(flow.f.numpy() - case["f0"]).max() <= tol * np.abs(case["f0"]).max()


def test_mask_construction_later_index_wins():
    """reference/lettuce/_simulation.py:57-86"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.Obstacle(ctx, [12, 8], reynolds_number=10, mach_number=0.05, domain_length_x=1)
    m = torch.zeros([12, 8], dtype=torch.bool)
    m[0:3, 2:5] = True                      # overlaps the inlet column x = 0
    flow.mask = m
    bnds = [None] + sorted(flow.boundaries, key=lambda b: str(b))
    names = [type(b).__name__ for b in bnds[1:]]
    # the stock list of obstacle.py:107-122 in the reference's sorted order
    assert names == ["AntiBounceBackOutlet", "BounceBackBoundary", "EquilibriumBoundaryPU"]
    ncm, nsm = lt.build_masks(flow, bnds, ctx)
    assert ncm.dtype == torch.uint8 and nsm.dtype == torch.uint8
    assert list(nsm.shape) == [9, 12, 8]
    assert (ncm[1:3, 2:5] == 2).all()                 # bounce-back
    assert (ncm[0, :] == 3).all()                     # inlet overwrote the overlap (later index wins)
    assert (ncm[-1, :] == 1).all()                    # the outlet plane
    # anti_bounce_back_outlet.py:88-93: the populations opposite to the outgoing ones stay on the plane
    kept = [q for q in range(9) if nsm[q].any()]
    assert kept == [3, 6, 7] and all(bool(nsm[q, -1, :].all()) and not bool(nsm[q, :-1, :].any()) for q in kept)
    assert lt.build_masks(flow, [None], ctx) == (None, None)


def test_broadcast_strides():
    from lettuce_b200._native import NativeError, broadcast_strides
    u = torch.zeros([2, 16, 1])
    assert broadcast_strides(u, 1, [16, 8]) == [1, 0, 0]
    assert broadcast_strides(torch.zeros([2]), 1, [16, 8]) == [0, 0, 0]
    assert broadcast_strides(torch.zeros([3, 4, 5, 6]), 1, [4, 5, 6]) == [30, 6, 1]
    assert broadcast_strides(torch.zeros([16, 8]), 0, [16, 8]) == [8, 1, 0]
    with pytest.raises(NativeError):
        broadcast_strides(torch.zeros([2, 3, 8]), 1, [16, 8])


def test_torch_gradient_order6():
    """reference/tests/util/test_torch_gradient.py: analytic derivative of sin on a periodic grid"""
    n = 64
    x = torch.linspace(0, 2 * np.pi * (1 - 1 / n), n, dtype=torch.float64)
    X, Y = torch.meshgrid(x, x, indexing="ij")
    g = lt.torch_gradient(torch.sin(X) * torch.cos(2 * Y), dx=2 * np.pi / n, order=6)
    assert torch.allclose(g[0], torch.cos(X) * torch.cos(2 * Y), atol=1e-6)
    assert torch.allclose(g[1], -2 * torch.sin(X) * torch.sin(2 * Y), atol=1e-5)
