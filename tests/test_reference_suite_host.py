"""The reference's host-side tests restated against lettuce_b200 (no GPU): tests/unit/*, tests/stencil/*,
tests/util/test_grid_fine_to_coarse.py, tests/util/test_torch_gradient.py -- same parameters and tolerances."""
import numpy as np
import pytest
import torch

import lettuce_b200 as lt

STENCILS = [lt.D1Q3, lt.D2Q9, lt.D3Q15, lt.D3Q19, lt.D3Q27]


def _units():      # tests/unit/__init__.py
    return lt.UnitConversion(reynolds_number=1000, mach_number=0.05, characteristic_length_lu=100,
                             characteristic_length_pu=2 * np.pi, characteristic_velocity_pu=2,
                             characteristic_density_pu=0.7)


def test_consistency():
    """tests/unit/test_consistency.py"""
    rho, u = 0.9, 0.1
    units = _units()
    assert (units.convert_density_to_pu(rho) * units.convert_velocity_to_pu(u) ** 2
            == pytest.approx(units.convert_energy_to_pu(rho * u * u)))
    assert units.convert_velocity_to_pu(u) ** 2 == pytest.approx(units.convert_incompressible_energy_to_pu(u * u))


def test_conversion_reversible():
    """tests/unit/test_conversion_reversible.py"""
    two = pytest.approx(2.0)
    units = _units()
    assert two == units.convert_velocity_to_lu(units.convert_velocity_to_pu(2.0))
    assert two == units.convert_time_to_lu(units.convert_time_to_pu(2.0))
    assert two == units.convert_length_to_lu(units.convert_length_to_pu(2.0))
    assert two == units.convert_density_to_lu(units.convert_density_to_pu(2.0))
    assert two == units.convert_pressure_to_lu(units.convert_pressure_to_pu(2.0))
    assert two == units.convert_density_lu_to_pressure_pu(units.convert_pressure_pu_to_density_lu(2.0))
    assert two == units.convert_energy_to_lu(units.convert_energy_to_pu(2.0))
    assert two == units.convert_incompressible_energy_to_lu(units.convert_incompressible_energy_to_pu(2.0))


def test_reynolds_number_consistent():
    """tests/unit/test_reynolds_number_consistent.py"""
    units = _units()
    re_lu = units.characteristic_velocity_lu * units.characteristic_length_lu / units.viscosity_lu
    re_pu = units.characteristic_velocity_pu * units.characteristic_length_pu / units.viscosity_pu
    assert re_lu == pytest.approx(re_pu)


@pytest.mark.parametrize("S", STENCILS)
def test_stencil_first_zero_opposite_symmetry_weights(S):
    """tests/stencil/test_first_zero.py, test_opposite.py, test_symmetry.py, test_weights.py"""
    s = S()
    assert s.e[0] == pytest.approx(np.zeros_like(s.e[0]))
    ts = lt.TorchStencil(s, lt.Context(device="cpu", dtype=torch.float64, use_native=False))
    assert torch.isclose(ts.e[s.opposite], -ts.e).all()
    assert np.array(s.e).sum() == pytest.approx(0.0)
    assert np.array(s.w).sum() == pytest.approx(1.0)


@pytest.mark.parametrize("dims", [2, 3])
def test_grid_fine_to_coarse(dims):
    """tests/util/test_grid_fine_to_coarse.py"""
    context = lt.Context(device='cpu', dtype=torch.float64, use_native=False)
    flow_f = lt.TaylorGreenVortex(context, [40] * dims, 1600, 0.15)
    flow_c = lt.TaylorGreenVortex(context, [20] * dims, 1600, 0.15)
    f_c = lt.grid_fine_to_coarse(flow_f, flow_f.f, flow_f.units.relaxation_parameter_lu,
                                 flow_c.units.relaxation_parameter_lu)
    p_init, u_init = flow_c.initial_pu()
    rho_init = flow_c.units.convert_pressure_pu_to_density_lu(p_init)
    u_init = flow_c.units.convert_velocity_to_lu(u_init)
    assert torch.isclose(flow_c.u(f_c), u_init).all()
    assert torch.isclose(flow_c.rho(f_c), rho_init).all()
    assert torch.isclose(f_c, flow_c.f).all()
    assert torch.isclose(flow_c.shear_tensor(flow_c.f), flow_c.shear_tensor(f_c)).all()


@pytest.mark.parametrize("dims", [2, 3])
@pytest.mark.parametrize("order", [2, 4, 6])
def test_torch_gradient(dims, order):
    """tests/util/test_torch_gradient.py (resolution 100 in 2-D, 48 in 3-D to keep the CPU suite short)"""
    context = lt.Context(device='cpu', dtype=torch.float64, use_native=False)
    n = 100 if dims == 2 else 48
    flow = lt.TaylorGreenVortex(context=context, resolution=[n] * dims, reynolds_number=1, mach_number=0.05)
    _, u = flow.initial_pu()
    dx = flow.units.convert_length_to_pu(1.0)
    u0_grad = lt.torch_gradient(u[0], dx=dx, order=order).numpy()
    u = u.numpy()
    u0_grad_np = np.array(np.gradient(u[0], dx))
    g = [x.numpy() for x in flow.grid]
    if dims == 2:
        x, y = g
        analytic = np.array([-np.sin(x) * np.sin(y), np.cos(x) * np.cos(y)])
        inner = (slice(None), slice(2, -2), slice(2, -2))
    else:
        x, y, z = g
        analytic = np.array([np.cos(x) * np.cos(y) * np.cos(z), np.sin(x) * np.sin(y) * (-np.cos(z)),
                             np.sin(x) * (-np.cos(y)) * np.sin(z)])
        inner = (slice(None), slice(2, -2), slice(2, -2), slice(2, -2))
    atol = 1e-3 if dims == 2 else 5e-3          # second order on the coarser 3-D grid: (2 pi / 48)^2 / 6 = 2.9e-3
    assert np.allclose(analytic, u0_grad, rtol=0.0, atol=atol)
    assert u0_grad_np[inner] == pytest.approx(u0_grad[inner], rel=0.0, abs=atol)
