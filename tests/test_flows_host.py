"""Host-side checks (no GPU) of the stock flows and the output side of the path
against golden files produced by the unmodified reference (tests/golden/make_golden_flows.py):
initial populations, masks, boundary values, energy spectrum, the fields VTKReporter writes, the
ErrorReporter row, and the reference's own tests of these pieces restated
(tests/flow/test_divergence.py, test_pressure_poisson.py, test_initialize_pressure.py,
tests/reporter/test_write_vtk.py, test_vtk_reporter_*.py, test_energy_spectrum.py)."""
import os
import types

import numpy as np
import pytest
import torch

import lettuce_b200 as lt
from golden_util import DT, golden_names, load_case

FLOW_CASES = [n for n in golden_names() if n.startswith("flow_")]


def build_flow(meta, device="cpu"):
    args = dict(meta["flow"]["args"])
    if "stencil" in args:
        args["stencil"] = getattr(lt, args["stencil"])
    ctx = lt.Context(device=device, dtype=DT[meta["dtype"]], use_native=(device != "cpu"))
    return getattr(lt, meta["flow"]["cls"])(ctx, **args)


def _tol(meta):
    return 2e-6 if meta["dtype"] == "f32" else 1e-12


def test_flow_cases_present():
    assert len(FLOW_CASES) == 12


@pytest.mark.parametrize("name", FLOW_CASES)
def test_flow_initialisation_matches_reference(name):
    """same constructor arguments -> same units, same initial populations, same p / u in PU"""
    case = load_case(name)
    meta = case["meta"]
    flow = build_flow(meta)
    un = meta["units"]
    assert float(flow.units.relaxation_parameter_lu) == pytest.approx(meta["tau"], rel=1e-14)
    assert float(flow.units.characteristic_length_lu) == pytest.approx(un["length_lu"], rel=1e-14)
    assert float(flow.units.characteristic_velocity_pu) == pytest.approx(un["velocity_pu"], rel=1e-12)
    assert list(flow.resolution) == meta["res"] and type(flow.stencil).__name__ == meta["stencil"]
    f = flow.f.numpy()
    assert f.dtype == case["f0"].dtype and f.shape == case["f0"].shape
    assert np.abs(f - case["f0"]).max() <= _tol(meta) * np.abs(case["f0"]).max()
    scale = max(np.abs(case["u_pu0"]).max(), 1e-3)
    assert np.abs(flow.u_pu.numpy() - case["u_pu0"]).max() <= 50 * _tol(meta) * scale
    pscale = max(np.abs(case["p_pu0"]).max(), 1e-3)
    assert np.abs(flow.p_pu.numpy() - case["p_pu0"]).max() <= 2e3 * _tol(meta) * pscale


@pytest.mark.parametrize("name", [n for n in FLOW_CASES if "cavity" in n or "poiseuille" in n])
def test_flow_boundaries_match_reference(name):
    """the flow's boundary list, sorted as Simulation sorts it, gives the reference's masks and
    lattice-unit boundary values"""
    case = load_case(name)
    meta = case["meta"]
    flow = build_flow(meta)
    bnds = [None] + sorted(flow.boundaries, key=lambda b: str(b))
    kinds = ["equilibrium" if isinstance(b, lt.EquilibriumBoundaryPU) else "bounce_back" for b in bnds[1:]]
    assert kinds == [b["kind"] for b in meta["boundaries"]]
    ncm, nsm = lt.build_masks(flow, bnds, flow.context)
    assert np.array_equal(ncm.numpy(), case["ncm"]) and np.array_equal(nsm.numpy(), case["nsm"])
    for i, b in enumerate(bnds[1:], start=1):
        if isinstance(b, lt.EquilibriumBoundaryPU):
            rho, u = b._lattice_units(flow)
            assert np.allclose(rho.numpy(), case[f"b{i}_rho"], rtol=1e-6, atol=0)
            assert np.allclose(u.numpy(), case[f"b{i}_u"], rtol=1e-6, atol=1e-12)
    if "poiseuille" in name:
        assert [float(a) for a in flow.acceleration] == pytest.approx(meta["force"]["acceleration"])


def test_couette_matches_reference_structure():
    """CouetteFlow2D: u_char_pu = 0 makes the reference's populations NaN (couette.py:21,42); the
    class is kept with the same boundary layout (moving equilibrium row y = 1, bounce-back row y = -1)."""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.CouetteFlow2D(ctx, 16, reynolds_number=10, mach_number=0.05)
    assert flow.resolution == [16, 16] and flow.units.characteristic_velocity_pu == 0
    eq, bb = flow.boundaries
    assert isinstance(eq, lt.EquilibriumBoundaryPU) and isinstance(bb, lt.BounceBackBoundary)
    assert bool(eq.make_no_collision_mask([16, 16], ctx)[:, 1].all())
    assert bool(bb.make_no_collision_mask([16, 16], ctx)[:, -1].all())
    assert flow.analytic_solution().shape == (1, 16, 16)


def test_flow_by_name_table():
    assert set(lt.flow_by_name) == {'taylor2d', 'taylor3d', 'poiseuille2d', 'shear2d', 'couette2d', 'decay2d'}
    assert lt.flow_by_name['taylor3d'] == (lt.TaylorGreenVortex, lt.D3Q19)
    assert lt.flow_by_name['decay2d'] == (lt.DecayingTurbulence, lt.D2Q9)
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    for key, (cls, stencil) in lt.flow_by_name.items():      # constructed as the reference's CLI does
        flow = cls(ctx, resolution=[16] * stencil().d, reynolds_number=100, mach_number=0.05, stencil=stencil)
        assert list(flow.f.shape) == [stencil().q] + [16] * stencil().d


# ---- output side --------------------------------------------------------------------------------
@pytest.mark.parametrize("name", [n for n in FLOW_CASES if "spectrum0" in load_case(n)])
def test_energy_spectrum_matches_reference(name):
    case = load_case(name)
    meta = case["meta"]
    flow = build_flow(meta)
    spec = lt.EnergySpectrum(flow)
    tol = 2e-4 if meta["dtype"] == "f32" else 1e-10
    for fkey, skey in (("f0", "spectrum0"), ("f_out", "spectrum_out")):
        flow.f = torch.from_numpy(case[fkey])
        got = spec().numpy()
        assert got.shape == case[skey].shape
        assert np.abs(got - case[skey]).max() <= tol * np.abs(case[skey]).max()


@pytest.mark.parametrize("key", ["taylor2d", "taylor3d", "poiseuille2d", "shear2d", "decay2d"])
def test_energy_spectrum_reference_test(key):
    """tests/reporter/test_energy_spectrum.py:5-28"""
    cls, stencil = lt.flow_by_name[key]
    ctx = lt.Context(device="cpu", dtype=torch.float32, use_native=False)
    flow = cls(context=ctx, resolution=[20] * stencil().d, reynolds_number=1600, mach_number=0.01, stencil=stencil)
    spectrum = lt.EnergySpectrum(flow)().numpy()
    obs = lt.IncompressibleKineticEnergy(flow)
    energy = obs.from_lattice(float(flow.incompressible_energy().sum()))
    if cls is lt.DecayingTurbulence:
        ek_ref, _ = flow.energy_spectrum
        assert spectrum == pytest.approx(ek_ref, rel=0.0, abs=0.1)
    if cls is lt.TaylorGreenVortex:
        ek_max = sorted(spectrum, reverse=True)
        assert ek_max[0] * 1e-5 > ek_max[1]
    assert energy == pytest.approx(np.sum(spectrum), rel=0.1, abs=0.0)


def _fake_simulation(flow, ncm=None):
    return types.SimpleNamespace(flow=flow, units=flow.units, no_collision_mask=ncm)


@pytest.mark.parametrize("name", ["flow_shear2d_d2q9_bgk_f32", "flow_decay3d_d3q19_bgk_f64",
                                  "flow_cavity2d_d2q9_bgk_f64"])
def test_vtk_reporter_writes_reference_fields(name, tmp_path):
    """the arrays in the .vtr file are flow.p_pu / flow.u_pu of the reference, in the layout of
    vtk_reporter.py:29-47 ([nx, ny, 1] in 2-D), readable back bit for bit"""
    case = load_case(name)
    meta = case["meta"]
    flow = build_flow(meta)
    flow.f = torch.from_numpy(case["f0"])
    rep = lt.VTKReporter(interval=1, filename_base=str(tmp_path / "out" / "fields"))
    assert os.path.isdir(tmp_path / "out")
    rep(_fake_simulation(flow))
    fn = tmp_path / "out" / "fields_00000000.vtr"
    assert os.path.isfile(fn)
    x, y, z, arrays = lt.read_vtr(fn)
    d = len(meta["res"])
    shape3 = tuple(meta["res"]) + (1,) * (3 - d)
    assert (x.size, y.size, z.size) == shape3 and x.dtype == np.int64 and np.array_equal(x, np.arange(shape3[0]))
    assert list(arrays) == ["p"] + ["ux", "uy", "uz"][:d]
    tol = 50 * _tol(meta)
    ref = {"p": case["p_pu0"][0]}
    ref.update({f"u{'xyz'[c]}": case["u_pu0"][c] for c in range(d)})
    for k, a in arrays.items():
        assert a.shape == shape3 and a.dtype == case["f0"].dtype
        assert np.array_equal(a, rep.point_dict[k])
        scale = max(np.abs(ref[k]).max(), 1e-3) * (40 if k == "p" else 1)
        assert np.abs(a.reshape(meta["res"]) - ref[k]).max() <= tol * scale
    flow.i = 3
    rep.interval = 2
    rep(_fake_simulation(flow))                                   # 3 % 2 != 0: nothing written
    assert not os.path.exists(tmp_path / "out" / "fields_00000003.vtr")


def test_vtr_container_layout(tmp_path):
    """the container: XML header, appended raw blocks with UInt64 sizes, x-fastest point order"""
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    fn = lt.write_vtk({"p": a, "q": (a * 2).astype(np.float64)}, id=7, filename_base=str(tmp_path / "o"))
    assert fn.endswith("o_00000007.vtr")
    raw = open(fn, "rb").read()
    head = raw[:raw.index(b"_", raw.index(b"<AppendedData")) + 1].decode()
    assert head.startswith('<?xml version="1.0"?>\n<VTKFile type="RectilinearGrid" version="1.0" '
                           'byte_order="LittleEndian" header_type="UInt64">')
    assert 'WholeExtent="0 1 0 2 0 3"' in head and '<PointData scalars="p">' in head
    # offsets: 3 coordinate blocks (int64) then p (float32) then q (float64), each with an 8-byte size
    offs = [8 + 2 * 8, 8 + 3 * 8, 8 + 4 * 8, 8 + 24 * 4]
    want = np.cumsum([0] + offs)
    import re
    assert [int(m) for m in re.findall(r'offset="(\d+)"', head)] == list(want)
    body = raw[len(head.encode()):]
    assert np.frombuffer(body, "<u8", 1, int(want[3]))[0] == 24 * 4
    p_flat = np.frombuffer(body, "<f4", 24, int(want[3]) + 8)
    assert np.array_equal(p_flat, a.transpose(2, 1, 0).reshape(-1))       # x fastest
    x, y, z, arrays = lt.read_vtr(fn)
    assert np.array_equal(arrays["p"], a) and np.array_equal(arrays["q"], a * 2.0)
    assert raw.endswith(b"\n</AppendedData>\n</VTKFile>\n")
    with pytest.raises(ValueError):
        lt.write_vtk({"p": a, "bad": np.zeros([2, 3])}, filename_base=str(tmp_path / "o"))


def test_write_vtk_reference_test(tmp_path):
    """tests/reporter/test_write_vtk.py:6-13"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=[16, 16], reynolds_number=10, mach_number=0.05)
    p, u = flow.initial_pu()
    lt.write_vtk({"p": ctx.convert_to_ndarray(p[0, ..., None])}, id=1, filename_base=tmp_path / "output")
    assert os.path.isfile(tmp_path / "output_00000001.vtr")


def test_vtk_reporter_mask(tmp_path):
    """tests/reporter/test_vtk_reporter_mask.py:6-20, without the time steps (those need the GPU)"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.PoiseuilleFlow2D(context=ctx, resolution=16, reynolds_number=10, mach_number=0.05)
    bnds = [None] + sorted(flow.boundaries, key=lambda b: str(b))
    ncm, _ = lt.build_masks(flow, bnds, ctx)
    rep = lt.VTKReporter(interval=1, filename_base=tmp_path / "output")
    rep.output_mask(_fake_simulation(flow, ncm))
    x, y, z, arrays = lt.read_vtr(tmp_path / "output_mask.vtr")
    assert arrays["mask"].shape == (16, 16, 1) and arrays["mask"].dtype == np.int64
    assert np.array_equal(arrays["mask"][..., 0], ncm.numpy())


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_error_reporter_matches_reference(dt):
    """ErrorReporter on the reference's final state gives the reference's last row"""
    case = load_case(f"flow_tgv2d_d2q9_bgk_{dt}")
    meta = case["meta"]
    flow = build_flow(meta)
    flow.f = torch.from_numpy(case["f_out"])
    flow.i = meta["n_steps"]
    rep = lt.ErrorReporter(flow.analytic_solution, interval=1, out=None)
    rep(_fake_simulation(flow))
    assert len(case["err_out"]) == meta["n_steps"] + 1
    tol = 1e-3 if dt == "f32" else 1e-9
    assert rep.out[0] == pytest.approx(list(case["err_out"][-1]), rel=tol)
    import io
    buf = io.StringIO()
    rep2 = lt.ErrorReporter(flow.analytic_solution, interval=3, out=buf)   # 8 % 3 != 0: header only
    rep2(_fake_simulation(flow))
    assert buf.getvalue() == "#error_u         error_p\n"


def test_hdf5_and_image_output_are_refused():
    """out of scope: the names exist and say so"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=[8, 8], reynolds_number=10, mach_number=0.05)
    with pytest.raises(NotImplementedError, match="out of scope"):
        lt.HDF5Reporter(flow, lt.BGKCollision(tau=0.6), interval=1)
    with pytest.raises(NotImplementedError, match="out of scope"):
        lt.LettuceDataset("nothing.h5")
    with pytest.raises(NotImplementedError, match="out of scope"):
        lt.write_image("x.png", np.zeros([4, 4]))


# ---- set-up helpers restated from the reference's tests -----------------------------------------
@pytest.mark.parametrize("stencil", [lt.D2Q9, lt.D3Q27])
def test_divergence(stencil):
    """tests/flow/test_divergence.py:5-36: the decaying-turbulence start is divergence free"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    d = stencil().d
    flow = lt.DecayingTurbulence(context=ctx, resolution=[50] * d, reynolds_number=1, mach_number=0.05,
                                 ic_energy=0.5, stencil=stencil, randseed=0,
                                 initialize_pressure=False)
    dx = flow.units.convert_length_to_pu(1.0)
    ekin = flow.units.convert_incompressible_energy_to_pu(torch.sum(flow.incompressible_energy())) * dx ** d
    u = flow.u_pu
    div = sum(float(lt.torch_gradient(u[c], dx=dx, order=6)[c].sum()) for c in range(d))
    assert flow.ic_energy == pytest.approx(float(ekin), rel=1)
    assert 0 == pytest.approx(div, abs=2e-3)


def test_pressure_poisson():
    """tests/flow/test_pressure_poisson.py:5-23 (D2Q9; the reference skips D3Q27)"""
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(context=ctx, resolution=32, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    p0, u = flow.initial_pu()
    u = flow.units.convert_velocity_to_lu(u)
    rho0 = flow.units.convert_pressure_pu_to_density_lu(p0)
    rho = lt.pressure_poisson(flow.units, u, torch.ones_like(rho0))
    pfinal = flow.units.convert_density_lu_to_pressure_pu(rho).numpy()
    assert pfinal == pytest.approx(p0.numpy(), rel=0.0, abs=0.05)


def test_grid_fine_to_coarse_and_get_subclasses():
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    fine = lt.TaylorGreenVortex(ctx, resolution=[32, 32], reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    coarse = lt.TaylorGreenVortex(ctx, resolution=[16, 16], reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    tf, tc = fine.units.relaxation_parameter_lu, coarse.units.relaxation_parameter_lu
    fc = lt.grid_fine_to_coarse(fine, fine.f, tf, tc)
    assert list(fc.shape) == [9, 16, 16]
    # density and velocity of every second node are kept
    assert torch.allclose(coarse.rho(fc), fine.rho(fine.f)[:, ::2, ::2], atol=1e-14)
    assert torch.allclose(coarse.u(fc), fine.u(fine.f)[:, ::2, ::2], atol=1e-14)
    with pytest.raises(lt.LettuceException):
        lt.grid_fine_to_coarse(fine, fine.f[0], tf, tc)
    import lettuce_b200._collision as cm
    names = {c.__name__ for c in lt.get_subclasses(lt.Collision, cm)}
    assert {"BGKCollision", "TRTCollision", "KBCCollision", "NoCollision"} <= names
