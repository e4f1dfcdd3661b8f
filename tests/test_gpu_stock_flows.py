"""GPU parity of the reference's STOCK FLOWS run through lettuce_b200's own flow classes
(DoublyPeriodicShear2D, Cavity2D, PoiseuilleFlow2D + Guo force, DecayingTurbulence, TaylorGreenVortex)
and of the output side (VTKReporter, ErrorReporter, EnergySpectrum) inside a real `lt.Simulation`,
against golden files from the unmodified reference (tests/golden/make_golden_flows.py).
The same files also run through tests/test_gpu_parity.py::test_golden with the boundary values taken
from the file; here the flow classes themselves build masks, boundary values and the force."""
import os

import numpy as np
import pytest
import torch

from golden_util import DT, TOL, golden_names, load_case, rel_err

pytestmark = pytest.mark.gpu

FLOW_CASES = [n for n in golden_names() if n.startswith("flow_")]


def _build(meta):
    import lettuce_b200 as lt
    args = dict(meta["flow"]["args"])
    if "stencil" in args:
        args["stencil"] = getattr(lt, args["stencil"])
    ctx = lt.Context(device="cuda", dtype=DT[meta["dtype"]], use_native=True)
    return lt, getattr(lt, meta["flow"]["cls"])(ctx, **args)


def _collision(lt, flow, meta):
    force = None
    if meta.get("force"):                         # tests/collision/test_force.py:12-22
        force = lt.Guo(flow, tau=meta["tau"], acceleration=flow.acceleration)
    return lt.BGKCollision(tau=meta["tau"], force=force)


@pytest.mark.parametrize("name", FLOW_CASES)
def test_stock_flow_matches_reference(name):
    case = load_case(name)
    meta = case["meta"]
    lt, flow = _build(meta)
    # (1) the flow's own initialisation on the device (lb_initialize; for the 2-D decaying turbulence
    # preceded by the pressure-Poisson Jacobi iteration in torch on the device)
    assert flow.f.is_cuda and flow.f.shape == case["f0"].shape
    init_tol = TOL[meta["dtype"]]
    if meta["flow"]["cls"] == "DecayingTurbulence" and meta["flow"]["args"].get("initialize_pressure", True) \
            and len(meta["res"]) == 2:
        init_tol = 1e-9      # Jacobi stops on a threshold: allow one iteration more or less than the CPU run
    assert rel_err(flow.f.cpu().numpy(), case["f0"]) <= init_tol
    assert float(flow.units.relaxation_parameter_lu) == pytest.approx(meta["tau"], rel=1e-12)
    # (2) identical initial field (SURVEY 8c parity procedure), then n steps through lt.Simulation
    flow.f.copy_(torch.from_numpy(case["f0"]))
    sim = lt.Simulation(flow, _collision(lt, flow, meta), [])
    if "ncm" in case:
        assert np.array_equal(sim.no_collision_mask.cpu().numpy(), case["ncm"])
        assert np.array_equal(sim.no_streaming_mask.cpu().numpy(), case["nsm"])
    sim(meta["n_steps"])
    assert flow.i == meta["n_steps"]
    assert rel_err(flow.f.cpu().numpy(), case["f_out"]) <= TOL[meta["dtype"]]


@pytest.mark.parametrize("name", [n for n in FLOW_CASES if "spectrum0" in load_case(n)])
def test_energy_spectrum_reporter(name):
    """ObservableReporter(EnergySpectrum) inside the simulation: rows at step 0 and n equal the
    reference's spectra (u from the moments kernel, FFT on the device)"""
    case = load_case(name)
    meta = case["meta"]
    lt, flow = _build(meta)
    flow.f.copy_(torch.from_numpy(case["f0"]))
    n = meta["n_steps"]
    rep = lt.ObservableReporter(lt.EnergySpectrum(flow), interval=n, out=None)
    lt.Simulation(flow, _collision(lt, flow, meta), [rep])(n)
    assert [r[0] for r in rep.out] == [0, n]
    tol = 2e-4 if meta["dtype"] == "f32" else 1e-10
    for row, key in ((rep.out[0], "spectrum0"), (rep.out[1], "spectrum_out")):
        got = np.asarray(row[2:])
        assert got.shape == case[key].shape
        assert np.abs(got - case[key]).max() <= tol * np.abs(case[key]).max()


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_error_reporter_in_simulation(dt):
    """ErrorReporter every step of the 2-D Taylor-Green run == the reference's rows"""
    case = load_case(f"flow_tgv2d_d2q9_bgk_{dt}")
    meta = case["meta"]
    lt, flow = _build(meta)
    flow.f.copy_(torch.from_numpy(case["f0"]))
    rep = lt.ErrorReporter(flow.analytic_solution, interval=1, out=None)
    lt.Simulation(flow, _collision(lt, flow, meta), [rep])(meta["n_steps"])
    got, ref = np.asarray(rep.out), case["err_out"]
    assert got.shape == ref.shape == (meta["n_steps"] + 1, 2)
    if dt == "f64":
        assert np.abs(got - ref).max() <= 1e-9 * np.abs(ref).max()
    else:
        # fp32: u agrees to round-off; p_pu = cs^2 (rho - 1) (u_pu / u_lu)^2 carries the round-off of
        # rho ~ 1 magnified by (u_pu / u_lu)^2 ~ 1200, a noise floor of ~ 2.4e-5 on err_p (measured 1.4e-5)
        un = meta["units"]
        floor = np.finfo(np.float32).eps * (un["velocity_pu"] / un["velocity_lu"]) ** 2 / 3
        assert np.abs(got[:, 0] - ref[:, 0]).max() <= 5e-3 * np.abs(ref[:, 0]).max()
        assert np.abs(got[:, 1] - ref[:, 1]).max() <= floor


@pytest.mark.parametrize("name", ["flow_cavity2d_d2q9_bgk_f32", "flow_decay3d_d3q19_bgk_f64"])
def test_vtk_reporter_in_simulation(name, tmp_path):
    """tests/reporter/test_vtk_reporter_no_mask.py / _mask.py: files for steps 0..2 and the mask file
    exist; beyond the reference's test, the arrays read back equal p / u of the state at that step."""
    case = load_case(name)
    meta = case["meta"]
    lt, flow = _build(meta)
    flow.f.copy_(torch.from_numpy(case["f0"]))
    rep = lt.VTKReporter(interval=1, filename_base=str(tmp_path / "data" / "output"))
    sim = lt.Simulation(flow, _collision(lt, flow, meta), [rep])
    d = len(meta["res"])
    shape3 = tuple(meta["res"]) + (1,) * (3 - d)
    if sim.no_collision_mask is not None:
        rep.output_mask(sim)
        _, _, _, m = lt.read_vtr(tmp_path / "data" / "output_mask.vtr")
        assert np.array_equal(m["mask"].reshape(meta["res"]), case["ncm"])
    sim(2)
    for i in range(3):
        assert os.path.isfile(tmp_path / "data" / f"output_{i:08d}.vtr")
    tol = 1e-4 if meta["dtype"] == "f32" else 1e-10
    # step 0: the reference's p_pu / u_pu of f0
    x, y, z, a0 = lt.read_vtr(tmp_path / "data" / "output_00000000.vtr")
    assert (x.size, y.size, z.size) == shape3 and list(a0) == ["p"] + ["ux", "uy", "uz"][:d]
    for c in range(d):
        ref = case["u_pu0"][c]
        assert np.abs(a0[f"u{'xyz'[c]}"].reshape(meta["res"]) - ref).max() <= tol * max(np.abs(case["u_pu0"]).max(), 1e-3)
    # step 2: the state the simulation holds now
    _, _, _, a2 = lt.read_vtr(tmp_path / "data" / "output_00000002.vtr")
    # (flow.p_pu / flow.u_pu: moments kernel + torch conversions; the file: lb_output_fields -- same values up
    # to the rounding of one fused scale factor)
    u_now, p_now = flow.u_pu.cpu().numpy(), flow.p_pu.cpu().numpy()
    ulp = 8 * float(np.finfo(case["f0"].dtype).eps)
    assert np.abs(a2["p"].reshape(meta["res"]) - p_now[0]).max() <= 200 * ulp * max(np.abs(p_now).max(), 1e-3) + 1e-30
    for c in range(d):
        assert np.abs(a2[f"u{'xyz'[c]}"].reshape(meta["res"]) - u_now[c]).max() <= ulp * np.abs(u_now).max()
        assert np.array_equal(rep.point_dict[f"u{'xyz'[c]}"], a2[f"u{'xyz'[c]}"])


def test_vtk_reporter_background_pipeline_equals_inline(tmp_path):
    """the overlapped report path (kernel on the stepping stream, D2H on a second stream into double-buffered
    pinned staging, file written by a background thread while the next steps run) writes byte for byte the
    files of the inline path -- 9 reports through 2 slots, i.e. with slot reuse and back-pressure"""
    import lettuce_b200 as lt
    files = {}
    for mode in (True, False):
        ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=[48, 40, 32], reynolds_number=400, mach_number=0.1, stencil=lt.D3Q19)
        base = tmp_path / ("bg" if mode else "inline") / "out"
        rep = lt.VTKReporter(interval=3, filename_base=str(base), asynchronous=mode, buffers=2)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [rep])
        sim(14)
        sim(10)                                   # a second call: the pipeline survives flush()
        rep.close()
        names = sorted(os.listdir(base.parent))
        assert names == [f"out_{i:08d}.vtr" for i in range(0, 25, 3)]
        files[mode] = [open(base.parent / n, "rb").read() for n in names]
        _, _, _, last = lt.read_vtr(base.parent / names[-1])
        assert np.array_equal(rep.point_dict["ux"], last["ux"])
    assert files[True] == files[False]


def test_initialize_fneq_improves_tgv():
    """tests/flow/test_initialize_fneq.py:64-87: with the non-equilibrium start the 2-D Taylor-Green run
    tracks the analytic solution better than with the equilibrium start"""
    import lettuce_b200 as lt
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    errs = []
    for fneq in (True, False):
        flow = lt.TaylorGreenVortex(ctx, resolution=32, reynolds_number=1000, mach_number=0.1, stencil=lt.D2Q9,
                                    initialize_fneq=fneq)
        rep = lt.ErrorReporter(flow.analytic_solution, interval=1, out=None)
        lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [rep])(10)
        errs.append(np.mean(np.abs(rep.out), axis=0))
    assert errs[0][0] < errs[1][0]


@pytest.mark.parametrize("stencil,res", [("D2Q9", [37, 45]), ("D2Q9", [64, 96]), ("D3Q19", [33, 7, 70]),
                                         ("D3Q27", [5, 40, 35]), ("D3Q19", [32, 32, 64]), ("D3Q27", [2, 3, 5])])
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_output_fields_kernel(stencil, res, dt):
    """lb_output_fields (one pass: moments, PU conversion, reordering to x-fastest) against flow.p_pu / flow.u_pu
    put into Fortran order on the host, on a random state, at sizes with partial 32 x 32 tiles"""
    import lettuce_b200 as lt
    ctx = lt.Context(device="cuda", dtype=DT[dt], use_native=True)
    S = getattr(lt, stencil)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=100, mach_number=0.05, stencil=S,
                                initialize_fneq=False)
    g = torch.Generator().manual_seed(11)
    w = torch.tensor(S().w, dtype=DT[dt]).reshape([S().q] + [1] * len(res))
    flow.f = (w * (1.0 + 0.1 * torch.rand([S().q, *res], generator=g, dtype=DT[dt]))).to("cuda").contiguous()
    out = flow.output_fields_pu()
    d = len(res)
    n = int(np.prod(res))
    assert out.is_cuda and list(out.shape) == [1 + d, n] and out.dtype == DT[dt]
    got = out.cpu().numpy()
    p, u = flow.p_pu.cpu().numpy(), flow.u_pu.cpu().numpy()
    eps = float(np.finfo(got.dtype).eps)
    # p = (rho - 1) * scale: the subtraction is exact in both routes, the scaling differs by a rounding or two
    assert np.abs(got[0] - p[0].reshape(-1, order="F")).max() <= 8 * eps * np.abs(p).max()
    for c in range(d):
        assert np.abs(got[1 + c] - u[c].reshape(-1, order="F")).max() <= 8 * eps * np.abs(u).max()
