"""Golden vectors for the reference's STOCK FLOWS (the callers of the hot path) and the output-side
observables, produced by RUNNING THE UNMODIFIED REFERENCE (build container only):

    python tests/golden/make_golden_flows.py

Each `flow_*.npz` has the format of make_golden.py (f0 = the populations the reference's flow
initialised, masks, boundary values in lattice units, f_out after n steps through
`lt.Simulation`), so the oracle and C-ABI parity tests consume it unchanged, plus

  meta["flow"]      class name and constructor arguments, so that tests can build the same flow with
                    lettuce_b200 and compare its initial populations, masks and boundary values
  meta["units"]     characteristic length / velocity in LU and PU of the reference's flow
  spectrum0/_out    EnergySpectrum (observable_reporter.py:71-137) of f0 / f_out   (periodic flows)
  p_pu0, u_pu0      flow.p_pu / flow.u_pu of f0: what VTKReporter writes (vtk_reporter.py:29-47)
  err_out           ErrorReporter row [err_u, err_p] after n steps (TGV 2-D only)

CouetteFlow2D is not among them: its characteristic PU velocity is 0 (couette.py:21,42), the LU
conversion divides by it and the reference's populations are NaN from the start.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.reference_shim import import_reference  # noqa: E402

lt = import_reference()
torch.set_num_threads(4)
DT = {"f32": torch.float32, "f64": torch.float64}


def run_flow(name, cls_name, dtype, args, n_steps, force=None, spectrum=False, error=False):
    ctx = lt.Context(device="cpu", dtype=DT[dtype], use_native=False)
    kw = dict(args)
    stencil_name = kw.pop("stencil", None)
    if stencil_name:
        kw["stencil"] = getattr(lt, stencil_name)
    cls = lt.Cavity2D if cls_name == "Cavity2D" else getattr(lt, cls_name)   # Cavity2D: module has no __all__
    flow = cls(ctx, **kw)
    tau = float(flow.units.relaxation_parameter_lu)
    meta = dict(name=name, stencil=type(flow.stencil).__name__, dtype=dtype, res=list(flow.resolution),
                collision="bgk", n_steps=n_steps, tau=tau, tau_minus=1.0, boundaries=[],
                flow=dict(cls=cls_name, args=args),
                units=dict(length_lu=float(flow.units.characteristic_length_lu),
                           length_pu=float(flow.units.characteristic_length_pu),
                           velocity_lu=float(flow.units.characteristic_velocity_lu),
                           velocity_pu=float(flow.units.characteristic_velocity_pu)))
    fobj = None
    if force == "guo":      # tests/collision/test_force.py:12-22
        acc = [float(a) for a in flow.acceleration]
        fobj = lt.Guo(flow, tau=tau, acceleration=flow.acceleration)
        meta["force"] = dict(kind="guo", tau=tau, acceleration=acc)
    reporters = []
    if error:
        reporters.append(lt.ErrorReporter(flow.analytic_solution, interval=1, out=None))
    sim = lt.Simulation(flow, lt.BGKCollision(tau=tau, force=fobj), reporters)
    out = {"f0": flow.f.numpy().copy(),
           "p_pu0": flow.p_pu.numpy().copy(), "u_pu0": flow.u_pu.numpy().copy()}
    if sim.no_collision_mask is not None:
        out["ncm"] = sim.no_collision_mask.numpy().copy()
        out["nsm"] = sim.no_streaming_mask.numpy().copy()
    for i, b in enumerate(sim.boundaries[1:], start=1):
        if isinstance(b, lt.EquilibriumBoundaryPU):
            out[f"b{i}_rho"] = np.asarray(flow.units.convert_pressure_pu_to_density_lu(b.pressure).numpy()).copy()
            out[f"b{i}_u"] = flow.units.convert_velocity_to_lu(b.velocity).numpy().copy()
            meta["boundaries"].append(dict(kind="equilibrium", index=i))
        else:
            meta["boundaries"].append(dict(kind="bounce_back", index=i))
    if spectrum:
        out["spectrum0"] = lt.EnergySpectrum(flow)().numpy().copy()
    sim(n_steps)
    out["f_out"] = flow.f.numpy().copy()
    if spectrum:
        out["spectrum_out"] = lt.EnergySpectrum(flow)().numpy().copy()
    if error:
        out["err_out"] = np.asarray(reporters[0].out, dtype=np.float64)
    assert np.isfinite(out["f_out"]).all(), name
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: tau {tau:.6f}  f_out sum {out['f_out'].sum():.12g}")


def main():
    for dt in ("f32", "f64"):
        run_flow(f"flow_shear2d_d2q9_bgk_{dt}", "DoublyPeriodicShear2D", dt,
                 dict(resolution=[32, 24], reynolds_number=500, mach_number=0.05, shear_layer_width=20,
                      initial_perturbation_magnitude=0.05), 10, spectrum=True)
        run_flow(f"flow_cavity2d_d2q9_bgk_{dt}", "Cavity2D", dt,
                 dict(resolution=[24, 20], reynolds_number=100, mach_number=0.05), 20)
        run_flow(f"flow_poiseuille2d_d2q9_bgkguo_{dt}", "PoiseuilleFlow2D", dt,
                 dict(resolution=16, reynolds_number=1, mach_number=0.02), 20, force="guo")
        run_flow(f"flow_tgv2d_d2q9_bgk_{dt}", "TaylorGreenVortex", dt,
                 dict(resolution=[24, 24], reynolds_number=100, mach_number=0.05, stencil="D2Q9"), 8,
                 spectrum=True, error=True)
    # decaying turbulence: seeded numpy draws; 2-D runs the pressure-Poisson Jacobi iteration (fp64)
    run_flow("flow_decay2d_d2q9_bgk_f64", "DecayingTurbulence", "f64",
             dict(resolution=[24, 24], reynolds_number=200, mach_number=0.05, k0=5, ic_energy=0.5,
                  randseed=7), 10, spectrum=True)
    run_flow("flow_decay2d_nopressure_d2q9_bgk_f32", "DecayingTurbulence", "f32",
             dict(resolution=[24, 24], reynolds_number=200, mach_number=0.05, k0=5, ic_energy=0.5,
                  randseed=7, initialize_pressure=False), 10, spectrum=True)
    run_flow("flow_decay3d_d3q19_bgk_f64", "DecayingTurbulence", "f64",
             dict(resolution=[12, 12, 12], reynolds_number=200, mach_number=0.05, k0=3, ic_energy=0.5,
                  randseed=11, stencil="D3Q19"), 6, spectrum=True)
    run_flow("flow_tgv3d_d3q27_bgk_f64", "TaylorGreenVortex", "f64",
             dict(resolution=[12, 12, 12], reynolds_number=100, mach_number=0.05, stencil="D3Q27"), 6,
             spectrum=True)


if __name__ == "__main__":
    main()
