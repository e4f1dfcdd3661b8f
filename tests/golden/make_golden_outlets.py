"""Golden vectors for the outlet boundaries, produced by the UNMODIFIED
reference in this container:  AntiBounceBackOutlet (lettuce/ext/_boundary/anti_bounce_back_outlet.py)
as in the stock `lt.Obstacle.boundaries` (obstacle.py:107-122) and EquilibriumOutletP
(equilibrium_outlet_p.py).  Same file format as make_golden.py; boundary entries of kind
'abb_outlet' / 'eq_outlet_p' carry `direction` (and `rho_outlet`).

    python tests/golden/make_golden_outlets.py        # needs /root/reference; writes outlet_*.npz
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (imports the reference through oracle/reference_shim.py)

lt = mg.lt
DT = mg.DT


def run_outlet_case(name, stencil, dtype, res, f0, collision, n_steps, build, tau=None, re=100.0, ma=0.05):
    """build(ctx, flow) -> list of reference boundary objects (any order; the reference sorts them)"""
    tdt = DT[dtype]
    ctx = lt.Context(device="cpu", dtype=tdt, use_native=False)
    flow = mg.GoldenFlow(ctx, res, getattr(lt, stencil), f0.to(tdt), re=re, ma=ma)
    if collision == "kbc" or tau is None:
        tau = flow.units.relaxation_parameter_lu
    flow.boundaries.extend(build(ctx, flow))
    sim = lt.Simulation(flow, mg.make_collision(collision, flow, tau, 1.0), [])
    out = {"f0": f0.to(tdt).numpy().copy(),
           "ncm": sim.no_collision_mask.numpy().copy(),
           "nsm": sim.no_streaming_mask.numpy().copy()}
    meta = dict(name=name, stencil=stencil, dtype=dtype, res=list(res), collision=collision,
                n_steps=n_steps, tau=tau, tau_minus=1.0, boundaries=[])
    for i, b in enumerate(sim.boundaries[1:], start=1):
        if isinstance(b, lt.EquilibriumOutletP):          # subclass of AntiBounceBackOutlet: test first
            meta["boundaries"].append(dict(kind="eq_outlet_p", index=i, direction=b._direction,
                                           rho_outlet=float(b.rho_outlet)))
        elif isinstance(b, lt.AntiBounceBackOutlet):
            meta["boundaries"].append(dict(kind="abb_outlet", index=i, direction=b._direction))
        elif isinstance(b, lt.EquilibriumBoundaryPU):
            out[f"b{i}_rho"] = flow.units.convert_pressure_pu_to_density_lu(b.pressure).numpy().copy()
            out[f"b{i}_u"] = flow.units.convert_velocity_to_lu(b.velocity).numpy().copy()
            meta["boundaries"].append(dict(kind="equilibrium", index=i))
        else:
            out[f"b{i}_mask"] = b._mask.numpy().copy() if hasattr(b, "_mask") else b.mask.numpy().copy()
            meta["boundaries"].append(dict(kind="bounce_back", index=i))
    sim(n_steps)
    out["f_out"] = flow.f.numpy().copy()
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: boundaries {[b['kind'] for b in meta['boundaries']]} f_out sum {out['f_out'].sum():.12g}")


def outlet(cls, direction, flow, **kw):
    b = cls(list(direction), flow, **kw)
    b._direction = [int(v) for v in direction]
    return b


def main():
    rng = np.random.default_rng(20261105)

    # ---- 2-D: the stock Obstacle boundary list (inlet equilibrium, ABB outlet on +x, bounce-back cylinder)
    for dt in ("f32", "f64"):
        res = [48, 24]
        ctx = lt.Context(device="cpu", dtype=DT[dt], use_native=False)
        ob = lt.Obstacle(ctx, res, reynolds_number=300, mach_number=0.05, domain_length_x=1.0, stencil=lt.D2Q9)
        x, y = ob.grid
        cyl = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.1 * y.max()) ** 2
        ob.mask = cyl
        ob.initialize()
        f0 = ob.f.clone()
        tau = max(ob.units.relaxation_parameter_lu, 0.55)
        inlet = torch.zeros(res, dtype=torch.bool)
        inlet[0, :] = True

        def stock(ctx, flow, cyl=cyl, inlet=inlet):
            return [lt.EquilibriumBoundaryPU(ctx, inlet, [1.0, 0.0], 0.0),
                    outlet(lt.AntiBounceBackOutlet, [1, 0], flow),
                    lt.BounceBackBoundary(cyl)]
        run_outlet_case(f"outlet_abb_d2q9_bgk_{dt}", "D2Q9", dt, res, f0, "bgk", 20, stock, tau=tau, re=300.0)

        def eqp(ctx, flow, cyl=cyl, inlet=inlet):
            return [lt.EquilibriumBoundaryPU(ctx, inlet, [1.0, 0.0], 0.0),
                    outlet(lt.EquilibriumOutletP, [1, 0], flow, rho_outlet=1.0),
                    lt.BounceBackBoundary(cyl)]
        run_outlet_case(f"outlet_eqp_d2q9_bgk_{dt}", "D2Q9", dt, res, f0, "bgk", 20, eqp, tau=tau, re=300.0)

    # ---- 2-D, outlet on the -y side (the contiguous axis of a 2-D lattice), TRT-free BGK on a random field
    S = lt.D2Q9()
    f0 = mg.positive_field(rng, 9, [14, 20], torch.float64, amp=0.02, stencil=S)
    run_outlet_case("outlet_abb_negy_d2q9_bgk_f64", "D2Q9", "f64", [14, 20], f0, "bgk", 6,
                    lambda ctx, flow: [outlet(lt.AntiBounceBackOutlet, [0, -1], flow)], tau=0.8)

    # ---- 3-D: outlet on +z (contiguous axis) with D3Q19 BGK, on -x with D3Q27 KBC; random solids
    for st, c, direction, mk in (("D3Q19", "bgk", [0, 0, 1], lt.AntiBounceBackOutlet),
                                 ("D3Q27", "kbc", [-1, 0, 0], lt.AntiBounceBackOutlet),
                                 ("D3Q19", "bgk", [0, 1, 0], lt.EquilibriumOutletP)):
        res = [12, 10, 8]
        S = getattr(lt, st)()
        f0 = mg.positive_field(rng, S.q, res, torch.float64, amp=0.02, stencil=S)
        solid = torch.tensor(rng.random(res) < 0.08)
        tag = "abb" if mk is lt.AntiBounceBackOutlet else "eqp"

        def build(ctx, flow, solid=solid, direction=direction, mk=mk):
            kw = dict(rho_outlet=1.01) if mk is lt.EquilibriumOutletP else {}
            return [lt.BounceBackBoundary(solid), outlet(mk, direction, flow, **kw)]
        run_outlet_case(f"outlet_{tag}_{st.lower()}_{c}_f64", st, "f64", res, f0, c, 5, build, tau=0.8, re=50.0)


if __name__ == "__main__":
    main()
