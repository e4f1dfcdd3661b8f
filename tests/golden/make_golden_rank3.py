"""Golden vectors for the collisions of SURVEY section 8f rank 3, produced by the UNMODIFIED reference in
this container: RegularizedCollision (lettuce/ext/_collision/regularized_collision.py) and
SmagorinskyCollision (smagorinsky_collision.py, force=None).  Same file format as make_golden.py;
collision = 'regularized' (meta tau = what the reference reads from the units, :19) or 'smagorinsky'
(meta tau, meta tau_minus = the Smagorinsky constant).

    python tests/golden/make_golden_rank3.py        # needs /root/reference; writes rank3_*.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (imports the reference through oracle/reference_shim.py)

lt = mg.lt
DT = mg.DT
_stock_make_collision = mg.make_collision


def make_collision(name, flow, tau, tau_minus):
    if name == "regularized":
        return lt.RegularizedCollision(tau=123.0)        # ignored by the reference (:19)
    if name == "smagorinsky":
        return lt.SmagorinskyCollision(tau=tau, smagorinsky_constant=tau_minus)
    return _stock_make_collision(name, flow, tau, tau_minus)


mg.make_collision = make_collision


def units_tau(res, re, ma):
    return lt.UnitConversion(re, ma, characteristic_length_lu=res[0]).relaxation_parameter_lu


def run_force_case(name, stencil, dtype, res, f0, n_steps, tau, kind, force_tau, acceleration, re=100.0, ma=0.05):
    """BGKCollision(tau, force=Guo | ShanChen(flow, force_tau, acceleration)) (bgk_collision.py:17-22);
    meta carries collision 'bgk' plus force = {kind, tau, acceleration}"""
    import json
    tdt = DT[dtype]
    ctx = lt.Context(device="cpu", dtype=tdt, use_native=False)
    flow = mg.GoldenFlow(ctx, res, getattr(lt, stencil), f0.to(tdt), re=re, ma=ma)
    force = (lt.Guo if kind == "guo" else lt.ShanChen)(flow, force_tau, acceleration)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=tau, force=force), [])
    out = {"f0": f0.to(tdt).numpy().copy()}
    meta = dict(name=name, stencil=stencil, dtype=dtype, res=list(res), collision="bgk", n_steps=n_steps,
                tau=tau, tau_minus=1.0, boundaries=[],
                force=dict(kind=kind, tau=force_tau, acceleration=[float(a) for a in acceleration]))
    sim(n_steps)
    out["f_out"] = flow.f.numpy().copy()
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: f_out sum {out['f_out'].sum():.12g}")


def main_forces():
    rng = np.random.default_rng(20261107)
    acc = {2: [2e-5, -1e-5], 3: [2e-5, -1e-5, 3e-5]}
    for st, res in (("D2Q9", [24, 20]), ("D3Q19", [10, 8, 12]), ("D3Q27", [8, 6, 8])):
        S = getattr(lt, st)()
        for dt in ("f32", "f64"):
            f0, tau = mg.tgv_field(getattr(lt, st), res if S.d == 3 else res, DT[dt], re=100.0, ma=0.05) \
                if res[0] == res[-1] else (mg.positive_field(rng, S.q, res, DT[dt], amp=0.02, stencil=S), 0.8)
            for kind in ("guo", "shan_chen"):
                # the Force's tau differs from the collision's in one of the two, to pin which is used where
                ftau = tau if kind == "guo" else 0.9
                run_force_case(f"rank3_force_{st.lower()}_{kind}_{dt}", st, dt, res, f0, 6, tau, kind, ftau, acc[S.d])
    # a strong force on a far-from-equilibrium field, one step (source term visible well above round-off)
    S = lt.D3Q19()
    f0 = mg.positive_field(rng, S.q, [6, 5, 7], torch.float64, amp=0.1, stencil=S)
    run_force_case("rank3_force_strong_d3q19_guo_f64", "D3Q19", "f64", [6, 5, 7], f0, 2, 0.7, "guo", 0.6, [1e-2, 2e-2, -1e-2])


def main():
    rng = np.random.default_rng(20261106)
    # TGV initial fields, 10 steps (the flows of BASELINE configs 1, 2, 4, reduced)
    for st, res in (("D2Q9", [32, 32]), ("D3Q19", [16, 16, 16]), ("D3Q27", [12, 12, 12])):
        for dt in ("f32", "f64"):
            f0, tau = mg.tgv_field(getattr(lt, st), res, DT[dt], re=100.0, ma=0.05)
            mg.run_case(f"rank3_tgv_{st.lower()}_regularized_{dt}", st, dt, res, f0, "regularized", 10,
                        tau=units_tau(res, 100.0, 0.05), re=100.0, ma=0.05)
            mg.run_case(f"rank3_tgv_{st.lower()}_smagorinsky_{dt}", st, dt, res, f0, "smagorinsky", 10,
                        tau=tau, tau_minus=0.17, re=100.0, ma=0.05)
    # random positive fields far from equilibrium, 2 steps: large shear -> tau_eff visibly above tau
    for st, res in (("D2Q9", [10, 14]), ("D3Q19", [6, 5, 7]), ("D3Q27", [5, 6, 7])):
        S = getattr(lt, st)()
        for dt in ("f32", "f64"):
            f0 = mg.positive_field(rng, S.q, res, DT[dt], amp=0.1, stencil=S)
            mg.run_case(f"rank3_rand_{st.lower()}_regularized_{dt}", st, dt, res, f0, "regularized", 2,
                        tau=units_tau(res, 20.0, 0.05), re=20.0, ma=0.05)
            mg.run_case(f"rank3_rand_{st.lower()}_smagorinsky_{dt}", st, dt, res, f0, "smagorinsky", 2,
                        tau=0.7, tau_minus=0.4)
    # with boundaries: bounce-back solids + equilibrium inlet, D3Q19 Smagorinsky / D2Q9 regularized
    res = [12, 10, 8]
    S = lt.D3Q19()
    f0 = mg.positive_field(rng, S.q, res, torch.float64, amp=0.02, stencil=S)
    solid = torch.tensor(rng.random(res) < 0.1)
    inlet = torch.zeros(res, dtype=torch.bool)
    inlet[0] = True
    spec = [dict(kind="bounce_back", mask=solid),
            dict(kind="equilibrium", mask=inlet, velocity_pu=[0.3, 0.0, 0.1], pressure_pu=0.002)]
    mg.run_case("rank3_box_d3q19_smagorinsky_f64", "D3Q19", "f64", res, f0, "smagorinsky", 5, tau=0.6,
                tau_minus=0.17, boundaries_spec=spec, re=50.0, ma=0.05)
    res = [20, 12]
    S = lt.D2Q9()
    f0 = mg.positive_field(rng, S.q, res, torch.float64, amp=0.02, stencil=S)
    solid = torch.tensor(rng.random(res) < 0.1)
    mg.run_case("rank3_box_d2q9_regularized_f64", "D2Q9", "f64", res, f0, "regularized", 5,
                tau=units_tau(res, 50.0, 0.05), boundaries_spec=[dict(kind="bounce_back", mask=solid)],
                re=50.0, ma=0.05)


if __name__ == "__main__":
    if "--forces-only" not in sys.argv:
        main()
    main_forces()
