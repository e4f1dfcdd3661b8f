"""Generate the golden vectors in tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Every case drives the reference through its public API (`lt.Simulation(flow, collision,
reporter)(n)` on `Context(device='cpu', use_native=False)`) from an explicit initial field
`f0`, and stores `f0`, the masks the reference built, the boundary parameters in lattice
units, and the reference's result after `n` steps (plus observables where noted).  The
fixtures pin `oracle/lbm_oracle.py` (tests/test_oracle_golden.py) and are compared with the
CUDA path directly (tests/test_gpu_parity.py).  The reference has no golden vectors of its
own for this path, so these are "outputs of the reference itself run
here".
"""
import json
import os
import sys
from typing import List, Optional

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.reference_shim import import_reference  # noqa: E402

lt = import_reference()
torch.set_num_threads(4)

DT = {"f32": torch.float32, "f64": torch.float64}


class GoldenFlow(lt.ExtFlow):
    """A flow whose initial populations are given explicitly."""

    def __init__(self, context, resolution, stencil, f0, re=100.0, ma=0.05):
        self._f0 = f0
        self._boundaries = []
        lt.ExtFlow.__init__(self, context, list(resolution), re, ma, stencil)

    def make_resolution(self, resolution, stencil=None):
        return resolution

    def make_units(self, reynolds_number, mach_number, resolution):
        return lt.UnitConversion(reynolds_number, mach_number,
                                 characteristic_length_lu=resolution[0])

    def initial_pu(self):
        ...

    def initialize(self):
        self.f = self.context.convert_to_tensor(self._f0).clone()

    @property
    def boundaries(self):
        return self._boundaries


class MaskedStreamingBounceBack(lt.BounceBackBoundary):
    """Bounce-back that ALSO exports a per-(q, node) streaming mask, to exercise the
    destination-masked streaming of _simulation.py:171-174 with a non-trivial mask."""

    def __init__(self, mask, nsm):
        super().__init__(mask)
        self._nsm = nsm

    def make_no_streaming_mask(self, shape, context):
        return self._nsm


class MaskedStreamingEquilibrium(lt.EquilibriumBoundaryPU):
    def __init__(self, context, mask, velocity, pressure, nsm):
        super().__init__(context, mask, velocity, pressure)
        self._nsm = nsm

    def make_no_streaming_mask(self, shape, context):
        return self._nsm


def positive_field(rng, q, res, dtype, amp=0.1, stencil=None):
    """w_q * (1 + amp * uniform): positive, near-equilibrium-ish populations."""
    w = np.asarray(stencil.w).reshape([q] + [1] * len(res))
    f = w * (1.0 + amp * rng.random([q, *res]))
    return torch.tensor(f, dtype=dtype)


def tgv_field(stencil_cls, res, dtype, re=100.0, ma=0.05):
    ctx = lt.Context(device="cpu", dtype=dtype, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=list(res), reynolds_number=re,
                                mach_number=ma, stencil=stencil_cls)
    return flow.f.clone(), flow.units.relaxation_parameter_lu


def make_collision(name, flow, tau, tau_minus):
    if name == "none":
        return lt.NoCollision()
    if name == "bgk":
        return lt.BGKCollision(tau=tau)
    if name == "trt":
        return lt.TRTCollision(tau=tau, tau_minus=tau_minus)
    if name == "kbc":
        return lt.KBCCollision()
    raise ValueError(name)


def run_case(name, stencil, dtype, res, f0, collision, n_steps, tau=None, tau_minus=1.0,
             boundaries_spec=(), observables=False, re=100.0, ma=0.05):
    """boundaries_spec: list of dicts {kind, mask, [velocity_pu, pressure_pu], [nsm]}"""
    tdt = DT[dtype]
    ctx = lt.Context(device="cpu", dtype=tdt, use_native=False)
    stencil_cls = getattr(lt, stencil)
    flow = GoldenFlow(ctx, res, stencil_cls, f0.to(tdt), re=re, ma=ma)
    if collision == "kbc":
        tau = flow.units.relaxation_parameter_lu       # kbc_collision.py:97-99
    for spec in boundaries_spec:
        nsm = spec.get("nsm")
        if spec["kind"] == "bounce_back":
            b = (MaskedStreamingBounceBack(spec["mask"], nsm) if nsm is not None
                 else lt.BounceBackBoundary(spec["mask"]))
        else:
            if nsm is not None:
                b = MaskedStreamingEquilibrium(ctx, spec["mask"], spec["velocity_pu"],
                                               spec["pressure_pu"], nsm)
            else:
                b = lt.EquilibriumBoundaryPU(ctx, spec["mask"], spec["velocity_pu"],
                                             spec["pressure_pu"])
        flow.boundaries.append(b)
    sim = lt.Simulation(flow, make_collision(collision, flow, tau, tau_minus), [])

    out = {"f0": f0.to(tdt).numpy().copy()}
    meta = dict(name=name, stencil=stencil, dtype=dtype, res=list(res), collision=collision,
                n_steps=n_steps, tau=tau, tau_minus=tau_minus, boundaries=[])
    if sim.no_collision_mask is not None:
        out["ncm"] = sim.no_collision_mask.numpy().copy()
        out["nsm"] = sim.no_streaming_mask.numpy().copy()
    # boundaries in the order the reference sorted them (index = position)
    for i, b in enumerate(sim.boundaries[1:], start=1):
        if isinstance(b, lt.EquilibriumBoundaryPU):
            rho_lu = flow.units.convert_pressure_pu_to_density_lu(b.pressure)
            u_lu = flow.units.convert_velocity_to_lu(b.velocity)
            out[f"b{i}_rho"] = rho_lu.numpy().copy()
            out[f"b{i}_u"] = u_lu.numpy().copy()
            meta["boundaries"].append(dict(kind="equilibrium", index=i))
        else:
            meta["boundaries"].append(dict(kind="bounce_back", index=i))
    if observables:
        u_conv = flow.units.characteristic_velocity_pu / flow.units.characteristic_velocity_lu
        dx = flow.units.convert_length_to_pu(1.0)
        meta["u_lu_to_pu"] = float(u_conv)
        meta["dx"] = float(dx)
        obs0 = [float(lt.IncompressibleKineticEnergy(flow)()),
                float(lt.MaximumVelocity(flow)()),
                float(lt.Enstrophy(flow)())]
        out["obs0"] = np.array(obs0, dtype=np.float64)
    sim(n_steps)
    out["f_out"] = flow.f.numpy().copy()
    if observables:
        obs1 = [float(lt.IncompressibleKineticEnergy(flow)()),
                float(lt.MaximumVelocity(flow)()),
                float(lt.Enstrophy(flow)())]
        out["obs_out"] = np.array(obs1, dtype=np.float64)
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: f_out sum {out['f_out'].sum():.12g}")


def pulses_2d(dtype):
    # tests/native/test_native_streaming.py:10-21
    f = torch.zeros([9, 16, 16], dtype=DT[dtype])
    for i in range(9):
        f[i, i + 1, i + 1] = float(i + 1)
    return f


def main():
    rng = np.random.default_rng(20261104)

    # ---- pure streaming: bit-exact cases -------------------------------------------------
    run_case("stream_d2q9_pulses_f32", "D2Q9", "f32", [16, 16], pulses_2d("f32"), "none", 1)
    for st, res in (("D2Q9", [12, 20]), ("D3Q19", [6, 10, 8]), ("D3Q27", [5, 7, 9])):
        S = getattr(lt, st)()
        for dt in ("f32", "f64"):
            f0 = torch.tensor(rng.random([S.q, *res]), dtype=DT[dt])
            run_case(f"stream_{st.lower()}_rand_{dt}", st, dt, res, f0, "none", 3)

    # ---- destination-masked streaming + bounce-back, bit-exact ---------------------------
    for st, res in (("D2Q9", [16, 12]), ("D3Q19", [8, 6, 12]), ("D3Q27", [6, 5, 8])):
        S = getattr(lt, st)()
        f0 = torch.tensor(rng.random([S.q, *res]), dtype=torch.float64)
        mask = torch.tensor(rng.random(res) < 0.2)
        nsm = torch.tensor(rng.random([S.q, *res]) < 0.15)
        run_case(f"maskstream_{st.lower()}_f64", st, "f64", res, f0, "none", 4,
                 boundaries_spec=[dict(kind="bounce_back", mask=mask, nsm=nsm)])
    # all-ones streaming mask (tests/native/test_native_no_streaming_mask.py)
    S = lt.D2Q9()
    f0 = torch.tensor(rng.random([9, 16, 16]), dtype=torch.float32)
    run_case("maskstream_allones_d2q9_f32", "D2Q9", "f32", [16, 16], f0, "none", 5,
             boundaries_spec=[dict(kind="bounce_back", mask=torch.zeros([16, 16], dtype=torch.bool),
                                   nsm=torch.ones([9, 16, 16], dtype=torch.bool))])

    # ---- collisions on TGV initial fields (BASELINE configs, reduced) --------------------
    for st, res, colls in (("D2Q9", [32, 32], ("bgk", "trt", "kbc")),
                           ("D3Q19", [16, 16, 16], ("bgk", "trt")),
                           ("D3Q27", [12, 12, 12], ("bgk", "trt", "kbc"))):
        for dt in ("f32", "f64"):
            f0, tau = tgv_field(getattr(lt, st), res, DT[dt], re=100.0, ma=0.05)
            for c in colls:
                run_case(f"tgv_{st.lower()}_{c}_{dt}", st, dt, res, f0, c, 10, tau=tau,
                         tau_minus=0.8 if c == "trt" else 1.0, observables=(c == "bgk"))

    # ---- collisions on random positive fields (single step, far from equilibrium) --------
    for st, res, colls in (("D2Q9", [10, 14], ("bgk", "trt", "kbc")),
                           ("D3Q19", [6, 5, 7], ("bgk", "trt")),
                           ("D3Q27", [5, 6, 7], ("bgk", "trt", "kbc"))):
        S = getattr(lt, st)()
        for dt in ("f32", "f64"):
            f0 = positive_field(rng, S.q, res, DT[dt], amp=0.1, stencil=S)
            for c in colls:
                run_case(f"rand_{st.lower()}_{c}_{dt}", st, dt, res, f0, c, 2, tau=0.7,
                         tau_minus=1.1 if c == "trt" else 1.0)

    # ---- obstacle-like: bounce-back + two equilibrium boundaries --------------------------
    # 2-D (BASELINE config 3, reduced): inlet column x=0, outlet column x=nx-1, cylinder
    for dt in ("f32", "f64"):
        res = [48, 24]
        S = lt.D2Q9()
        ctx = lt.Context(device="cpu", dtype=DT[dt], use_native=False)
        ob = lt.Obstacle(ctx, res, reynolds_number=300, mach_number=0.05, domain_length_x=1.0,
                         stencil=lt.D2Q9)
        x, y = ob.grid
        cyl = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.1 * y.max()) ** 2
        ob.mask = cyl
        ob.initialize()
        f0 = ob.f.clone()
        inlet = torch.zeros(res, dtype=torch.bool); inlet[0, :] = True
        outlet = torch.zeros(res, dtype=torch.bool); outlet[-1, :] = True
        tau = ob.units.relaxation_parameter_lu
        # GoldenFlow uses UnitConversion(re, ma, characteristic_length_lu=res[0]); pass PU values directly
        spec = [dict(kind="equilibrium", mask=inlet, velocity_pu=[1.0, 0.0], pressure_pu=0.0),
                dict(kind="equilibrium", mask=outlet,
                     velocity_pu=np.stack([np.linspace(0.9, 1.1, res[1])[None, :],
                                           np.zeros([1, res[1]])]),          # [2,1,ny]
                     pressure_pu=np.linspace(-0.01, 0.01, res[1])[None, None, :]),  # [1,1,ny]
                dict(kind="bounce_back", mask=cyl)]
        run_case(f"obstacle_d2q9_bgk_{dt}", "D2Q9", dt, res, f0, "bgk", 20, tau=max(tau, 0.55),
                 boundaries_spec=spec, re=300.0, ma=0.05)

    # 3-D with D3Q19 / D3Q27, full-shape velocity fields, KBC on D3Q27
    for st, c in (("D3Q19", "bgk"), ("D3Q27", "kbc")):
        res = [12, 10, 8]
        S = getattr(lt, st)()
        f0 = positive_field(rng, S.q, res, torch.float64, amp=0.02, stencil=S)
        solid = torch.tensor(rng.random(res) < 0.1)
        inlet = torch.zeros(res, dtype=torch.bool); inlet[0] = True
        top = torch.zeros(res, dtype=torch.bool); top[:, :, -1] = True    # overlaps inlet: later index wins
        spec = [dict(kind="bounce_back", mask=solid),
                dict(kind="equilibrium", mask=inlet,
                     velocity_pu=0.5 * rng.random([3, *res]), pressure_pu=0.01 * rng.random([1, *res])),
                dict(kind="equilibrium", mask=top, velocity_pu=[0.3, 0.0, 0.1], pressure_pu=0.002)]
        run_case(f"box_{st.lower()}_{c}_f64", st, "f64", res, f0, c, 5, tau=0.8,
                 boundaries_spec=spec, re=50.0, ma=0.05)


if __name__ == "__main__":
    main()
