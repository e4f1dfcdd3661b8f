"""MLUPS and HBM-roofline fraction of every BASELINE.json config that fits one GPU, plus a few
kernel variants.  Device-resident timing with CUDA events (10 warm-up steps, then `steps`).

    python scripts/bench_configs.py [--quick]
Prints one JSON line per case."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

PEAK = 6553.0
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
quick = "--quick" in sys.argv
only = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--only=")]


def timed(sim, steps, warm=10):
    sim(warm)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sim(steps)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 1e3


def wanted(name):
    return not only or any(o in name for o in only)


def report(name, flow, seconds, steps, extra=None):
    n = flow.f[0].numel()
    q = flow.stencil.q
    esz = flow.f.element_size()
    mlups = n * steps / 1e6 / seconds
    gbs = mlups * 1e6 * 2 * q * esz / 1e9
    line = {"case": name, "mlups": round(mlups, 1), "us_per_step": round(seconds / steps * 1e6, 2),
            "bytes_per_update": 2 * q * esz, "achieved_gbs": round(gbs, 1), "frac_of_measured_peak": round(gbs / PEAK, 4)}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def tgv_case(name, res, stencil, dtype, collision, steps, **kw):
    import gc
    if not wanted(name):
        return
    flow, sim, t = tgv(res, stencil, dtype, collision, steps=steps, **kw)
    report(name, flow, t, steps)
    del flow, sim
    gc.collect()
    torch.cuda.empty_cache()


def tgv(res, stencil, dtype, collision, re=1600, ma=0.1, reporters=None, steps=100):
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=re, mach_number=ma, stencil=stencil)
    tau = flow.units.relaxation_parameter_lu
    coll = {"bgk": lambda: lt.BGKCollision(tau), "trt": lambda: lt.TRTCollision(tau), "kbc": lambda: lt.KBCCollision(),
            "none": lambda: lt.NoCollision(), "regularized": lambda: lt.RegularizedCollision(),
            "smagorinsky": lambda: lt.SmagorinskyCollision(tau),
            "bgk_guo": lambda: lt.BGKCollision(tau, force=lt.Guo(flow, tau, [1e-6] * flow.stencil.d))}[collision]()
    reps = reporters(flow) if reporters else []
    sim = lt.Simulation(flow, coll, reps)
    return flow, sim, timed(sim, steps)


class EquilibriumOutletObstacle(lt.Obstacle):
    """BASELINE config 3: "bounce-back + equilibrium inlet/outlet" -- equilibrium boundaries on the
    x = 0 and x = nx-1 columns (all node-local), bounce-back on the obstacle mask"""

    @property
    def boundaries(self):
        x = self.grid[0]
        u_in = self.units.characteristic_velocity_pu * self._unit_vector()
        return [lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x - x.max()) < 1e-6, velocity=u_in),
                lt.BounceBackBoundary(self.mask)]


def main():
    import gc
    S = 20 if quick else 100
    # config 1: TGV2D D2Q9 BGK fp64 256^2, 1000 steps (launch-latency bound)
    tgv_case("cfg1 TGV2D D2Q9 BGK f64 256^2 x1000", 256, lt.D2Q9, torch.float64, "bgk", 1000, re=100, ma=0.05)
    # config 2 and its neighbours
    for coll in ("bgk", "trt", "none"):
        tgv_case(f"cfg2 TGV3D D3Q19 {coll} f32 256^3", 256, lt.D3Q19, torch.float32, coll, S)
    for coll in ("regularized", "smagorinsky"):
        tgv_case(f"rank3 TGV3D D3Q19 {coll} f32 256^3", 256, lt.D3Q19, torch.float32, coll, S)
    tgv_case("rank3 TGV3D D3Q27 smagorinsky f32 256^3", 256, lt.D3Q27, torch.float32, "smagorinsky", S)
    tgv_case("rank3 TGV3D D3Q27 regularized f32 256^3", 256, lt.D3Q27, torch.float32, "regularized", S)
    tgv_case("rank3 TGV3D D3Q19 bgk+guo f32 256^3", 256, lt.D3Q19, torch.float32, "bgk_guo", S)
    tgv_case("TGV3D D3Q19 bgk f64 256^3", 256, lt.D3Q19, torch.float64, "bgk", S)
    for coll in ("bgk", "trt", "kbc"):
        tgv_case(f"TGV3D D3Q27 {coll} f32 256^3", 256, lt.D3Q27, torch.float32, coll, S)
    tgv_case("TGV3D D3Q27 kbc f64 256^3", 256, lt.D3Q27, torch.float64, "kbc", S)
    # config 3: obstacle 2048x1024 D2Q9 BGK, bounce-back cylinder + equilibrium inlet/outlet
    for dtype, nm in ((torch.float32, "f32"), (torch.float64, "f64")):
        name = f"cfg3 Obstacle D2Q9 BGK {nm} 2048x1024 (bounce-back + 2 equilibrium)"
        if not wanted(name):
            continue
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = EquilibriumOutletObstacle(ctx, [2048, 1024], reynolds_number=300, mach_number=0.01, domain_length_x=1)
        x, y = flow.grid
        flow.mask = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.05 * y.max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
        t = timed(sim, 5 * S)
        report(name, flow, t, 5 * S)
        del flow, sim
        gc.collect()
    # the stock lt.Obstacle boundary list: AntiBounceBackOutlet on +x (per-population streaming mask ->
    # the masked-streaming kernel)
    name = "stock lt.Obstacle D2Q9 BGK f32 2048x1024 (equilibrium inlet, anti-bounce-back outlet, bounce-back)"
    if wanted(name):
        ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
        flow = lt.Obstacle(ctx, [2048, 1024], reynolds_number=300, mach_number=0.01, domain_length_x=1)
        x, y = flow.grid
        flow.mask = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.05 * y.max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
        t = timed(sim, 5 * S)
        report(name, flow, t, 5 * S)
        del flow, sim
        gc.collect()
    # the same boundary list in 3-D: a sphere in a 384 x 192 x 192 channel, D3Q19 and D3Q27
    for st, nm in ((lt.D3Q19, "D3Q19"), (lt.D3Q27, "D3Q27")):
        name = f"stock lt.Obstacle {nm} BGK f32 384x192x192 (equilibrium inlet, anti-bounce-back outlet, bounce-back)"
        if not wanted(name):
            continue
        ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
        flow = lt.Obstacle(ctx, [384, 192, 192], reynolds_number=300, mach_number=0.01, domain_length_x=2, stencil=st)
        x, y, z = flow.grid
        flow.mask = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2 + (z - 0.5 * z.max()) ** 2) < (0.1 * y.max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
        t = timed(sim, S)
        report(name, flow, t, S)
        del flow, sim
        gc.collect()
    tgv_case("TGV2D D2Q9 BGK f32 2048x1024 (no masks)", [2048, 1024], lt.D2Q9, torch.float32, "bgk", 5 * S)
    tgv_case("TGV2D D2Q9 KBC f32 2048x1024", [2048, 1024], lt.D2Q9, torch.float32, "kbc", 5 * S)
    # config 4: TGV3D D3Q27 KBC f32 512^3 with the fused energy reporter, interval 1 / 10 / 100 / none
    if not quick:
        for interval in (1, 10, 100, None):
            reps = (lambda fl, k=interval: [lt.ObservableReporter(lt.IncompressibleKineticEnergy(fl), interval=k, out=None)]) \
                if interval else None
            tgv_case(f"cfg4 TGV3D D3Q27 KBC f32 512^3 energy reporter interval={interval}", 512, lt.D3Q27,
                     torch.float32, "kbc", 100, reporters=reps)
        tgv_case("TGV3D D3Q19 bgk f32 512^3", 512, lt.D3Q19, torch.float32, "bgk", 50)


if __name__ == "__main__":
    main()
