"""What does a fused every-step energy reporter cost on the big lattices?  Device resident, CUDA events.
    python scripts/bench_obs_overhead.py"""
import gc
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402
from scripts.bench_configs import timed, PEAK  # noqa: E402


def case(name, res, stencil, dtype, coll, steps):
    out = {"case": name}
    for label, interval in (("no_reporter", None), ("energy_every_step", 1), ("energy_and_max_velocity_every_step", -1)):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=stencil)
        tau = flow.units.relaxation_parameter_lu
        c = {"bgk": lambda: lt.BGKCollision(tau), "trt": lambda: lt.TRTCollision(tau), "kbc": lambda: lt.KBCCollision()}[coll]()
        reps = []
        if interval:
            reps.append(lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None))
        if interval == -1:
            reps.append(lt.ObservableReporter(lt.MaximumVelocity(flow), interval=1, out=None))
        sim = lt.Simulation(flow, c, reps)
        t = timed(sim, steps)
        n, q, esz = flow.f[0].numel(), flow.stencil.q, flow.f.element_size()
        out[label] = {"us_per_step": round(t / steps * 1e6, 2), "frac": round(n * steps * 2 * q * esz / t / 1e9 / PEAK, 4)}
        del sim, flow, reps
        gc.collect()
        torch.cuda.empty_cache()
    print(json.dumps(out), flush=True)


f32, f64 = torch.float32, torch.float64
case("TGV3D D3Q19 BGK f32 256^3 (config 2)", 256, lt.D3Q19, f32, "bgk", 100)
case("TGV3D D3Q19 TRT f32 256^3", 256, lt.D3Q19, f32, "trt", 100)
case("TGV3D D3Q19 BGK f64 256^3", 256, lt.D3Q19, f64, "bgk", 50)
case("TGV3D D3Q27 BGK f32 256^3", 256, lt.D3Q27, f32, "bgk", 100)
case("TGV3D D3Q27 KBC f32 256^3", 256, lt.D3Q27, f32, "kbc", 100)
case("TGV2D D2Q9 BGK f32 2048x1024", [2048, 1024], lt.D2Q9, f32, "bgk", 300)
