"""A/B of the persistent per-thread prefetch (cp.async) step kernel against the classic one (lb_desc.kernel_flags),
device-resident, CUDA events, one JSON line per case:  python scripts/bench_prefetch.py [--quick]"""
import gc
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native
PEAK = 6548.8
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
quick = "--quick" in sys.argv


def timed(sim, steps, warm=10):
    sim(warm)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    sim(steps)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / 1e3


def case(name, res, stencil, dtype, coll, steps, interval=None):
    out = {"case": name}
    variants = [("classic", nat.LB_KF_NO_TMA | nat.LB_KF_NO_PREFETCH), ("prefetch", nat.LB_KF_FORCE_PREFETCH)]
    if stencil is lt.D2Q9:
        variants.append(("tma", nat.LB_KF_FORCE_TMA | nat.LB_KF_NO_PREFETCH))
    for label, flag in variants:
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=stencil)
        tau = flow.units.relaxation_parameter_lu
        c = {"bgk": lambda: lt.BGKCollision(tau), "trt": lambda: lt.TRTCollision(tau), "kbc": lambda: lt.KBCCollision(),
             "smagorinsky": lambda: lt.SmagorinskyCollision(tau), "regularized": lambda: lt.RegularizedCollision()}[coll]()
        reps = [lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=interval, out=None)] if interval else []
        sim = lt.Simulation(flow, c, reps)
        sim.kernel_flags = flag
        t = timed(sim, steps)
        n, q, esz = flow.f[0].numel(), flow.stencil.q, flow.f.element_size()
        mlups = n * steps / 1e6 / t
        out[label] = {"mlups": round(mlups, 1), "us_per_step": round(t / steps * 1e6, 2),
                      "frac": round(mlups * 1e6 * 2 * q * esz / 1e9 / PEAK, 4)}
        del sim, flow, reps
        gc.collect()
        torch.cuda.empty_cache()
    out["prefetch_speedup"] = round(out["prefetch"]["mlups"] / out["classic"]["mlups"], 4)
    if "tma" in out:
        out["tma_speedup"] = round(out["tma"]["mlups"] / out["classic"]["mlups"], 4)
    print(json.dumps(out), flush=True)


f32, f64 = torch.float32, torch.float64
S = 30 if quick else 100
if "--cfg4" in sys.argv:
    case("cfg4 TGV3D D3Q27 KBC f32 512^3", 512, lt.D3Q27, f32, "kbc", 30)
    case("cfg4 TGV3D D3Q27 KBC f32 512^3 energy every step", 512, lt.D3Q27, f32, "kbc", 30, interval=1)
    sys.exit(0)
case("TGV3D D3Q27 KBC f32 256^3", 256, lt.D3Q27, f32, "kbc", S)
case("TGV3D D3Q27 KBC f32 256^3 energy every step", 256, lt.D3Q27, f32, "kbc", S, interval=1)
case("TGV3D D3Q27 TRT f32 256^3", 256, lt.D3Q27, f32, "trt", S)
case("TGV3D D3Q27 BGK f32 256^3", 256, lt.D3Q27, f32, "bgk", S)
case("TGV3D D3Q27 smagorinsky f32 256^3", 256, lt.D3Q27, f32, "smagorinsky", S)
case("TGV3D D3Q27 regularized f32 256^3", 256, lt.D3Q27, f32, "regularized", S)
case("TGV3D D3Q27 KBC f64 256^3", 256, lt.D3Q27, f64, "kbc", S)
case("TGV3D D3Q27 BGK f64 256^3", 256, lt.D3Q27, f64, "bgk", S)
case("TGV3D D3Q27 TRT f64 256^3", 256, lt.D3Q27, f64, "trt", S)
case("TGV2D D2Q9 BGK f32 2048x1024", [2048, 1024], lt.D2Q9, f32, "bgk", 5 * S)
case("TGV2D D2Q9 KBC f32 2048x1024", [2048, 1024], lt.D2Q9, f32, "kbc", 5 * S)
case("TGV2D D2Q9 BGK f64 2048x1024", [2048, 1024], lt.D2Q9, f64, "bgk", 5 * S)
case("TGV2D D2Q9 KBC f64 2048x1024", [2048, 1024], lt.D2Q9, f64, "kbc", 5 * S)
case("TGV2D D2Q9 BGK f32 4096x4096 (302 MB of populations: beyond L2)", [4096, 4096], lt.D2Q9, f32, "bgk", S)
case("TGV2D D2Q9 KBC f32 4096x4096", [4096, 4096], lt.D2Q9, f32, "kbc", S)
if not quick:
    case("cfg4 TGV3D D3Q27 KBC f32 512^3", 512, lt.D3Q27, f32, "kbc", 50)
    case("cfg4 TGV3D D3Q27 KBC f32 512^3 energy every step", 512, lt.D3Q27, f32, "kbc", 50, interval=1)
    case("cfg4 TGV3D D3Q27 KBC f32 512^3 energy every 10 steps", 512, lt.D3Q27, f32, "kbc", 50, interval=10)
