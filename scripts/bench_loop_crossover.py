"""Where does the cooperative multi-step kernel (grid barrier per step) stop beating one launch per step (programmatic
dependent launch)?  Device resident, CUDA events; LB_KF_NO_RESIDENT so that the 2-D cases measure the loop kernel.
    python scripts/bench_loop_crossover.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native


def run(stencil, res, dtype, flags, steps):
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=100, mach_number=0.05, stencil=stencil)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
    sim.kernel_flags = flags
    sim(steps)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    sim(steps)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / steps


f32, f64 = torch.float32, torch.float64
cases = [(lt.D2Q9, [256, 256], f64), (lt.D2Q9, [320, 320], f64), (lt.D2Q9, [384, 384], f64), (lt.D2Q9, [512, 512], f64),
         (lt.D2Q9, [512, 512], f32), (lt.D2Q9, [640, 640], f32), (lt.D2Q9, [768, 768], f32), (lt.D2Q9, [1024, 1024], f32),
         (lt.D3Q19, [32, 32, 32], f32), (lt.D3Q19, [48, 48, 48], f32), (lt.D3Q19, [64, 64, 64], f32),
         (lt.D3Q19, [80, 80, 80], f32), (lt.D3Q19, [96, 96, 96], f32), (lt.D3Q19, [48, 48, 48], f64),
         (lt.D3Q19, [64, 64, 64], f64), (lt.D3Q27, [64, 64, 64], f32), (lt.D3Q27, [80, 80, 80], f32)]
for stencil, res, dtype in cases:
    q = stencil().q
    n = 1
    for r in res:
        n *= r
    mb = n * q * (4 if dtype == f32 else 8) / 2 ** 20
    per_step = run(stencil, res, dtype, nat.LB_KF_NO_LOOP | nat.LB_KF_NO_TMA, 500)
    loop = run(stencil, res, dtype, nat.LB_KF_NO_RESIDENT | nat.LB_KF_NO_TMA, 500)
    print(json.dumps({"case": f"{stencil.__name__} {'x'.join(map(str, res))} {str(dtype)[6:]}", "populations_MB": round(mb, 1),
                      "launch_per_step_us": round(per_step, 2), "loop_kernel_us": round(loop, 2),
                      "loop_wins": loop < per_step}), flush=True)
