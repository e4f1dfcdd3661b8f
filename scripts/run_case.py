"""Run one configuration for a few steps (profiling target): python scripts/run_case.py D3Q27 kbc f32 256 20"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

stencil, coll, dt, res, steps = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
ctx = lt.Context(device="cuda", dtype={"f32": torch.float32, "f64": torch.float64}[dt], use_native=True)
S = getattr(lt, stencil)
flow = lt.TaylorGreenVortex(ctx, resolution=[res] * S().d, reynolds_number=1600, mach_number=0.1, stencil=S)
tau = flow.units.relaxation_parameter_lu
c = {"bgk": lambda: lt.BGKCollision(tau), "trt": lambda: lt.TRTCollision(tau), "kbc": lambda: lt.KBCCollision(),
     "none": lambda: lt.NoCollision()}[coll]()
sim = lt.Simulation(flow, c, [])
print(sim(steps), "MLUPS (incl. launch)")
