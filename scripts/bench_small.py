"""BASELINE config 1 (TGV2D D2Q9 BGK fp64 256^2, 1000 steps per call) and its fp32 / 512^2 relatives with the three
multi-step strategies of the library (lb_desc.kernel_flags): one launch per step (PDL), the cooperative loop
kernel (grid barrier per step) and the tile-resident kernel (state in shared memory, one grid barrier per h steps;
LB_RES_H=<h> forces h).  (Round 2 also measured a barrier-free "chain" kernel here -- populations handed from
step to step as tagged words through L2; it was 1.3x - 2.9x SLOWER than the loop kernel and was removed:
profiles/r2_chain_kernel_experiment_*.jsonl)
    python scripts/bench_small.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native


def run(res, dtype, flags, steps=1000):
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
    sim.kernel_flags = flags
    sim(steps)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    sim(steps)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / steps, flow.f.clone()


CASES = (([256, 256], torch.float64), ([256, 256], torch.float32), ([512, 512], torch.float64), ([128, 128], torch.float64),
         ([512, 512], torch.float32), ([64, 64], torch.float64))
if "--cfg1" in sys.argv:
    CASES = CASES[:2]
for res, dtype in CASES:
    out = {"LB_RES_H": os.environ.get("LB_RES_H"), "case": f"TGV2D D2Q9 BGK {str(dtype)[6:]} {res[0]}x{res[1]} x1000 steps"}
    ref = None
    for name, flags in (("launch_per_step", nat.LB_KF_NO_LOOP | nat.LB_KF_NO_TMA), ("loop_kernel", nat.LB_KF_NO_RESIDENT),
                        ("resident_kernel", 0)):
        us, f = run(res, dtype, flags)
        out[name + "_us_per_step"] = round(us, 3)
        if ref is None:
            ref = f
        out[name + "_bit_exact"] = bool(torch.equal(f, ref))
    print(json.dumps(out), flush=True)
