"""Does the result depend on the nodes-per-thread width the launcher picks?  Runs the same case in two processes,
one with the vector kernels and one forced to one node per thread (LB_VEC=1), and compares bit for bit.
    python scripts/check_vec_width.py"""
import os
import subprocess
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, ROOT)
    import lettuce_b200 as lt
    stencil, dtype, coll, steps, out = sys.argv[2], getattr(torch, sys.argv[3]), sys.argv[4], int(sys.argv[5]), sys.argv[6]
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=[16, 12, 32] if stencil != "D2Q9" else [48, 64], reynolds_number=400,
                                mach_number=0.1, stencil=getattr(lt, stencil))
    tau = flow.units.relaxation_parameter_lu
    c = {"bgk": lambda: lt.BGKCollision(tau), "none": lambda: lt.NoCollision(), "kbc": lambda: lt.KBCCollision(),
         "trt": lambda: lt.TRTCollision(tau)}[coll]()
    sim = lt.Simulation(flow, c, [])
    sim.kernel_flags = lt.native.LB_KF_NO_LOOP | lt.native.LB_KF_NO_TMA
    sim(steps)
    torch.save(flow.f.cpu(), out)
    sys.exit(0)

ok = True
with tempfile.TemporaryDirectory() as tmp:
    for stencil, dtype, coll in (("D3Q19", "float32", "none"), ("D3Q19", "float32", "bgk"), ("D3Q19", "float64", "bgk"),
                                 ("D3Q27", "float32", "kbc"), ("D2Q9", "float32", "trt"), ("D3Q27", "float32", "bgk")):
        for steps in (1, 10):
            outs = []
            for vec in ("", "1"):
                env = dict(os.environ)
                if vec:
                    env["LB_VEC"] = vec
                path = os.path.join(tmp, f"o{vec}.pt")
                subprocess.run([sys.executable, __file__, "child", stencil, dtype, coll, str(steps), path], env=env, check=True,
                               stdout=subprocess.DEVNULL)
                outs.append(torch.load(path))
            a, b = outs
            diff = (a - b).abs()
            same = bool(torch.equal(a, b))
            ok = ok and same
            print(f"vector kernels vs one node per thread: {stencil} {dtype} {coll} {steps} steps: bit-exact={same} "
                  f"max diff {float(diff.max()):.3e} ({int((diff > 0).sum())} of {a.numel()} elements differ)", flush=True)
print("check_vec_width", "OK" if ok else "MISMATCH")
