"""Small invocations of every kernel family, for compute-sanitizer (memcheck / racecheck on the
small parity cases; the design rests on "every output element is written exactly once", shared-memory staging,
mbarrier-driven bulk copies, and cooperative multi-step launches):

    compute-sanitizer --tool memcheck  python scripts/sanitize_cases.py
    compute-sanitizer --tool racecheck python scripts/sanitize_cases.py      (shared-memory hazards)
    compute-sanitizer --tool synccheck python scripts/sanitize_cases.py

Each case also checks its result (against the plain path / the other kernel variant), so a sanitizer run is a
parity run as well.  Prints one line per case and `sanitize_cases OK`."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import lettuce_b200 as lt  # noqa: E402
import gpu_util as gu  # noqa: E402

nat = lt.native
L = nat.lib()
st = torch.cuda.current_stream().cuda_stream


def tgv(res, stencil, dtype, coll="bgk", flags=0, reporters=False, steps=4):
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=stencil)
    tau = flow.units.relaxation_parameter_lu
    c = {"bgk": lambda: lt.BGKCollision(tau), "kbc": lambda: lt.KBCCollision(), "trt": lambda: lt.TRTCollision(tau)}[coll]()
    reps = [lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None),
            lt.ObservableReporter(lt.Enstrophy(flow), interval=2, out=None)] if reporters else []
    sim = lt.Simulation(flow, c, reps)
    sim.kernel_flags = flags
    sim(steps)
    torch.cuda.synchronize()
    return flow.f.clone(), [r[2] for r in reps[0].out] if reporters else None


def say(name, ok):
    print(f"sanitize case {name}: {'ok' if ok else 'MISMATCH'}", flush=True)
    assert ok, name


# 1. classic kernel (one launch per step) == cooperative multi-step kernel, with fused reductions + enstrophy
a, ea = tgv([32, 24, 32], lt.D3Q19, torch.float32, flags=nat.LB_KF_NO_LOOP | nat.LB_KF_NO_TMA, reporters=True)
b, eb = tgv([32, 24, 32], lt.D3Q19, torch.float32, flags=nat.LB_KF_NO_TMA, reporters=True)
say("classic per-step launches vs cooperative loop kernel (OBS, enstrophy)", torch.equal(a, b) and np.allclose(ea, eb, rtol=1e-6))

# 2. persistent TMA-prefetch kernel (mbarrier + bulk copies + shared memory) == classic kernel
for stencil, dtype, coll, res in ((lt.D3Q27, torch.float32, "kbc", [64, 64, 128]), (lt.D3Q19, torch.float64, "bgk", [64, 64, 64])):
    a, ea = tgv(res, stencil, dtype, coll, flags=nat.LB_KF_NO_LOOP | nat.LB_KF_NO_TMA, reporters=True, steps=3)
    b, eb = tgv(res, stencil, dtype, coll, flags=nat.LB_KF_NO_LOOP | nat.LB_KF_FORCE_TMA, reporters=True, steps=3)
    say(f"TMA kernel vs classic kernel {stencil.__name__} {coll} {dtype}", torch.equal(a, b) and np.allclose(ea, eb, rtol=1e-6))

# 3. masked streaming + outlet + bounce-back (stock obstacle), 2-D and 3-D, batched vs per-step seam
for res, stencil in (([48, 32], lt.D2Q9), ([32, 16, 24], lt.D3Q27)):
    outs = []
    for seam in (False, True):
        ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
        flow = lt.Obstacle(ctx, res, reynolds_number=100, mach_number=0.05, domain_length_x=1, stencil=stencil)
        g = flow.grid
        flow.mask = sum((c - 0.4 * c.max()) ** 2 for c in g) < (0.15 * g[1].max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(tau=max(float(flow.units.relaxation_parameter_lu), 0.6)), [])
        if seam:
            for _ in range(6):
                sim._collide_and_stream(sim)
        else:
            sim(6)
        torch.cuda.synchronize()
        outs.append(flow.f.clone())
    say(f"stock obstacle {res} masked streaming, batched vs seam", torch.equal(outs[0], outs[1]) and bool(torch.isfinite(outs[0]).all()))

# 4. z-slab halo paths on one GPU with the slab as its own neighbour: LL (tagged words, staged in shared memory,
#    polled) and REMOTE (+ handshake kernel), against the periodic run
for stencil, dtype, res in ((lt.D3Q19, torch.float32, [16, 12, 64]), (lt.D3Q27, torch.float64, [8, 6, 16]), (lt.D3Q19, torch.float32, [6, 5, 7])):
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=stencil)
    f0 = flow.f.clone()
    sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
    desc = sim._desc()
    sim(6)
    want = flow.f.clone()
    for mode in ("ll", "remote"):
        A, B = f0.clone(), torch.empty_like(f0)
        flags = torch.zeros([8], dtype=torch.int32, device="cuda")
        h = nat.LbHalo()
        lines, nf = int(np.prod(res[:-1])), L.lb_zface_count(desc.stencil)
        if mode == "ll":
            nb = L.lb_halo_ll_bytes(desc)
            lo = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
            hi = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
            h.mode = nat.LB_HALO_LL
        else:
            lo = [torch.zeros([nf, lines], dtype=dtype, device="cuda") for _ in range(2)]
            hi = [torch.zeros([nf, lines], dtype=dtype, device="cuda") for _ in range(2)]
            h.mode = nat.LB_HALO_REMOTE
            h.signal_up, h.signal_down = flags.data_ptr(), flags.data_ptr() + 4
        for k in range(len(lo)):
            h.recv_lo[k], h.recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
            h.peer_up_recv_lo[k], h.peer_down_recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
        h.wait_slots = flags.data_ptr()
        fl = nat.LbFields()
        fl.f, fl.f_next = A.data_ptr(), B.data_ptr()
        fl.halo = C.pointer(h)
        nat.check(L.lb_step(desc, fl, 6, st))
        torch.cuda.synchronize()
        say(f"halo {mode} self-neighbour {stencil.__name__} {dtype} {res}", torch.equal(A, want) and int(flags[2].item()) == 0)

# 5. the small kernels: initialise (device), moments, observables, output fields, mask packing
ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
flow = lt.TaylorGreenVortex(ctx, resolution=[20, 18, 16], reynolds_number=100, mach_number=0.05, stencil=lt.D3Q19)
rho, u = flow.rho_u()
fields = flow.output_fields_pu()
e = float(lt.IncompressibleKineticEnergy(flow)(flow.f))
w = float(lt.Enstrophy(flow)(flow.f))
torch.cuda.synchronize()
say("initialize / moments / observables / enstrophy / output fields", bool(torch.isfinite(fields).all()) and e > 0 and w > 0
    and abs(float(rho.mean()) - 1.0) < 1e-3)
print("sanitize_cases OK")
