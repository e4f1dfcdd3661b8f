"""The kernels either side of the step (SURVEY 8f rank 4 and the reporter tail) on the GPU box:
    python scripts/bench_output.py [res] [--no-vtk-runs]

One JSON line per measurement, D3Q19 fp32 Taylor-Green lattice of res^3 nodes (default 256):
  * every small kernel alone -- output_fields_kernel (lb_output_fields), moments_kernel (lb_moments, rho + u),
    moments_kernel<OBS> (lb_observables), enstrophy_kernel (lb_enstrophy), initialize_kernel (lb_initialize):
    CUDA events around 20 back-to-back launches into preallocated buffers, algorithmic bytes per node as stated,
    fraction of the measured HBM peak (MEASURED_PEAKS.json);
  * a whole VTKReporter call inline (device + d2h + file write) and its parts;
  * MLUPS of simulation(1000) without a reporter and with VTKReporter(interval = 1000 / 100 / 10), background
    pipeline (kernel on the stepping stream, D2H on a second stream, file written by a thread) and inline.
Not a bench.py line; side measurements for DESIGN.md sections 3.5 / 3.6."""
import ctypes as C
import json
import os
import sys
import tempfile
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native
args = [a for a in sys.argv[1:] if not a.startswith("--")]
res = int(args[0]) if args else 256
PEAK = 6548.8
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
flow = lt.TaylorGreenVortex(ctx, resolution=[res] * 3, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
n, d, q, esz = res ** 3, 3, 19, 4
L = nat.lib()
st = torch.cuda.current_stream().cuda_stream


def cuda_us(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def line(kernel, values_per_node, us, **more):
    gbs = values_per_node * esz * n / us / 1e3
    out = {"kernel": kernel, "res": res, "us": round(us, 1), "values_per_node": values_per_node,
           "achieved_gbs": round(gbs, 1), "frac_of_measured_peak": round(gbs / PEAK, 3)}
    out.update(more)
    print(json.dumps(out), flush=True)


desc = nat.make_desc("D3Q19", torch.float32, [res] * 3, tau=float(flow.units.relaxation_parameter_lu))
f = flow.f
fields = torch.empty([1 + d, n], dtype=f.dtype, device="cuda")
rho = torch.empty([1, res, res, res], dtype=f.dtype, device="cuda")
u = torch.empty([d, res, res, res], dtype=f.dtype, device="cuda")
obs = torch.zeros(nat.LB_OBS_COUNT, dtype=torch.float64, device="cuda")
scratch = torch.empty(L.lb_obs_scratch_bytes(desc), dtype=torch.uint8, device="cuda")
fnew = torch.empty_like(f)

line("output_fields_kernel (lb_output_fields: p, u in file order)", q + 1 + d,
     cuda_us(lambda: flow.output_fields_pu(out=fields)))
line("moments_kernel (lb_moments: rho + u)", q + 1 + d,
     cuda_us(lambda: nat.check(L.lb_moments(desc, f.data_ptr(), rho.data_ptr(), u.data_ptr(), st))))
line("moments_kernel<OBS> (lb_observables: energy, max|u|^2, mass)", q,
     cuda_us(lambda: nat.check(L.lb_observables(desc, f.data_ptr(), 0, None, scratch.data_ptr(), obs.data_ptr(), st))))
line("enstrophy_kernel (lb_enstrophy: 6th-order curl of u, reduction)", d,
     cuda_us(lambda: nat.check(L.lb_enstrophy(desc, u.data_ptr(), scratch.data_ptr(), obs.data_ptr(), st))),
     note="reads d values per node once from HBM; the 18 stencil taps per node come from L1/L2")
line("initialize_kernel (lb_initialize: feq - fneq from rho, u)", q + 1 + d,
     cuda_us(lambda: nat.check(L.lb_initialize(desc, rho.data_ptr(), u.data_ptr(), 0, 1, fnew.data_ptr(), st))))
del fnew, rho, u

# ---- one inline report and its parts -----------------------------------------------------------------
stage = torch.empty(fields.shape, dtype=fields.dtype, pin_memory=True)
d2h_us = cuda_us(lambda: stage.copy_(fields, non_blocking=True), 5)
host_f = torch.empty(f.shape, dtype=f.dtype, pin_memory=True)
full_us = cuda_us(lambda: host_f.copy_(f, non_blocking=True), 3)
del host_f
sim_ns = types.SimpleNamespace(flow=flow, units=flow.units)
with tempfile.TemporaryDirectory() as tmp:
    rep = lt.VTKReporter(interval=1, filename_base=os.path.join(tmp, "out"), asynchronous=False)
    rep(sim_ns)
    t0 = time.perf_counter()
    for i in range(3):
        flow.i = i
        rep(sim_ns)
    torch.cuda.synchronize()
    report_ms = (time.perf_counter() - t0) / 3 * 1e3
    size = os.path.getsize(os.path.join(tmp, "out_00000000.vtr"))
flow.i = 0
print(json.dumps({"what": "one inline VTK report, D3Q19 fp32", "res": res, "fields_d2h_ms": round(d2h_us / 1e3, 2),
                  "report_ms": round(report_ms, 1), "file_bytes": size, "fields_bytes": int(stage.numel() * 4),
                  "full_f_d2h_ms": round(full_us / 1e3, 2), "full_f_bytes": int(f.numel() * 4),
                  "note": "full_f = what the reference's reporter path moves to the host instead (flow.f, q values per node)"}),
      flush=True)
del stage, fields

# ---- reports inside a run ---------------------------------------------------------------------------------
if "--no-vtk-runs" not in sys.argv:
    T = 1000

    def run(interval, asynchronous, T=T):
        fl = lt.TaylorGreenVortex(ctx, resolution=[res] * 3, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
        with tempfile.TemporaryDirectory() as tmp:
            reps = [] if interval is None else [lt.VTKReporter(interval=interval, filename_base=os.path.join(tmp, "o"),
                                                               asynchronous=asynchronous)]
            sim = lt.Simulation(fl, lt.BGKCollision(tau=fl.units.relaxation_parameter_lu), reps)
            sim(50)                                    # warm-up (one report at i = 0)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            sim(T)                                     # flushes its reports before it returns
            dt = time.perf_counter() - t0
            files = len(os.listdir(tmp))
            for r in reps:
                r.close()
        return n * T / 1e6 / dt, files

    base, _ = run(None, True)
    print(json.dumps({"what": f"simulation({T}) without reporter", "res": res, "mlups": round(base, 1)}), flush=True)
    for interval in (1000, 100, 10):
        for asynchronous in (True, False):
            if interval == 10 and not asynchronous and res >= 256:
                continue                               # 100 x 96 ms of inline reports: not informative
            steps = 200 if interval == 10 else T          # 268 MB per file at 256^3: keep the temporary directory small
            m, files = run(interval, asynchronous, steps)
            print(json.dumps({"what": f"simulation({steps}) with VTKReporter(interval={interval})",
                              "pipeline": "background (second stream + writer thread)" if asynchronous else "inline",
                              "res": res, "mlups": round(m, 1), "loss_pct": round(100 * (1 - m / base), 2),
                              "files_written": files}), flush=True)
