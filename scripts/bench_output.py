"""Cost of one VTK report on the GPU box: python scripts/bench_output.py [res] [reports]

Times, for a D3Q19 fp32 Taylor-Green lattice of res^3 nodes (default 256),
  device  : lb_output_fields -- one pass over f: moments, PU conversion, reordering into file point order
            (CUDA events); torch_route = the lb_moments + torch passes it replaced
  d2h     : the (1 + d) fields into the pinned staging buffer
  report  : a whole VTKReporter call (device + d2h + file write to a tmp directory)
  full_f  : what the reference's reporter path moves instead -- the populations to the host
            (flow.f D2H, q values per node), WITHOUT the host-side moments it would still compute
and prints one JSON line.  Not a bench.py line; a side measurement for DESIGN.md section 3.6."""
import json
import os
import sys
import tempfile
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

res = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reports = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
flow = lt.TaylorGreenVortex(ctx, resolution=[res] * 3, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
sim = types.SimpleNamespace(flow=flow, units=flow.units)
d = 3


def device_part():
    return flow.output_fields_pu()            # lb_output_fields: one kernel


def torch_route():
    """what the kernel replaced: lb_moments + PU conversion + reordering with torch"""
    rho, u = flow.rho_u()
    fields = torch.cat([flow.units.convert_density_lu_to_pressure_pu(rho), flow.units.convert_velocity_to_pu(u)])
    return fields.permute(0, *range(d, 0, -1)).reshape(1 + d, -1)


def cuda_ms(fn, n):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n, out


dev_ms, ordered = cuda_ms(device_part, 5)
torch_ms, _ = cuda_ms(torch_route, 5)
stage = torch.empty(ordered.shape, dtype=ordered.dtype, pin_memory=True)
d2h_ms, _ = cuda_ms(lambda: stage.copy_(ordered, non_blocking=True), 5)
host_f = torch.empty(flow.f.shape, dtype=flow.f.dtype, pin_memory=True)
full_ms, _ = cuda_ms(lambda: host_f.copy_(flow.f, non_blocking=True), 3)

with tempfile.TemporaryDirectory() as tmp:
    rep = lt.VTKReporter(interval=1, filename_base=os.path.join(tmp, "out"))
    rep(sim)                                   # allocates the staging buffer
    t0 = time.perf_counter()
    for i in range(reports):
        flow.i = i
        rep(sim)
    torch.cuda.synchronize()
    report_ms = (time.perf_counter() - t0) / reports * 1e3
    size = os.path.getsize(os.path.join(tmp, "out_00000000.vtr"))

n = res ** 3
print(json.dumps({"what": "one VTK report, D3Q19 fp32", "res": res, "nodes": n,
                  "device_ms": round(dev_ms, 3), "torch_route_ms": round(torch_ms, 3), "d2h_ms": round(d2h_ms, 3), "report_ms": round(report_ms, 2),
                  "file_bytes": size, "fields_bytes": int(stage.numel() * 4),
                  "full_f_d2h_ms": round(full_ms, 2), "full_f_bytes": int(flow.f.numel() * 4),
                  "device_gbs": round((19 + 1 + d) * 4 * n / dev_ms / 1e6, 1)}))
