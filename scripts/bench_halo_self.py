"""The z-slab halo path of the step kernel on ONE GPU with the slab as its own neighbour (the "peer" pointers are
the rank's own receive buffers): the kernel-side cost of a transport without NVLink and without a second rank.
    python scripts/bench_halo_self.py [res] [steps] [mode: ll | remote | none]
Checks the result bit for bit against the plain periodic run and prints one JSON line."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native
res = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
mode = sys.argv[3] if len(sys.argv) > 3 else "ll"
ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
flow = lt.TaylorGreenVortex(ctx, resolution=[res] * 3, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
f0 = flow.f.clone()
sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
L = nat.lib()
desc = sim._desc()
st = torch.cuda.current_stream().cuda_stream


def timed(fn):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / steps


if mode == "none":
    us = timed(lambda: sim(steps))
    ok = True
else:
    lines, nf = res * res, 5
    A, B = f0.clone(), torch.empty_like(f0)
    flags = torch.zeros([8], dtype=torch.int32, device="cuda")
    h = nat.LbHalo()
    if mode == "ll":
        nb = L.lb_halo_ll_bytes(desc)
        lo = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
        hi = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
        h.mode = nat.LB_HALO_LL
        for k in range(3):
            h.recv_lo[k], h.recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
            h.peer_up_recv_lo[k], h.peer_down_recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
    else:
        lo = [torch.zeros([nf, lines], dtype=torch.float32, device="cuda") for _ in range(2)]
        hi = [torch.zeros([nf, lines], dtype=torch.float32, device="cuda") for _ in range(2)]
        h.mode = nat.LB_HALO_REMOTE
        for k in range(2):
            h.recv_lo[k], h.recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
            h.peer_up_recv_lo[k], h.peer_down_recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
        h.signal_up, h.signal_down = flags.data_ptr(), flags.data_ptr() + 4    # our own wait slots
    h.wait_slots = flags.data_ptr()
    epoch = [0]

    def call():
        fl = nat.LbFields()
        fl.f, fl.f_next = A.data_ptr(), B.data_ptr()
        h.epoch = epoch[0]
        fl.halo = C.pointer(h)
        nat.check(L.lb_step(desc, fl, steps, st))
        epoch[0] += steps if mode == "ll" else steps + 1
    assert steps % 2 == 0
    us = timed(call)
    sim(2 * steps)
    ok = bool(torch.equal(A, flow.f)) and int(flags[2].item()) == 0
print(json.dumps({"what": f"D3Q19 BGK f32 {res}^3, slab = its own neighbour, halo mode {mode}", "us_per_step": round(us, 2),
                  "bit_exact_vs_periodic": ok}))
