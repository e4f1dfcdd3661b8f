"""Single-GPU timing of the step kernel's halo variants (D3Q19 BGK fp32 256^3 by default):
where does the multi-GPU overhead come from?  Raw C-ABI calls, CUDA-event timing."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402

nat = lt.native
L = nat.lib()
res = [int(os.environ.get("RES", 256))] * 3
if os.environ.get("RES3"):
    res = [int(v) for v in os.environ["RES3"].split("x")]
steps = int(os.environ.get("STEPS", 100))
ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
tau = flow.units.relaxation_parameter_lu
desc = nat.make_desc("D3Q19", torch.float32, res, nat.LB_COLL_BGK, tau)
A, B = flow.f, torch.empty_like(flow.f)
lines = res[0] * res[1]
send = [torch.zeros([5, lines], device="cuda") for _ in range(2)]
recv = [torch.zeros([5, lines], device="cuda") for _ in range(2)]
flags = torch.full([8], 1 << 30, dtype=torch.int32, device="cuda")
flags[2:] = 0
st = torch.cuda.current_stream().cuda_stream


def run(name, make_halo):
    global A, B
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(2):
        torch.cuda.synchronize()
        e0.record()
        for s in range(steps):
            fl = nat.LbFields()
            fl.f, fl.f_next = A.data_ptr(), B.data_ptr()
            h = make_halo(B)
            if h is not None:
                fl.halo = C.pointer(h)
            nat.check(L.lb_step(desc, fl, 1, st))
            A, B = B, A
        e1.record()
        torch.cuda.synchronize()
    print(f"{name:45s} {e0.elapsed_time(e1) / steps * 1e3:8.1f} us/step", flush=True)


def packed(halo_in):
    def mk(out):
        h = nat.LbHalo()
        h.mode = nat.LB_HALO_PACKED
        h.send_hi[0], h.send_lo[0] = send[0].data_ptr(), send[1].data_ptr()
        if halo_in:
            h.recv_lo[0], h.recv_hi[0] = recv[0].data_ptr(), recv[1].data_ptr()
        return h
    return mk


def direct(out):
    h = nat.LbHalo()
    h.mode = nat.LB_HALO_DIRECT
    h.peer_up_f[0] = h.peer_down_f[0] = out.data_ptr()      # ourselves: same stores as the periodic wrap
    h.signal_up = h.signal_down = flags.data_ptr() + 16
    h.wait_slots = flags.data_ptr()
    return h


run("periodic (no halo)", lambda out: None)
run("PACKED edge stores", packed(False))
run("PACKED edge stores + halo_in loads", packed(True))
run("DIRECT edge stores to own buffer + sync kernel", direct)
run("periodic (no halo) again", lambda out: None)
