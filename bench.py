#!/usr/bin/env python
"""bench.py -- MLUPS of the stream-collide hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--inner T] [--impl reference]

Workload (BASELINE.json configs[1]): TaylorGreenVortex3D, D3Q19, BGK, fp32, 256^3 per GPU.
For N > 1 the domain is 256 x 256 x (256 N), slab-decomposed along z (weak scaling, one
process per GPU under torchrun), single-plane halos exchanged every time step.

One bench "step" = one `simulation(T)` call = T lattice time steps (default T = 1000, the step
count BASELINE.json configs[0] / examples/00_simplest_TGV quote).  MLUPS = nodes * T / seconds,
as lettuce/_simulation.py:207 defines it.
  value : device-resident -- K steps timed with CUDA events on the launching stream, barrier +
          synchronize on both sides, max over ranks.
  e2e   : the same through the public API with HOST buffers: every step copies the initial
          populations from pinned host memory to the device, runs simulation(T) and copies the
          resulting populations back to pinned host memory, all inside the timed region.
The roofline object is for the step kernel: 2*Q*sizeof(dtype) bytes per node per launch
 over the average launch duration measured here, against the measured copy
bandwidth in MEASURED_PEAKS.json.  `--impl reference` times the CPU restatement of the
reference's torch path (oracle/lbm_oracle.py) on the host cores, one time step of the same
workload per bench step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

RES = 256
Q = 19
WORKLOAD = "TaylorGreenVortex3D D3Q19 BGK fp32 256^3 per GPU (BASELINE.json configs[1])"
FALLBACK_HBM_GBS = 6650.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--inner", type=int, default=1000, help="lattice time steps per bench step")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--res", type=int, default=RES, help="override the per-GPU cube edge (debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def profiled_traffic():
    """dram bytes per launch of the step kernel from the committed ncu capture, or None"""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("d3q19_f32_256_bytes_per_launch")
        except Exception:
            pass
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_oracle_mlups(res, n_steps, warm=1):
    """the reference's torch CPU path, restated (oracle/lbm_oracle.py), all host threads"""
    import torch
    from oracle import lbm_oracle as orc
    import lettuce_b200 as lt
    torch.set_num_threads(os.cpu_count() or 1)
    ctx = lt.Context(device="cpu", dtype=torch.float32, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
    lat = orc.OracleLattice("D3Q19", torch.float32)
    params = dict(tau=flow.units.relaxation_parameter_lu)
    f = flow.f
    for _ in range(warm):
        f = orc.step(lat, f, "bgk", params)
    t0 = time.perf_counter()
    for _ in range(n_steps):
        f = orc.step(lat, f, "bgk", params)
    dt = time.perf_counter() - t0
    assert bool(torch.isfinite(f).all())
    return res ** 3 * n_steps / 1e6 / dt, dt, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    res = args.res
    import torch
    from oracle import lbm_oracle as orc
    import lettuce_b200 as lt
    torch.set_num_threads(os.cpu_count() or 1)
    ctx = lt.Context(device="cpu", dtype=torch.float32, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
    lat = orc.OracleLattice("D3Q19", torch.float32)
    params = dict(tau=flow.units.relaxation_parameter_lu)
    f = flow.f
    for _ in range(args.warmup):
        f = orc.step(lat, f, "bgk", params)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        f = orc.step(lat, f, "bgk", params)          # one bench step = ONE time step (bounded sample)
    dt = time.perf_counter() - t0
    mlups = res ** 3 * args.steps / 1e6 / dt
    cores = torch.get_num_threads()
    sample = f"{args.steps} time steps of the {res}^3 lattice, one per bench step (native arm runs {args.inner} per step)"
    print(json.dumps({
        "impl": "reference", "metric": "MLUPS", "value": mlups, "unit": "MLUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("256^3", f"{res}^3"), "timesteps_per_step": 1,
                   "device": "host CPU"},
        "cpu_baseline": {"value": mlups, "unit": "MLUPS", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": mlups, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def run_native(args):
    import torch
    import torch.distributed as dist
    import lettuce_b200 as lt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    res, T = args.res, args.inner
    ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
    if world == 1:
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
    else:
        from lettuce_b200.slab import SlabTaylorGreenVortex, SlabSimulation
        flow = SlabTaylorGreenVortex(ctx, [res, res, res * world], reynolds_number=1600, mach_number=0.1,
                                     stencil=lt.D3Q19, rank=rank, world=world)
        sim = SlabSimulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [],
                             transport=os.environ.get("LB_SLAB_TRANSPORT", "auto"))
    nodes_local = res ** 3
    f_bytes = flow.f.numel() * flow.f.element_size()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident: value -----------------------------------------------------------------
    for _ in range(args.warmup):
        sim(T)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        sim(T)
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = args.steps * T                      # step kernels per rank in the timed region
    # z-slabs add, per simulation(T) call and rank: one handshake kernel up front, one after every
    # step, and the final scatter of the last received halo (lb_capi.cu step_typed)
    per_call_extra = 1 if (world > 1 and sim.transport == 'p2p') else 2   # direct stores: no final scatter
    launches_all = launches if world == 1 else args.steps * (2 * T + per_call_extra)
    mlups = nodes_local * world * launches / 1e6 / (ms_total / 1e3)
    assert bool(torch.isfinite(flow.f).all()), "populations became non-finite"

    # ---- end to end through the public API with host buffers: e2e -----------------------------------
    e2e = None
    if not args.no_e2e:
        host_in = flow.f.detach().to("cpu").pin_memory()
        host_out = torch.empty_like(host_in).pin_memory()
        e2e_steps = max(1, min(args.steps, 3))

        def one():
            flow.f.copy_(host_in, non_blocking=True)         # H2D of the step's input
            sim(T)                                          # the call a user makes
            host_out.copy_(flow.f, non_blocking=True)       # D2H of the step's result
        one()
        barrier()
        ev0.record()
        for _ in range(e2e_steps):
            one()
        ev1.record()
        barrier()
        ms2 = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        e2e = {"value": nodes_local * world * e2e_steps * T / 1e6 / (float(ms2.item()) / 1e3),
               "unit": "MLUPS", "h2d_bytes_per_step": f_bytes * world, "d2h_bytes_per_step": f_bytes * world,
               "steps": e2e_steps}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    bytes_per_launch = 2 * Q * 4 * nodes_local                      # 2*Q*sizeof(f32) per node
    avg_launch_s = ms_total / 1e3 / launches
    achieved = bytes_per_launch / avg_launch_s / 1e9
    line = {
        "metric": "MLUPS", "value": mlups, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("256^3", f"{res}^3"), "timesteps_per_step": T,
                   "parallelism": "single GPU" if world == 1 else
                   f"z-slabs x{world}, halo exchange per step ({sim.transport} transport)",
                   "l2": f"inputs larger than L2: {2 * f_bytes / 1e9:.2f} GB of populations per GPU vs 126 MB"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": profiled_traffic(), "peak_source": peak_src,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_us": avg_launch_s * 1e6,
                     "kernel": "lb::step_kernel<D3Q19,float,V=4,BGK>"},
        "clocks": clocks, "gpu_launches": launches_all,
    }
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        n_cpu = 12
        v, dt, cores = cpu_oracle_mlups(res, n_cpu)
        line["cpu_baseline"] = {"value": v, "unit": "MLUPS", "cores": cores, "kind": "port",
                                "sample": f"{n_cpu} time steps of the same {res}^3 lattice ({dt:.1f} s of CPU work), "
                                          "torch-CPU restatement of the reference path, all host threads"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
