#!/usr/bin/env python
"""bench.py -- MLUPS of the stream-collide hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--inner T] [--impl reference] [--no-extra]

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
          resulting populations back to pinned host memory, all inside the timed region.  Beside it:
          the same with ONE time step per call (PCIe-bound by construction) and through the C ABI's
          host-buffer entry lb_run_host.
  parity: before anything is timed, the SAME 256^3 initial populations are stepped by the CPU checker
          (the unmodified reference from baseline/_ref when installed, else the oracle port) and by the GPU;
          rel_err = max|a-b| / max|b| must be <= 1e-5 (fp32 bar of BASELINE.json) or the bench fails.
          For N > 1: a 64^2 x 64N slab run must equal the undivided single-GPU run bit for bit.
  extra : the other BASELINE configs measured in the same run (device-resident, CUDA events): N = 1 ->
          configs 1, 3 (fp32 / fp64, all-local and stock boundary lists), 4 (reporter interval 1 / 10 / none);
          N > 1 -> config 5: 1024 x 1024 x 128 per GPU (weak) and the global 1024^3 (strong).
The roofline object is for the step kernel: 2*Q*sizeof(dtype) bytes per node per launch
 over the average launch duration measured here, against the measured copy
bandwidth in MEASURED_PEAKS.json.  `--impl reference` times the reference's own CPU PyTorch path
(baseline/_ref, installed by baseline/install_ref.sh; the oracle port when that is absent) on the host
cores, one time step of the same workload per bench step.
"""
import argparse
import gc
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
PARITY_TOL = 1e-5


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--inner", type=int, default=1000, help="lattice time steps per bench step")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--res", type=int, default=RES, help="override the per-GPU cube edge (debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU leg (then no parity object either)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the other BASELINE configs")
    ap.add_argument("--cpu-steps", type=int, default=8, help="time steps of the CPU checker (parity + cpu_baseline)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cited_traffic():
    """dram bytes per launch of the step kernel from the committed ncu capture (a citation, not measured here)"""
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


# ---------------------------------------------------------------------------------------------------
# the CPU leg: the reference's own torch path when baseline/_ref is installed, else the oracle port
# ---------------------------------------------------------------------------------------------------
def cpu_reference_run(res, n_steps, warm=1, want_state=False):
    """config 2 on the host cores.  Returns dict(mlups, seconds, cores, kind, f0, f_end, steps_total):
    f0 / f_end (numpy, only when want_state) = the initial populations and the state after `warm + n_steps`
    time steps -- what the GPU parity check replays."""
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
    from oracle import reference_shim
    out = {"cores": torch.get_num_threads()}
    if reference_shim.reference_available():
        import contextlib
        import io
        rlt = reference_shim.import_reference()
        ctx = rlt.Context(device=torch.device("cpu"), dtype=torch.float32, use_native=False)
        with contextlib.redirect_stdout(io.StringIO()):      # initialize_f_neq prints the device (_flow.py:326)
            flow = rlt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1,
                                         stencil=rlt.D3Q19)
        sim = rlt.Simulation(flow, rlt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
        f0 = flow.f.detach().clone().numpy() if want_state else None
        if warm:
            sim(warm)
        t0 = time.perf_counter()
        sim(n_steps)                                         # the reference's own hot loop (_simulation.py:195-207)
        dt = time.perf_counter() - t0
        out.update(kind="reference", f0=f0, f_end=flow.f.detach().numpy() if want_state else None,
                   what="the unmodified reference's lt.Simulation on Context(device='cpu', use_native=False) "
                        "(baseline/_ref)")
    else:
        from oracle import lbm_oracle as orc
        import lettuce_b200 as lt
        ctx = lt.Context(device="cpu", dtype=torch.float32, use_native=False)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
        lat = orc.OracleLattice("D3Q19", torch.float32)
        params = dict(tau=flow.units.relaxation_parameter_lu)
        f = flow.f
        f0 = f.clone().numpy() if want_state else None
        for _ in range(warm):
            f = orc.step(lat, f, "bgk", params)
        t0 = time.perf_counter()
        for _ in range(n_steps):
            f = orc.step(lat, f, "bgk", params)
        dt = time.perf_counter() - t0
        out.update(kind="port", f0=f0, f_end=f.numpy() if want_state else None,
                   what="torch-CPU restatement of the reference path (oracle/lbm_oracle.py); baseline/_ref not installed")
    out.update(mlups=res ** 3 * n_steps / 1e6 / dt, seconds=dt, steps_total=warm + n_steps)
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    res = args.res
    r = cpu_reference_run(res, args.steps, warm=args.warmup)
    sample = (f"{args.steps} time steps of the {res}^3 lattice, one per bench step (the native arm runs {args.inner} "
              f"per step); {r['what']}")
    print(json.dumps({
        "impl": "reference", "metric": "MLUPS", "value": r["mlups"], "unit": "MLUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["seconds"] / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("256^3", f"{res}^3"), "timesteps_per_step": 1,
                   "device": "host CPU"},
        "cpu_baseline": {"value": r["mlups"], "unit": "MLUPS", "cores": r["cores"], "kind": r["kind"], "sample": sample},
        "e2e": {"value": r["mlups"], "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------------
# the other BASELINE configs (device-resident, CUDA events); every line carries mlups, us_per_step, frac
# ---------------------------------------------------------------------------------------------------
def _timed(sim, steps, warm, dev, world=1):
    import torch
    import torch.distributed as dist
    sim(warm)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sim(steps)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()) / 1e3


def _extra_line(name, nodes, q, esz, seconds, steps, peak, **more):
    mlups = nodes * steps / 1e6 / seconds
    gbs = mlups * 1e6 * 2 * q * esz / 1e9
    line = {"case": name, "mlups": round(mlups, 1), "us_per_step": round(seconds / steps * 1e6, 2),
            "bytes_per_update": 2 * q * esz, "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4)}
    line.update(more)
    return line


def extras_single_gpu(dev, peak):
    """BASELINE configs 1, 3 and 4 on one GPU"""
    import torch
    import lettuce_b200 as lt
    out = []

    def free():
        gc.collect()
        torch.cuda.empty_cache()

    def guarded(fn):
        try:
            fn()
        except Exception as e:                        # an extra must never take the headline down
            out.append({"case": fn.__name__, "error": f"{type(e).__name__}: {e}"[:300]})
        free()

    def cfg1():   # examples/00_simplest_TGV.py:9-21 at 256^2
        ctx = lt.Context(device=dev, dtype=torch.float64, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=256, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
        t = _timed(sim, 1000, 1000, dev)
        out.append(_extra_line("cfg1 TGV2D D2Q9 BGK f64 256^2 x1000 steps", 256 * 256, 9, 8, t, 1000, peak,
                               note="4.7 MB of state, never HBM-bound: tile-resident kernel (tiles in shared memory, one grid "
                                    "barrier per ~6 time steps); the limit is the arithmetic of the tiles incl. their halo rings"))

    def cfg2_reporter():   # the headline lattice with the reporter the reference's examples attach (03_outputs_TGV.py)
        ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=256, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
        rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [rep])
        t = _timed(sim, 200, 20, dev)
        out.append(_extra_line("cfg2 TGV3D D3Q19 BGK f32 256^3 with a fused energy reporter every step", 256 ** 3, 19, 4, t,
                               200, peak))

    class LocalObstacle(lt.Obstacle):
        """BASELINE config 3, "bounce-back + equilibrium inlet/outlet" read as all node-local boundaries
        (SURVEY 8d): equilibrium boundaries on the x = 0 and x = nx-1 columns, bounce-back on the mask"""

        @property
        def boundaries(self):
            x = self.grid[0]
            u_in = self.units.characteristic_velocity_pu * self._unit_vector()
            return [lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                    lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x - x.max()) < 1e-6, velocity=u_in),
                    lt.BounceBackBoundary(self.mask)]

    def obstacle(dtype, stock):
        ctx = lt.Context(device=dev, dtype=dtype, use_native=True)
        cls = lt.Obstacle if stock else LocalObstacle
        flow = cls(ctx, [2048, 1024], reynolds_number=300, mach_number=0.01, domain_length_x=1)
        x, y = flow.grid
        flow.mask = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.05 * y.max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
        t = _timed(sim, 500, 100, dev)
        esz = 4 if dtype == torch.float32 else 8
        nm = "f32" if esz == 4 else "f64"
        what = ("stock lt.Obstacle list: equilibrium inlet, AntiBounceBackOutlet, bounce-back (masked streaming)" if stock
                else "bounce-back cylinder + equilibrium inlet and outlet columns (all node-local)")
        out.append(_extra_line(f"cfg3 Obstacle2D D2Q9 BGK {nm} 2048x1024, {what}", 2048 * 1024, 9, esz, t, 500, peak))

    def cfg3():
        for dtype in (torch.float32, torch.float64):
            for stock in (False, True):
                obstacle(dtype, stock)
                free()

    def cfg4():
        for interval in (None, 10, 1):
            ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
            flow = lt.TaylorGreenVortex(ctx, resolution=512, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q27)
            reps = [lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=interval, out=None)] if interval else []
            sim = lt.Simulation(flow, lt.KBCCollision(), reps)
            t = _timed(sim, 50, 10, dev)
            line = _extra_line(f"cfg4 TGV3D D3Q27 KBC f32 512^3, fused energy reporter interval={interval}",
                               512 ** 3, 27, 4, t, 50, peak)
            if reps:
                e = [r[2] for r in reps[0].out]
                line["energy_first_last"] = [e[0], e[-1]]
                line["energy_monotone"] = all(b <= a * (1 + 1e-6) for a, b in zip(e, e[1:]))
            out.append(line)
            del sim, flow, reps
            free()

    def cfg5_one_gpu():   # the single-GPU rate the config-5 efficiencies are quoted against
        ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=[1024, 1024, 128], reynolds_number=1600, mach_number=0.1,
                                    stencil=lt.D3Q19)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
        t = _timed(sim, 30, 5, dev)
        out.append(_extra_line("cfg5 slab geometry on ONE GPU: TGV3D D3Q19 BGK f32 1024x1024x128 (periodic by itself)",
                               1024 * 1024 * 128, 19, 4, t, 30, peak))

    for fn in (cfg1, cfg2_reporter, cfg3, cfg4, cfg5_one_gpu):
        guarded(fn)
    return out


def extras_multi_gpu(dev, rank, world, peak):
    """BASELINE config 5: TGV3D D3Q19 BGK fp32, weak = 1024 x 1024 x 128 per GPU, strong = the global 1024^3"""
    import torch
    import lettuce_b200 as lt
    from lettuce_b200.slab import SlabSimulation, SlabTaylorGreenVortex
    out = []
    cases = [("weak", [1024, 1024, 128 * world])]
    if world != 8:                                     # at 8 GPUs the two geometries coincide
        cases.append(("strong", [1024, 1024, 1024]))
    for mode, res in cases:
        try:
            ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
            flow = SlabTaylorGreenVortex(ctx, res, 1600, 0.1, lt.D3Q19, rank=rank, world=world)
            sim = SlabSimulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [],
                                 transport=os.environ.get("LB_SLAB_TRANSPORT", "auto"))
            torch.cuda.empty_cache()
            steps = 30 if res[2] // world <= 128 else 12
            t = _timed(sim, steps, 5, dev, world)
            finite = bool(torch.isfinite(flow.f).all())
            n = res[0] * res[1] * res[2]
            name = (f"cfg5 {mode}{'+strong' if world == 8 else ''} TGV3D D3Q19 BGK f32 {res[0]}x{res[1]}x{res[2]} on {world} GPUs "
                    f"({res[2] // world} planes per GPU, {sim.transport})")
            out.append(_extra_line(name, n, 19, 4, t, steps, peak * world, n_gpus=world,
                                   mlups_per_gpu=round(n * steps / 1e6 / t / world, 1), finite=finite))
            sim.close()
            del sim, flow
        except Exception as e:
            out.append({"case": f"cfg5 {mode}", "error": f"{type(e).__name__}: {e}"[:300]})
        gc.collect()
        torch.cuda.empty_cache()
    return out


def slab_bit_exact(dev, rank, world):
    """64 x 64 x 64N TGV, D3Q19 BGK fp32, 21 steps in three calls: the gathered slabs of the default transport
    against the undivided run on rank 0's GPU.  True / False on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist
    import lettuce_b200 as lt
    from lettuce_b200.slab import SlabSimulation, SlabTaylorGreenVortex
    res = [64, 64, 64 * world]
    ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
    flow = SlabTaylorGreenVortex(ctx, res, 400, 0.1, lt.D3Q19, rank=rank, world=world)
    sim = SlabSimulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [],
                         transport=os.environ.get("LB_SLAB_TRANSPORT", "auto"))
    sim(1); sim(18); sim(2)
    parts = [torch.empty_like(flow.f) for _ in range(world)]
    dist.all_gather(parts, flow.f.contiguous())
    ok = None
    if rank == 0:
        g = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=lt.D3Q19)
        lt.Simulation(g, lt.BGKCollision(tau=g.units.relaxation_parameter_lu), [])(21)
        ok = bool(torch.equal(torch.cat(parts, dim=-1), g.f))
    sim.close()
    return ok


def _bind_to_gpu_numa_node(local):
    """Run this rank on the CPUs of the NUMA node its GPU hangs off, so that the pinned host buffers of the end-to-end
    leg (allocated below: first touch) and the copies do not cross the socket interconnect.  Returns
    {"node": n, "cpus": k} or None where the box has no such information (single node, container without sysfs).
    LB_NO_NUMA_BIND=1 leaves the affinity alone (A/B)."""
    if os.environ.get("LB_NO_NUMA_BIND"):
        return None
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return {"node": node, "cpus": len(allowed)}
    except Exception:
        return None


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
    numa = _bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    res, T = args.res, args.inner
    ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)

    # ---- parity first: nothing is timed on a path that does not reproduce the reference -----------------
    parity, cpu = None, None
    bit_exact = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_run(res, args.cpu_steps, warm=1, want_state=True)
        f0 = torch.from_numpy(cpu["f0"]).to(dev)
        chk_flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
        init_err = float((chk_flow.f - f0).abs().max() / f0.abs().max())   # lb_initialize vs the reference's f0
        chk_flow.f = f0
        lt.Simulation(chk_flow, lt.BGKCollision(tau=chk_flow.units.relaxation_parameter_lu), [])(cpu["steps_total"])
        want = torch.from_numpy(cpu["f_end"]).to(dev)
        rel = float((chk_flow.f - want).abs().max() / want.abs().max())
        parity = {"rel_err": rel, "tol": PARITY_TOL, "steps": cpu["steps_total"], "ok": rel <= PARITY_TOL,
                  "against": cpu["kind"], "lattice": f"{res}^3", "initial_state_rel_err": init_err}
        del chk_flow, f0, want
        cpu.pop("f0"), cpu.pop("f_end")
        gc.collect()
        torch.cuda.empty_cache()
    if world > 1:
        bit_exact = slab_bit_exact(dev, rank, world)

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
    # z-slabs add per simulation(T) call and rank: 'p2p-ll' one scatter of the last received halo; the
    # handshake transports one handshake kernel up front and one after every step (+ the scatter, packed ones)
    if world == 1:
        launches_all = launches
    elif sim.transport == 'p2p-ll':
        launches_all = args.steps * (T + 1)
    else:
        launches_all = args.steps * (2 * T + (1 if sim.transport == 'p2p' else 2))
    mlups = nodes_local * world * launches / 1e6 / (ms_total / 1e3)
    assert bool(torch.isfinite(flow.f).all()), "populations became non-finite"

    # ---- end to end through the public API with host buffers: e2e -----------------------------------
    e2e = None
    if not args.no_e2e:
        host_in = flow.f.detach().to("cpu").pin_memory()
        host_out = torch.empty_like(host_in).pin_memory()
        e2e_steps = max(1, min(args.steps, 3))

        def one(t_steps):
            flow.f.copy_(host_in, non_blocking=True)         # H2D of the step's input
            sim(t_steps)                                    # the call a user makes
            host_out.copy_(flow.f, non_blocking=True)       # D2H of the step's result

        def timed_e2e(t_steps):
            one(t_steps)
            barrier()
            ev0.record()
            for _ in range(e2e_steps):
                one(t_steps)
            ev1.record()
            barrier()
            ms2 = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
            return nodes_local * world * e2e_steps * t_steps / 1e6 / (float(ms2.item()) / 1e3)

        e2e = {"value": timed_e2e(T), "unit": "MLUPS", "h2d_bytes_per_step": f_bytes * world,
               "d2h_bytes_per_step": f_bytes * world, "steps": e2e_steps, "timesteps_per_call": T,
               "one_timestep_per_call": {"value": timed_e2e(1), "unit": "MLUPS",
                                         "note": "same copies around ONE time step: PCIe-bound by construction "
                                                 "(2 x %.2f GB per call)" % (f_bytes / 1e9)}}
        if world == 1:
            # the C ABI's own host-buffer entry (plain pointers, no torch): lb_run_host
            import ctypes as C
            nat = lt.native
            desc = sim._desc()
            buf = host_in.clone().pin_memory()
            for t_steps, key in ((T, "capi_lb_run_host"), (1, "capi_lb_run_host_one_timestep")):
                nat.check(nat.lib().lb_run_host(desc, buf.data_ptr(), None, None, None, 0, min(t_steps, 2), None))  # warm
                t0 = time.perf_counter()
                nat.check(nat.lib().lb_run_host(desc, buf.data_ptr(), None, None, None, 0, t_steps, None))
                dt = time.perf_counter() - t0                # lb_run_host synchronises before it returns
                e2e[key] = {"value": nodes_local * t_steps / 1e6 / dt, "unit": "MLUPS", "timesteps_per_call": t_steps,
                            "timer": "host wall clock around the blocking C call"}
        del host_in, host_out

    extra = None
    if not args.no_extra:
        peak_x, _ = measured_peak()
        if world > 1:
            sim.close()
        sim = flow = None
        gc.collect()
        torch.cuda.empty_cache()
        extra = extras_single_gpu(dev, peak_x) if world == 1 else extras_multi_gpu(dev, rank, world, peak_x)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    bytes_per_launch = 2 * Q * 4 * nodes_local                      # 2*Q*sizeof(f32) per node
    avg_launch_s = ms_total / 1e3 / launches
    achieved = bytes_per_launch / avg_launch_s / 1e9
    transport = os.environ.get("LB_SLAB_TRANSPORT", "auto")
    if world > 1 and transport == "auto":
        from lettuce_b200.slab import auto_transport
        transport = auto_transport([res, res, res])
    line = {
        "metric": "MLUPS", "value": mlups, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("256^3", f"{res}^3"), "timesteps_per_step": T,
                   "parallelism": "single GPU" if world == 1 else
                   f"z-slabs x{world}, halo exchange per step ({transport} transport)",
                   "l2": f"inputs larger than L2: {2 * f_bytes / 1e9:.2f} GB of populations per GPU vs 126 MB"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "traffic_cited": cited_traffic(),
                     "traffic_note": "traffic_cited = dram__bytes_read.sum + dram__bytes_write.sum of one launch from the "
                                     "committed ncu --set full capture (profiles/roofline_traffic.json), not measured in this run",
                     "peak_source": peak_src,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_us": avg_launch_s * 1e6,
                     "kernel": "lb::step_kernel<D3Q19,float,V=4,BGK>"},
        "clocks": clocks, "gpu_launches": launches_all,
    }
    if e2e is not None:
        e2e["host_numa_binding"] = numa   # rank 0's: {"node", "cpus"} or null where the box offers none
        line["e2e"] = e2e
    if parity is not None:
        line["parity"] = parity
    if bit_exact is not None:
        line["slab_bit_exact"] = bit_exact
    if cpu is not None:
        line["cpu_baseline"] = {"value": cpu["mlups"], "unit": "MLUPS", "cores": cpu["cores"], "kind": cpu["kind"],
                                "sample": f"{args.cpu_steps} time steps of the same {res}^3 lattice "
                                          f"({cpu['seconds']:.1f} s of CPU work, all host threads): {cpu['what']}"}
    if extra is not None:
        line["extra"] = extra
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"parity FAILED: rel_err {parity['rel_err']:.3e} > {PARITY_TOL}")
    if bit_exact is False:
        raise SystemExit("slab run is not bit-exact against the undivided run")


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
