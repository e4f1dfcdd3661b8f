"""BASELINE config 5 under torchrun (one rank per GPU): TaylorGreenVortex3D D3Q19 BGK fp32,
  --mode strong : the global 1024^3 lattice cut into WORLD_SIZE z-slabs (2 / 4 / 8 GPUs),
  --mode weak   : 1024 x 1024 x 128 per GPU.
Each slab is initialised on its own GPU by lb_initialize (ghost planes, no global field anywhere).
Device-resident timing: CUDA events, barrier + synchronize on both sides, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29519 scripts/bench_cfg5.py --mode strong"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402
from lettuce_b200.slab import SlabSimulation, SlabTaylorGreenVortex  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="strong", choices=["strong", "weak"])
ap.add_argument("--nxy", type=int, default=1024)
ap.add_argument("--nz", type=int, default=None, help="global (strong) / per-GPU (weak) extent along z")
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--warmup", type=int, default=10)
ap.add_argument("--transport", default="p2p-packed")
a = ap.parse_args()

world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nz_global = (a.nz or 1024) if a.mode == "strong" else (a.nz or 128) * world
res = [a.nxy, a.nxy, nz_global]
ctx = lt.Context(device=dev, dtype=torch.float32, use_native=True)
if world == 1:
    flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [])
else:
    flow = SlabTaylorGreenVortex(ctx, res, 1600, 0.1, lt.D3Q19, rank=rank, world=world)
    sim = SlabSimulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [], transport=a.transport)
torch.cuda.empty_cache()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


sim(a.warmup)
barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
sim(a.steps)
e1.record()
barrier()
ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
finite = bool(torch.isfinite(flow.f).all())
if rank == 0:
    n = res[0] * res[1] * res[2]
    sec = float(ms.item()) / 1e3
    mlups = n * a.steps / 1e6 / sec
    print(json.dumps({"case": f"cfg5 {a.mode} TGV3D D3Q19 BGK f32 {res[0]}x{res[1]}x{res[2]} on {world} GPU(s), {a.transport}",
                      "mlups": round(mlups, 1), "us_per_step": round(sec / a.steps * 1e6, 1), "n_gpus": world,
                      "mlups_per_gpu": round(mlups / world, 1), "nodes_per_gpu": n // world, "finite": finite,
                      "peak_mem_gb": round(torch.cuda.max_memory_allocated(dev) / 1e9, 2)}), flush=True)
if world > 1:
    dist.destroy_process_group()
