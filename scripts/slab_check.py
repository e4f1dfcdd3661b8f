"""Run under torchrun (one rank per GPU): z-slab run of a Taylor-Green vortex with each
transport, gathered and compared with the undivided run on rank 0's GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 scripts/slab_check.py [--res 64] [--steps 20]
Prints one line `slab_check OK ...` on success, exits non-zero otherwise."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402
from lettuce_b200.slab import SlabSimulation, SlabTaylorGreenVortex  # noqa: E402


def obstacle_check(a, rank, world, dev):
    """the stock 3-D lt.Obstacle (equilibrium inlet, AntiBounceBackOutlet with its no-streaming mask, bounce-back
    sphere) in z-slabs against the undivided run: masks are sharded like the populations"""
    from lettuce_b200.slab import SlabObstacle
    ok = True
    nx = 2 * a.res
    res = [nx, a.res, a.res]
    for stencil_name in ("D3Q19", "D3Q27"):
        stencil = getattr(lt, stencil_name)
        for dtype in (torch.float32, torch.float64):
            ctx = lt.Context(device=dev, dtype=dtype, use_native=True)

            def sphere(x, y, z, ref):
                gx, gy, gz = ref
                return ((x - 0.3 * gx.max()) ** 2 + (y - 0.5 * gy.max()) ** 2 + (z - 0.45 * gz.max()) ** 2) < (0.2 * gy.max()) ** 2
            want = None
            if rank == 0:
                g = lt.Obstacle(ctx, res, reynolds_number=100, mach_number=0.05, domain_length_x=2, stencil=stencil)
                g.mask = sphere(*g.grid, g.grid)
                lt.Simulation(g, lt.BGKCollision(tau=g.units.relaxation_parameter_lu), [])(a.steps)
                want = g.f
            for transport in ("p2p-ll", "p2p-packed", "nccl"):
                flow = SlabObstacle(ctx, res, 100, 0.05, 2, rank=rank, world=world, stencil=stencil)
                flow.mask = sphere(*flow.grid, flow.global_grid)
                sim = SlabSimulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [], transport=transport)
                assert sim._pack_streaming_bits(sim._desc()) is not None    # the outlet's mask is really there
                sim(1); sim(a.steps - 3); sim(2)
                parts = [torch.empty_like(flow.f) for _ in range(world)]
                dist.all_gather(parts, flow.f.contiguous())
                if rank == 0:
                    got = torch.cat(parts, dim=-1)
                    exact = bool(torch.equal(got, want))
                    print(f"slab_check stock Obstacle {res} {stencil_name} BGK {dtype} world={world} {transport}: "
                          f"bit-exact={exact} finite={bool(torch.isfinite(got).all())}", flush=True)
                    ok = ok and exact
                sim.close()
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=64)
    ap.add_argument("--steps", type=int, default=21)
    ap.add_argument("--stencil", default="D3Q19")
    ap.add_argument("--collision", default="bgk")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    stencil = getattr(lt, a.stencil)
    res = [a.res, a.res, a.res]
    ok = True
    for dtype in (torch.float32, torch.float64):
        ctx = lt.Context(device=dev, dtype=dtype, use_native=True)

        def collision(flow):
            tau = flow.units.relaxation_parameter_lu
            return {"bgk": lt.BGKCollision(tau), "trt": lt.TRTCollision(tau, 0.9), "kbc": lt.KBCCollision(),
                    "none": lt.NoCollision()}[a.collision]
        want = e_want = None
        if rank == 0:
            g = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=stencil)
            rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(g), interval=7, out=None)
            lt.Simulation(g, collision(g), [rep])(a.steps)
            want, e_want = g.f, [r[2] for r in rep.out]
        ens_want = None
        if rank == 0:     # enstrophy of the undivided run (3-plane velocity halo across slabs)
            g2 = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=stencil)
            rep2 = lt.ObservableReporter(lt.Enstrophy(g2), interval=5, out=None)
            lt.Simulation(g2, collision(g2), [rep2])(10)
            ens_want = [r[2] for r in rep2.out]
        for transport in ("p2p-ll", "p2p", "p2p-packed", "p2p-push", "nccl"):
            if transport == "p2p-ll":
                flow = SlabTaylorGreenVortex(ctx, res, 400, 0.1, stencil, rank=rank, world=world)
                rep = lt.ObservableReporter(lt.Enstrophy(flow), interval=5, out=None)
                sim = SlabSimulation(flow, collision(flow), [rep], transport=transport)
                sim(10)
                if rank == 0:
                    tol = 1e-5 if dtype == torch.float32 else 1e-12
                    ens = [r[2] for r in rep.out]
                    ens_ok = len(ens) == len(ens_want) and all(abs(x - y) <= tol * abs(y) for x, y in zip(ens, ens_want))
                    print(f"slab_check {a.stencil} {a.collision} {dtype} world={world} enstrophy across slabs: "
                          f"{ens} vs undivided {ens_want} ok={ens_ok}", flush=True)
                    ok = ok and ens_ok
                sim.close()
            flow = SlabTaylorGreenVortex(ctx, res, 400, 0.1, stencil, rank=rank, world=world)
            rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=7, out=None)
            sim = SlabSimulation(flow, collision(flow), [rep], transport=transport)
            # several calls of different (odd / even) lengths: exercises the ping-pong bookkeeping
            sim(1); sim(a.steps - 3); sim(2)
            parts = [torch.empty_like(flow.f) for _ in range(world)]
            dist.all_gather(parts, flow.f.contiguous())
            if rank == 0:
                got = torch.cat(parts, dim=-1)
                exact = bool(torch.equal(got, want))
                err = float((got - want).abs().max() / want.abs().max())
                e_got = [r[2] for r in rep.out]
                e_ok = len(e_got) == len(e_want) and all(abs(x - y) <= 1e-6 * abs(y) for x, y in zip(e_got, e_want))
                print(f"slab_check {a.stencil} {a.collision} {dtype} world={world} {transport}: "
                      f"bit-exact={exact} rel_err={err:.2e} energy_ok={e_ok}", flush=True)
                ok = ok and err <= (1e-5 if dtype == torch.float32 else 1e-12) and e_ok
            sim.close()
    ok = obstacle_check(a, rank, world, dev) and ok
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    if rank == 0 and flag.item():
        print("slab_check OK")
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
