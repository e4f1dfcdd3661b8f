"""world_size-2 (and 3) gloo tests on CPU of the multi-GPU host logic: slab partition, ring
neighbours, packed-halo exchange, unpacking.  The local update is the oracle (injected stepper),
so the only thing under test is lettuce_b200.slab's exchange logic: the gathered slabs must equal
the oracle run on the undivided lattice (bit for bit for pure streaming)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, stencil, res, n_steps, collision, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import lettuce_b200 as lt
        from lettuce_b200.slab import SlabSimulation, slab_range
        from oracle import lbm_oracle as orc
        torch.set_num_threads(1)
        torch.manual_seed(1234)
        S = getattr(lt, stencil)()
        lat = orc.OracleLattice(stencil, torch.float64)
        w = torch.tensor(S.w, dtype=torch.float64).reshape([S.q] + [1] * len(res))
        f_global = w * (1.0 + 0.1 * torch.rand([S.q, *res], dtype=torch.float64))
        z0, z1 = slab_range(res[-1], rank, world)

        class LocalFlow:            # the minimum of the Flow interface the host path touches
            pass
        flow = LocalFlow()
        flow.context = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
        flow.stencil = S
        flow.resolution = list(res[:-1]) + [z1 - z0]
        flow.f = f_global[..., z0:z1].contiguous()
        flow.i = 0
        params = dict(tau=0.7) if collision != "none" else {}
        sim = SlabSimulation(flow, None, [], transport="host",
                             local_stepper=lambda f: orc.step(lat, f, collision, params))
        assert (sim.up, sim.down) == ((rank + 1) % world, (rank - 1) % world)
        sim(n_steps)
        assert flow.i == n_steps
        parts = [torch.empty_like(flow.f) for _ in range(world)]
        dist.all_gather(parts, flow.f.contiguous())
        if rank == 0:
            ref = orc.run(lat, f_global, n_steps, collision, params)
            got = torch.cat(parts, dim=-1)
            q.put(float((got - ref).abs().max()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,stencil,res,collision", [
    (2, "D3Q19", [4, 5, 8], "bgk"),
    (2, "D3Q27", [3, 4, 6], "none"),
    (3, "D2Q9", [6, 9], "bgk"),
    (2, "D2Q9", [5, 2], "none"),        # one-plane slabs: both faces are the same plane
])
def test_slab_exchange_matches_undivided_oracle(world, stencil, res, collision):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, stencil, res, 5, collision, q))
             for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    err = q.get(timeout=10)
    # pure streaming is data movement: exact.  With a collision the torch matmul behind `j` may
    # block differently for a slab than for the whole lattice: round-off only.
    assert err == 0.0 if collision == "none" else err < 1e-14


def _masked_worker(rank, world, port, stencil, res, n_steps, q):
    """bounce-back solids + a random per-(q, node) no-streaming mask, both sharded along the last axis: the
    destination bits of a population that crosses a slab face live on the RECEIVING rank"""
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import lettuce_b200 as lt
        from lettuce_b200.slab import SlabSimulation, slab_range
        from oracle import lbm_oracle as orc
        torch.set_num_threads(1)
        torch.manual_seed(4321)
        S = getattr(lt, stencil)()
        lat = orc.OracleLattice(stencil, torch.float64)
        w = torch.tensor(S.w, dtype=torch.float64).reshape([S.q] + [1] * len(res))
        f_global = w * (1.0 + 0.1 * torch.rand([S.q, *res], dtype=torch.float64))
        solid = torch.rand(res) < 0.15
        nsm = (torch.rand([S.q, *res]) < 0.1).to(torch.uint8)
        ncm = solid.to(torch.uint8)
        bnd = [orc.OracleBoundary(kind="bounce_back", mask=solid)]
        z0, z1 = slab_range(res[-1], rank, world)

        class LocalFlow:
            pass
        flow = LocalFlow()
        flow.context = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
        flow.stencil = S
        flow.resolution = list(res[:-1]) + [z1 - z0]
        flow.f = f_global[..., z0:z1].contiguous()
        flow.i = 0
        params = dict(tau=0.8)
        bnd_l = [orc.OracleBoundary(kind="bounce_back", mask=solid[..., z0:z1])]
        ncm_l, nsm_l = ncm[..., z0:z1].contiguous(), nsm[..., z0:z1].contiguous()
        sim = SlabSimulation(flow, None, [], transport="host",
                             local_stepper=(lambda f: orc.collide(lat, f, "bgk", params, ncm=ncm_l, boundaries=bnd_l),
                                            lambda f: orc.stream(lat, f, None)),
                             host_no_streaming_mask=nsm_l)
        sim(n_steps)
        parts = [torch.empty_like(flow.f) for _ in range(world)]
        dist.all_gather(parts, flow.f.contiguous())
        # the 3-plane velocity halo of the enstrophy reporter (SlabSimulation._exchange_velocity_halo)
        u_global = torch.arange(S.d * f_global[0].numel(), dtype=torch.float64).reshape([S.d, *res])
        ok_u = True
        g = sim.GRADIENT_RADIUS
        if z1 - z0 >= g:
            ext = sim._exchange_velocity_halo(u_global[..., z0:z1].contiguous())
            idx = torch.arange(z0 - g, z1 + g) % res[-1]
            ok_u = bool(torch.equal(ext, u_global[..., idx]))
        oks = [None] * world
        dist.all_gather_object(oks, ok_u)
        if rank == 0:
            ref = orc.run(lat, f_global, n_steps, "bgk", params, ncm=ncm, nsm=nsm, boundaries=bnd)
            got = torch.cat(parts, dim=-1)
            q.put((float((got - ref).abs().max()), all(oks)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,stencil,res", [(2, "D3Q19", [4, 5, 8]), (3, "D2Q9", [6, 9]), (2, "D3Q27", [3, 4, 2])])
def test_slab_masked_streaming_and_velocity_halo(world, stencil, res):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_masked_worker, args=(r, world, port, stencil, res, 4, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    err, u_ok = q.get(timeout=10)
    assert err < 1e-14 and u_ok


def test_slab_obstacle_is_the_slice_of_the_undivided_flow():
    import lettuce_b200 as lt
    from lettuce_b200._simulation import build_masks
    from lettuce_b200.slab import SlabObstacle
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    res = [16, 8, 12]
    g = lt.Obstacle(ctx, res, reynolds_number=50, mach_number=0.05, domain_length_x=2, stencil=lt.D3Q19)
    x, y, z = g.grid
    solid = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2 + (z - 0.5 * z.max()) ** 2) < (0.2 * y.max()) ** 2
    g.mask = solid
    gb = [None] + sorted(g.boundaries, key=str)
    ncm_g, nsm_g = build_masks(g, gb, ctx)
    for world in (2, 3):
        for rank in range(world):
            s = SlabObstacle(ctx, res, 50, 0.05, 2, rank=rank, world=world, stencil=lt.D3Q19)
            s.mask = solid                      # the whole-lattice mask is cut to the slab
            assert s.resolution == [16, 8, 12 // world]
            assert s.units.relaxation_parameter_lu == g.units.relaxation_parameter_lu
            assert torch.equal(s.f, g.f[..., s.z0:s.z1])
            for a, b in zip(s.grid, g.grid):
                assert torch.equal(a, b[..., s.z0:s.z1])
            sb = [None] + sorted(s.boundaries, key=str)
            assert [type(b).__name__ for b in sb[1:]] == [type(b).__name__ for b in gb[1:]]
            ncm_s, nsm_s = build_masks(s, sb, ctx)
            assert torch.equal(ncm_s, ncm_g[..., s.z0:s.z1]) and torch.equal(nsm_s, nsm_g[..., s.z0:s.z1])


def test_auto_transport_is_the_handshake_free_one():
    from lettuce_b200.slab import auto_transport
    assert auto_transport([256, 256, 256]) == 'p2p-ll'            # bench.py --gpus N: 256^3 per GPU
    assert auto_transport([1024, 1024, 128]) == 'p2p-ll'          # config 5
    assert auto_transport([2048, 64]) == 'p2p-ll'                 # 2-D: y-slabs, faces are lines


def test_slab_helpers():
    import lettuce_b200 as lt
    from lettuce_b200.slab import slab_range, zface_populations
    assert slab_range(1024, 3, 8) == (384, 512)
    with pytest.raises(ValueError):
        slab_range(10, 0, 4)
    hi, lo = zface_populations(lt.D3Q19())
    assert hi == [5, 7, 10, 11, 14] and lo == [6, 8, 9, 12, 13]
    hi27, _ = zface_populations(lt.D3Q27())
    assert hi27 == [5, 7, 10, 11, 14, 19, 22, 23, 26]
    assert len(zface_populations(lt.D2Q9())[0]) == 3
    L = lt.native.lib()
    assert [L.lb_zface_count(i) for i in range(3)] == [3, 5, 9]


def test_slab_tgv_initial_state_equals_global_slice():
    import lettuce_b200 as lt
    from lettuce_b200.slab import SlabTaylorGreenVortex
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    g = lt.TaylorGreenVortex(ctx, resolution=[12, 10, 16], reynolds_number=100, mach_number=0.05, stencil=lt.D3Q19)
    for world in (2, 4):
        for rank in range(world):
            s = SlabTaylorGreenVortex(ctx, [12, 10, 16], 100, 0.05, lt.D3Q19, rank=rank, world=world)
            assert s.resolution == [12, 10, 16 // world]
            assert s.units.relaxation_parameter_lu == g.units.relaxation_parameter_lu
            assert torch.allclose(s.f, g.f[..., s.z0:s.z1], rtol=0, atol=1e-15)


def _reduce_worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import lettuce_b200 as lt
        from lettuce_b200 import _native as nat
        from lettuce_b200.slab import SlabSimulation
        from oracle import lbm_oracle as orc
        torch.set_num_threads(1)
        S = lt.D2Q9()
        lat = orc.OracleLattice("D2Q9", torch.float64)
        torch.manual_seed(7)
        w = torch.tensor(S.w, dtype=torch.float64).reshape(9, 1, 1)
        f_global = w * (1.0 + 0.1 * torch.rand([9, 6, 4 * world], dtype=torch.float64))

        class LocalFlow:
            pass
        flow = LocalFlow()
        flow.context = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
        flow.stencil, flow.resolution, flow.i = S, [6, 4], 0
        flow.f = f_global[..., 4 * rank:4 * rank + 4].contiguous()
        sim = SlabSimulation(flow, None, [], transport="host", local_stepper=lambda f: f)
        # per-rank partial observables of two report steps, as the step kernel delivers them
        obs = torch.zeros([2, nat.LB_OBS_COUNT], dtype=torch.float64)
        for row, f in enumerate((flow.f, 2 * flow.f)):
            obs[row, nat.LB_OBS_ENERGY] = orc.energy_lu(lat, f)
            obs[row, nat.LB_OBS_MAXU2] = orc.max_velocity_lu(lat, f) ** 2
            obs[row, nat.LB_OBS_MASS] = f.sum()
        keep = obs.clone()
        out = sim._finalize_obs(obs)
        assert torch.equal(obs, keep)                       # the partials are not modified in place
        assert flow._obs_allreduce is not None              # stand-alone observables use the same hook
        if rank == 0:
            e = float(orc.energy_lu(lat, f_global))
            m = float(orc.max_velocity_lu(lat, f_global)) ** 2
            q.put([abs(float(out[0, nat.LB_OBS_ENERGY]) - e) / e, abs(float(out[0, nat.LB_OBS_MAXU2]) - m) / m,
                   abs(float(out[0, nat.LB_OBS_MASS]) - float(f_global.sum())),
                   abs(float(out[1, nat.LB_OBS_MASS]) - 2 * float(f_global.sum()))])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_observables_reduce_globally(world):
    """fused-reporter partials: energy / mass are summed over ranks, max |u|^2 is the maximum"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_reduce_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    errs = q.get(timeout=10)
    assert max(errs) <= 1e-12
