"""GPU tests of the round-2 kernel work: the persistent TMA-prefetch step kernel against the classic one (bit for
bit), and the refusals / fixes that came out of the round-1 review (forced collision + outlet, enstrophy on axes
shorter than the stencil, Collision.__call__ with an outlet boundary)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lt():
    import gpu_util as gu
    import lettuce_b200 as lt
    return gu, lt


def _collision(lt, flow, name):
    tau = float(flow.units.relaxation_parameter_lu)
    return {"bgk": lambda: lt.BGKCollision(tau), "trt": lambda: lt.TRTCollision(tau, 0.9), "kbc": lambda: lt.KBCCollision(),
            "smagorinsky": lambda: lt.SmagorinskyCollision(tau), "regularized": lambda: lt.RegularizedCollision(),
            "none": lambda: lt.NoCollision()}[name]()


@pytest.mark.parametrize("stencil,dtype,res,collision", [
    ("D3Q27", torch.float32, [64, 64, 128], "kbc"), ("D3Q27", torch.float32, [64, 64, 128], "trt"),
    ("D3Q19", torch.float32, [64, 96, 128], "bgk"), ("D3Q19", torch.float64, [64, 64, 64], "bgk"),
    ("D3Q27", torch.float64, [48, 64, 64], "kbc"), ("D2Q9", torch.float32, [1024, 512], "bgk"),
    ("D3Q19", torch.float32, [64, 64, 128], "none"), ("D3Q27", torch.float32, [64, 64, 128], "smagorinsky"),
])
def test_tma_kernel_equals_classic_kernel(stencil, dtype, res, collision):
    """lb_desc.kernel_flags: LB_KF_FORCE_TMA (persistent CTAs, populations through cp.async.bulk + mbarrier) and
    LB_KF_NO_TMA (one tile per CTA, direct loads) give the same populations bit for bit, and the same fused
    energy / max-velocity series up to the order of the partial sums"""
    gu, lt = _lt()
    nat = lt.native
    out = {}
    for flag in (nat.LB_KF_NO_TMA, nat.LB_KF_FORCE_TMA):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=800, mach_number=0.1, stencil=getattr(lt, stencil))
        reps = [lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None),
                lt.ObservableReporter(lt.MaximumVelocity(flow), interval=2, out=None)]
        sim = lt.Simulation(flow, _collision(lt, flow, collision), reps)
        sim.kernel_flags = flag | nat.LB_KF_NO_LOOP
        sim(5)
        plain = lt.Simulation(flow, _collision(lt, flow, collision), [])     # and without the fused reductions
        plain.kernel_flags = flag | nat.LB_KF_NO_LOOP
        plain(3)
        out[flag] = (flow.f.clone(), [r[2] for r in reps[0].out], [r[2] for r in reps[1].out])
    a, b = out[nat.LB_KF_NO_TMA], out[nat.LB_KF_FORCE_TMA]
    assert torch.equal(a[0], b[0])
    tol = 1e-6 if dtype == torch.float32 else 1e-13
    assert a[1] == pytest.approx(b[1], rel=tol) and a[2] == pytest.approx(b[2], rel=tol)


def test_forced_collision_with_outlet_is_refused():
    gu, lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    flow = lt.Obstacle(ctx, [32, 16], reynolds_number=50, mach_number=0.05, domain_length_x=1)
    force = lt.Guo(flow, 0.8, [1e-6, 0.0])
    with pytest.raises(NotImplementedError, match="forcing scheme"):
        lt.Simulation(flow, lt.BGKCollision(0.8, force=force), [])
    # and the C ABI itself
    nat = lt.native
    import ctypes as C
    desc = nat.make_desc("D2Q9", torch.float64, [32, 16], nat.LB_COLL_BGK, 0.8, force=nat.LB_FORCE_GUO,
                         acceleration=[1e-6, 0.0], force_tau=0.8)
    f = torch.ones([9, 32, 16], dtype=torch.float64, device="cuda") / 9
    g = torch.empty_like(f)
    ncm = torch.zeros([32, 16], dtype=torch.uint8, device="cuda")
    bnd = (nat.LbBoundary * 1)()
    bnd[0].kind, bnd[0].direction = nat.LB_BND_ABB_OUTLET, 1
    fl = nat.LbFields()
    fl.f, fl.f_next, fl.no_collision_mask = f.data_ptr(), g.data_ptr(), ncm.data_ptr()
    fl.boundaries, fl.n_boundaries = C.cast(bnd, C.POINTER(nat.LbBoundary)), 1
    assert nat.lib().lb_step(desc, fl, 1, None) == -3          # LB_EUNSUPPORTED
    assert b"forcing scheme" in nat.lib().lb_last_error()


def test_enstrophy_needs_three_nodes_per_axis():
    gu, lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    f0 = torch.rand([9, 16, 2], dtype=torch.float64) * 0.1 + 0.1
    flow = gu.FieldFlow(ctx, [16, 2], lt.D2Q9, f0)
    with pytest.raises(lt.native.NativeError, match="at least 3 nodes"):
        lt.Enstrophy(flow)(flow.f)
    rep = lt.ObservableReporter(lt.Enstrophy(flow), interval=1, out=None)
    flow.i = 1                                           # skip the initial (standalone) report
    with pytest.raises(NotImplementedError, match="at least 3 nodes"):
        lt.Simulation(flow, lt.BGKCollision(0.8), [rep])(2)


def test_collision_call_with_outlet_boundary_is_deterministic_and_right():
    """Simulation._collide (Collision.__call__ + boundaries, no streaming) with the stock obstacle list: the outlet
    nodes read their neighbours' populations, so the call runs out of place -- against the oracle's collide"""
    from oracle import lbm_oracle as orc
    gu, lt = _lt()
    res = [40, 24]
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    flow = lt.Obstacle(ctx, res, reynolds_number=100, mach_number=0.05, domain_length_x=1)
    x, y = flow.grid
    cyl = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.12 * y.max()) ** 2
    flow.mask = cyl
    g = torch.Generator(device="cuda").manual_seed(5)
    flow.f = flow.f * (1.0 + 0.05 * torch.rand(flow.f.shape, device="cuda", dtype=torch.float64, generator=g))
    f0 = flow.f.cpu().clone()
    sim = lt.Simulation(flow, lt.BGKCollision(tau=0.7), [])
    eq = sim.boundaries[3]
    rho_b, u_b = eq._lattice_units(flow)
    inlet = torch.zeros(res, dtype=torch.bool)
    inlet[0] = True
    lat = orc.OracleLattice("D2Q9", torch.float64)
    obnds = [orc.OracleBoundary("abb_outlet", direction=[1, 0]), orc.OracleBoundary("bounce_back", cyl.cpu()),
             orc.OracleBoundary("equilibrium", inlet, rho=rho_b.cpu().double(), u=u_b.cpu().double())]
    ncm, _ = orc.build_masks(lat, res, obnds)
    ref = orc.collide(lat, f0, "bgk", dict(tau=0.7), ncm=ncm, boundaries=obnds)
    got = sim._collide().cpu()
    assert float((got - ref).abs().max() / ref.abs().max()) <= 1e-12
    flow.f = f0.to("cuda")
    assert torch.equal(sim._collide().cpu(), got)       # and again: the same bits


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("dims", [2, 3])
def test_tma_kernel_with_no_collision_mask(dtype, dims):
    """bounce-back solid + two equilibrium boundaries (no masked streaming): the persistent TMA kernel brings the
    mask bytes of a tile in with the populations -- same bits as the classic kernel"""
    gu, lt = _lt()
    nat = lt.native
    res = [1024, 512] if dims == 2 else [64, 64, 128]
    stencil = lt.D2Q9 if dims == 2 else lt.D3Q19
    out = []
    for flag in (nat.LB_KF_NO_TMA, nat.LB_KF_FORCE_TMA):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)

        class LocalObstacle(lt.Obstacle):
            @property
            def boundaries(self):
                x = self.grid[0]
                u_in = self.units.characteristic_velocity_pu * self._unit_vector()
                return [lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                        lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x - x.max()) < 1e-6, velocity=u_in),
                        lt.BounceBackBoundary(self.mask)]
        flow = LocalObstacle(ctx, res, reynolds_number=200, mach_number=0.05, domain_length_x=1, stencil=stencil)
        g = flow.grid
        flow.mask = sum((c - 0.4 * c.max()) ** 2 for c in g) < (0.1 * g[1].max()) ** 2
        rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=max(float(flow.units.relaxation_parameter_lu), 0.55)), [rep])
        sim.kernel_flags = flag | nat.LB_KF_NO_LOOP
        sim(6)
        out.append((flow.f.clone(), [r[2] for r in rep.out]))
    assert torch.equal(out[0][0], out[1][0]) and bool(torch.isfinite(out[0][0]).all())
    assert out[0][1] == pytest.approx(out[1][1], rel=1e-6 if dtype == torch.float32 else 1e-13)


@pytest.mark.parametrize("case", ["plain", "ncm", "sbits", "tma", "halo_ll"])
def test_every_output_element_is_written(case):
    """the push scheme writes every (q, node) of the output exactly once: pre-fill the output buffer with NaN, step
    once, no NaN may be left (compute-sanitizer is closed on this pool: this is the write-coverage check)"""
    import ctypes as C
    gu, lt = _lt()
    nat = lt.native
    L = nat.lib()
    res = [64, 64, 128] if case == "tma" else [12, 10, 24]
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    if case in ("ncm", "sbits"):
        flow = lt.Obstacle(ctx, res, reynolds_number=100, mach_number=0.05, domain_length_x=1, stencil=lt.D3Q19)
        g = flow.grid
        flow.mask = sum((c - 0.4 * c.max()) ** 2 for c in g) < (0.15 * g[1].max()) ** 2
        if case == "ncm":
            flow.__class__ = type("B", (lt.Obstacle,), {"boundaries": property(lambda self: [lt.BounceBackBoundary(self.mask)])})
    else:
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=400, mach_number=0.1, stencil=lt.D3Q19)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=0.7), [])
    sim.kernel_flags = nat.LB_KF_FORCE_TMA if case == "tma" else nat.LB_KF_NO_TMA
    flow.f_next.fill_(float("nan"))
    if case != "halo_ll":
        sim._collide_and_stream(sim)
        assert bool(torch.isfinite(flow.f).all())
        return
    desc = sim._desc()
    nb = L.lb_halo_ll_bytes(desc)
    lo = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
    hi = [torch.zeros([nb], dtype=torch.uint8, device="cuda") for _ in range(3)]
    flags = torch.zeros([8], dtype=torch.int32, device="cuda")
    h = nat.LbHalo()
    h.mode = nat.LB_HALO_LL
    for k in range(3):
        h.recv_lo[k], h.recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
        h.peer_up_recv_lo[k], h.peer_down_recv_hi[k] = lo[k].data_ptr(), hi[k].data_ptr()
    h.wait_slots = flags.data_ptr()
    A, B = flow.f.clone(), torch.full_like(flow.f, float("nan"))
    fl = nat.LbFields()
    fl.f, fl.f_next = A.data_ptr(), B.data_ptr()
    fl.halo = C.pointer(h)
    nat.check(L.lb_step(desc, fl, 1, torch.cuda.current_stream().cuda_stream))
    assert bool(torch.isfinite(B).all()) and int(flags[2].item()) == 0


@pytest.mark.parametrize("stencil,dtype,res,collision", [
    ("D3Q27", torch.float32, [64, 64, 128], "kbc"), ("D3Q27", torch.float32, [64, 64, 128], "trt"),
    ("D3Q27", torch.float32, [64, 64, 128], "bgk"), ("D3Q27", torch.float32, [64, 64, 128], "smagorinsky"),
    ("D3Q27", torch.float32, [64, 64, 128], "regularized"), ("D3Q27", torch.float64, [48, 64, 64], "bgk"),
    ("D3Q27", torch.float32, [37, 64, 256], "kbc"), ("D3Q27", torch.float64, [48, 64, 64], "kbc"),
    ("D3Q27", torch.float32, [64, 64, 65], "kbc"), ("D3Q27", torch.float32, [66, 64, 66], "trt"), ("D2Q9", torch.float32, [1024, 512], "kbc"),
    ("D2Q9", torch.float64, [999, 512], "bgk"), ("D2Q9", torch.float32, [1024, 512], "none"),
])
def test_prefetch_kernel_equals_classic_kernel(stencil, dtype, res, collision):
    """LB_KF_FORCE_PREFETCH (every thread copies its own part of the CTA's next tile into shared memory with cp.async
    while it works on the current one; a CTA takes consecutive tiles) against the classic kernel: the same
    populations bit for bit, the same fused energy / max-velocity series up to the order of the partial sums.
    D3Q27 and D2Q9 (where it is compiled), vector and one-node-per-thread widths (D3Q27 KBC fp64, odd line
    lengths), tile counts that are not a multiple of the tiles per CTA."""
    gu, lt = _lt()
    nat = lt.native
    out = {}
    for flag in (nat.LB_KF_NO_PREFETCH | nat.LB_KF_NO_TMA, nat.LB_KF_FORCE_PREFETCH):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=800, mach_number=0.1, stencil=getattr(lt, stencil))
        reps = [lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None),
                lt.ObservableReporter(lt.MaximumVelocity(flow), interval=2, out=None)]
        sim = lt.Simulation(flow, _collision(lt, flow, collision), reps)
        sim.kernel_flags = flag | nat.LB_KF_NO_LOOP
        sim(5)
        plain = lt.Simulation(flow, _collision(lt, flow, collision), [])     # and without the fused reductions
        plain.kernel_flags = flag | nat.LB_KF_NO_LOOP
        plain(4)
        out[flag] = (flow.f.clone(), [r[2] for r in reps[0].out], [r[2] for r in reps[1].out])
    a, b = out[nat.LB_KF_NO_PREFETCH | nat.LB_KF_NO_TMA], out[nat.LB_KF_FORCE_PREFETCH]
    assert torch.equal(a[0], b[0]) and bool(torch.isfinite(a[0]).all())
    tol = 1e-6 if dtype == torch.float32 else 1e-13
    assert a[1] == pytest.approx(b[1], rel=tol) and a[2] == pytest.approx(b[2], rel=tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64], ids=["f32", "f64"])
def test_prefetch_kernel_with_no_collision_mask(dtype):
    """bounce-back solid + two equilibrium boundaries on D3Q27: the prefetch kernel loads a tile's mask word one tile
    ahead into a register -- same bits as the classic kernel; and every output element is written (NaN pre-fill)"""
    gu, lt = _lt()
    nat = lt.native
    out = []
    for flag in (nat.LB_KF_NO_PREFETCH | nat.LB_KF_NO_TMA, nat.LB_KF_FORCE_PREFETCH):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)

        class LocalObstacle(lt.Obstacle):
            @property
            def boundaries(self):
                x = self.grid[0]
                u_in = self.units.characteristic_velocity_pu * self._unit_vector()
                return [lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                        lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x - x.max()) < 1e-6, velocity=u_in),
                        lt.BounceBackBoundary(self.mask)]
        flow = LocalObstacle(ctx, [64, 64, 128], reynolds_number=200, mach_number=0.05, domain_length_x=1, stencil=lt.D3Q27)
        g = flow.grid
        flow.mask = sum((c - 0.4 * c.max()) ** 2 for c in g) < (0.1 * g[1].max()) ** 2
        rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None)
        sim = lt.Simulation(flow, lt.BGKCollision(tau=max(float(flow.units.relaxation_parameter_lu), 0.55)), [rep])
        sim.kernel_flags = flag | nat.LB_KF_NO_LOOP
        sim(5)
        flow.f_next.fill_(float("nan"))
        sim._collide_and_stream(sim)
        out.append((flow.f.clone(), [r[2] for r in rep.out]))
    assert torch.equal(out[0][0], out[1][0]) and bool(torch.isfinite(out[0][0]).all())
    assert out[0][1] == pytest.approx(out[1][1], rel=1e-6 if dtype == torch.float32 else 1e-13)


@pytest.mark.parametrize("dtype,res,collision,steps", [
    (torch.float64, [256, 256], "bgk", 50), (torch.float64, [256, 256], "bgk", 37), (torch.float32, [256, 256], "bgk", 64),
    (torch.float64, [128, 128], "kbc", 21), (torch.float32, [96, 160], "trt", 30), (torch.float64, [100, 36], "bgk", 2),
    (torch.float64, [100, 36], "bgk", 3), (torch.float32, [64, 64], "none", 70), (torch.float64, [97, 64], "regularized", 9),
    (torch.float32, [16, 16], "bgk", 25), (torch.float64, [384, 320], "bgk", 12), (torch.float32, [256, 256], "smagorinsky", 11),
])
def test_tile_resident_kernel_equals_one_launch_per_step(dtype, res, collision, steps):
    """small 2-D lattices: the multi-step call keeps tiles of the lattice in shared memory for h steps per grid
    barrier (lb_resident.cuh; halo nodes recomputed by the neighbouring CTAs).  Same populations, bit for bit, as one
    launch per step (LB_KF_NO_LOOP) and as the cooperative loop kernel (LB_KF_NO_RESIDENT); two calls in a row (the
    state must end in the buffer the host expects for odd and even step counts)."""
    gu, lt = _lt()
    nat = lt.native
    out = []
    for flag in (nat.LB_KF_NO_LOOP | nat.LB_KF_NO_TMA, nat.LB_KF_NO_RESIDENT, 0):
        ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
        flow = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
        sim = lt.Simulation(flow, _collision(lt, flow, collision), [])
        sim.kernel_flags = flag
        sim(steps)
        a = flow.f.clone()
        sim(steps + 1)
        out.append((a, flow.f.clone(), flow.i))
    for o in out[1:]:
        assert torch.equal(o[0], out[0][0]) and torch.equal(o[1], out[0][1]) and o[2] == out[0][2]
    assert bool(torch.isfinite(out[0][1]).all())

