"""GPU parity at BASELINE.json's OWN sizes, against the CPU oracle on identical initial
populations -- the cases the small seeded shapes of test_gpu_parity.py cannot stand in for:

  config 1  TaylorGreenVortex2D D2Q9 BGK fp64 256 x 256, 1000 steps (examples/00_simplest_TGV.py:9-21), <= 1e-12
  config 2  TaylorGreenVortex3D D3Q19 BGK fp32 256^3, 3 steps, <= 1e-5
  config 3  Obstacle2D D2Q9 BGK 2048 x 1024, fp32 and fp64, with the all-local boundary list (two equilibrium
            columns + bounce-back cylinder) and with the stock list of lt.Obstacle (equilibrium inlet,
            AntiBounceBackOutlet, bounce-back; obstacle.py:107-122), 8 steps
  config 4  TaylorGreenVortex3D D3Q27 KBC fp32 at 128^3 with the every-step fused energy reporter against the
            oracle's energy of every state (512^3 itself is 14.5 GB of populations per copy: the oracle cannot
            hold it in seconds; interval independence at that size is a bench.py extra)

Tolerance: rel = max|a - b| / max|b| <= 1e-12 (fp64) / 1e-5 (fp32), the bars of BASELINE.json's north_star."""
import numpy as np
import pytest
import torch

from golden_util import rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float32: 1e-5, torch.float64: 1e-12}


def _mods():
    import gpu_util as gu
    import lettuce_b200 as lt
    from oracle import lbm_oracle as orc
    torch.set_num_threads(max(1, torch.get_num_threads()))
    return gu, lt, orc


def test_config1_tgv2d_d2q9_bgk_fp64_256_1000_steps():
    gu, lt, orc = _mods()
    cpu = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    ref_flow = lt.TaylorGreenVortex(cpu, resolution=256, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    f0 = ref_flow.f.clone()                      # torch-path initial state (pinned to the reference by tgv_* goldens)
    tau = ref_flow.units.relaxation_parameter_lu
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=256, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    assert rel_err(flow.f.cpu().numpy(), f0.numpy()) <= 1e-14          # lb_initialize vs the torch initialisation
    flow.f = f0.to("cuda")                                            # identical initial populations, byte for byte
    sim = lt.Simulation(flow, lt.BGKCollision(tau=tau), [])
    mlups = sim(1000)
    assert flow.i == 1000 and mlups > 0
    lat = orc.OracleLattice("D2Q9", torch.float64)
    ref = orc.run(lat, f0, 1000, "bgk", dict(tau=tau))
    err = rel_err(flow.f.cpu().numpy(), ref.numpy())
    print(f"config 1: rel err after 1000 steps {err:.3e}")
    assert err <= 1e-12
    # the batched call (one cooperative multi-step launch) == 1000 calls of the per-step seam, bit for bit
    flow2 = lt.TaylorGreenVortex(ctx, resolution=256, reynolds_number=100, mach_number=0.05, stencil=lt.D2Q9)
    flow2.f = f0.to("cuda")
    sim2 = lt.Simulation(flow2, lt.BGKCollision(tau=tau), [])
    for _ in range(1000):
        sim2._collide_and_stream(sim2)
    assert torch.equal(flow2.f, flow.f)


def test_config2_tgv3d_d3q19_bgk_fp32_256_3_steps():
    gu, lt, orc = _mods()
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=256, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q19)
    f0 = flow.f.cpu().clone()
    tau = flow.units.relaxation_parameter_lu
    lt.Simulation(flow, lt.BGKCollision(tau=tau), [])(3)
    lat = orc.OracleLattice("D3Q19", torch.float32)
    ref = orc.run(lat, f0, 3, "bgk", dict(tau=tau))
    err = rel_err(flow.f.cpu().numpy(), ref.numpy())
    print(f"config 2: rel err after 3 steps at 256^3 {err:.3e}")
    assert err <= 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("stock", [False, True], ids=["all-local", "stock-abb-outlet"])
def test_config3_obstacle2d_2048x1024(dtype, stock):
    gu, lt, orc = _mods()
    res = [2048, 1024]
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)

    class LocalObstacle(lt.Obstacle):
        @property
        def boundaries(self):
            x = self.grid[0]
            u_in = self.units.characteristic_velocity_pu * self._unit_vector()
            return [lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x) < 1e-6, velocity=u_in),
                    lt.EquilibriumBoundaryPU(context=self.context, mask=torch.abs(x - x.max()) < 1e-6, velocity=u_in),
                    lt.BounceBackBoundary(self.mask)]

    flow = (lt.Obstacle if stock else LocalObstacle)(ctx, res, reynolds_number=300, mach_number=0.01, domain_length_x=1)
    x, y = flow.grid
    cyl = ((x - 0.3 * x.max()) ** 2 + (y - 0.5 * y.max()) ** 2) < (0.05 * y.max()) ** 2
    flow.mask = cyl
    # a perturbed start, so that the 8 steps carry signal through the cylinder and both end columns
    g = torch.Generator(device="cuda").manual_seed(11)
    flow.f = flow.f * (1.0 + 0.02 * torch.rand(flow.f.shape, device="cuda", dtype=dtype, generator=g))
    f0 = flow.f.cpu().clone()
    tau = float(flow.units.relaxation_parameter_lu)
    sim = lt.Simulation(flow, lt.BGKCollision(tau=tau), [])
    kinds = [type(b).__name__ for b in sim.boundaries[1:]]
    lat = orc.OracleLattice("D2Q9", dtype)
    inlet = torch.zeros(res, dtype=torch.bool)
    inlet[0] = True
    outlet = torch.zeros(res, dtype=torch.bool)
    outlet[-1] = True
    eqs = [b for b in sim.boundaries[1:] if isinstance(b, lt.EquilibriumBoundaryPU)]
    rho_b, u_b = eqs[0]._lattice_units(flow)
    rho_b, u_b = rho_b.cpu().to(dtype), u_b.cpu().to(dtype)
    if stock:
        assert kinds == ["AntiBounceBackOutlet", "BounceBackBoundary", "EquilibriumBoundaryPU"]
        obnds = [orc.OracleBoundary("abb_outlet", direction=[1, 0]), orc.OracleBoundary("bounce_back", cyl.cpu()),
                 orc.OracleBoundary("equilibrium", inlet, rho=rho_b, u=u_b)]
    else:
        assert kinds == ["BounceBackBoundary", "EquilibriumBoundaryPU", "EquilibriumBoundaryPU"]
        # the two equilibrium boundaries are ordered by object address (sorted(key=str)): take the masks as built
        m1 = sim.boundaries[2]._mask.cpu()
        m2 = sim.boundaries[3]._mask.cpu()
        obnds = [orc.OracleBoundary("bounce_back", cyl.cpu()), orc.OracleBoundary("equilibrium", m1, rho=rho_b, u=u_b),
                 orc.OracleBoundary("equilibrium", m2, rho=rho_b, u=u_b)]
        assert bool((m1 | m2).eq(inlet | outlet).all())
    ncm, nsm = orc.build_masks(lat, res, obnds)
    assert torch.equal(ncm, sim.no_collision_mask.cpu())
    if stock:
        assert torch.equal(nsm.to(torch.uint8), sim.no_streaming_mask.cpu().to(torch.uint8))
    sim(8)
    ref = orc.run(lat, f0, 8, "bgk", dict(tau=tau), ncm=ncm, nsm=nsm, boundaries=obnds)
    err = rel_err(flow.f.cpu().numpy(), ref.numpy())
    print(f"config 3 ({'stock' if stock else 'all-local'}, {dtype}): rel err after 8 steps {err:.3e}")
    assert err <= TOL[dtype]


def test_config4_tgv3d_d3q27_kbc_fp32_128_energy_every_step():
    gu, lt, orc = _mods()
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=128, reynolds_number=1600, mach_number=0.1, stencil=lt.D3Q27)
    f0 = flow.f.cpu().clone()
    rep = lt.ObservableReporter(lt.IncompressibleKineticEnergy(flow), interval=1, out=None)
    n = 6
    lt.Simulation(flow, lt.KBCCollision(), [rep])(n)
    tau = flow.units.relaxation_parameter_lu
    lat = orc.OracleLattice("D3Q27", torch.float32)
    scale = lt.IncompressibleKineticEnergy(flow).from_lattice(1.0)
    f = f0
    assert [r[0] for r in rep.out] == list(range(n + 1))
    for i in range(n + 1):
        assert rep.out[i][2] == pytest.approx(float(orc.energy_lu(lat, f)) * scale, rel=2e-5)
        if i < n:
            f = orc.step(lat, f, "kbc", dict(tau=tau))
    err = rel_err(flow.f.cpu().numpy(), f.numpy())
    print(f"config 4 (128^3): rel err after {n} steps {err:.3e}")
    assert err <= 1e-5
