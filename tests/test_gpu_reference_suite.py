"""The reference's OWN tests for this path, restated against lettuce_b200: same flows,
same parameters, same tolerances.  Where the reference compares its native kernel with its CPU torch path
(`tests/native/*`), the CPU side here is the oracle, which is pinned to that torch path.

    reference/tests/native/test_native_{streaming,bgk_collision,bounce_back,equilibrium_pu,no_streaming_mask}.py
    reference/tests/collision/test_collision_{conserves_mass,conserves_momentum,fixpoint_2x,
                                              relaxes_shear_moments,optimizes_pseudo_entropy}.py
    reference/tests/boundary/test_bounceback_bc.py, tests/test_equilibrium.py, tests/test_checkpoint.py,
    reference/tests/reporter/test_generic_reporters.py, tests/flow/test_obstacle.py
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DTYPES = [torch.float32, torch.float64]
STENCILS = ["D2Q9", "D3Q19", "D3Q27"]


def _lt():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import lettuce_b200 as lt
    return lt


def _test_flow(lt, ctx, stencil_name, resolution=16, re=100, ma=0.1):
    """reference/tests/conftest.py:195-232 TestFlow: uniform u = 1.01, p = 0.01 (PU), settable boundaries"""

    class TestFlow(lt.ExtFlow):
        def __init__(self):
            self._boundaries = []
            lt.ExtFlow.__init__(self, ctx, resolution, re, ma, getattr(lt, stencil_name)())

        def make_resolution(self, resolution, stencil=None):
            return [resolution] * stencil.d if isinstance(resolution, int) else resolution

        def make_units(self, reynolds_number, mach_number, resolution):
            return lt.UnitConversion(reynolds_number, mach_number, characteristic_length_lu=resolution[0])

        def initial_pu(self):
            return (0.01 * np.ones([1] + self.resolution), 1.01 * np.ones([self.stencil.d] + self.resolution))

        @property
        def boundaries(self):
            return self._boundaries

        @boundaries.setter
        def boundaries(self, b):
            self._boundaries = b

    return TestFlow()


def _dummy_flow(lt, ctx, init, boundaries=()):
    """reference/tests/conftest.py:243-266 DummyFlow: 16 x 16 D2Q9, Re = Ma = 1, hand-written initial f"""

    class Dummy(lt.ExtFlow):
        def __init__(self):
            lt.ExtFlow.__init__(self, ctx, 16, 1.0, 1.0)

        def make_resolution(self, resolution, stencil=None):
            return [resolution, resolution]

        def make_units(self, reynolds_number, mach_number, _):
            return lt.UnitConversion(reynolds_number=reynolds_number, mach_number=mach_number)

        def initial_pu(self):
            ...

        def initialize(self):
            self.f = torch.zeros([9, 16, 16], dtype=ctx.dtype, device=ctx.device)
            init(self.f)

        @property
        def boundaries(self):
            return list(boundaries)

    return Dummy()


def _conserving_collisions(lt, flow):
    """reference/tests/conftest.py:120-137 fix_conserving_collision"""
    out = [lt.BGKCollision(0.6), lt.TRTCollision(0.6), lt.RegularizedCollision(0.6), lt.SmagorinskyCollision(0.6)]
    if type(flow.stencil).__name__ in ("D2Q9", "D3Q27"):
        out.append(lt.KBCCollision(0.6))
    return out


# ---- tests/native/* --------------------------------------------------------------------------------
def test_native_streaming():
    """test_native_streaming.py:24-51: nine tagged pulses, one NoCollision step, native == CPU"""
    lt = _lt()
    from oracle import lbm_oracle as orc
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)

    def init(f):
        for i in range(9):
            f[i, i + 1, i + 1] = float(i + 1)
    flow = _dummy_flow(lt, ctx, init)
    f0 = flow.f.cpu().clone()
    lt.Simulation(flow, lt.NoCollision(), [])(1)
    ref = orc.run(orc.OracleLattice("D2Q9", torch.float32), f0, 1, "none", {})
    assert torch.equal(flow.f.cpu(), ref)
    e = lt.D2Q9().e
    for i in range(9):   # each pulse moved by its own velocity
        assert flow.f[i, i + 1 + e[i][0], i + 1 + e[i][1]].item() == float(i + 1)


def test_native_bgk_collision():
    """test_native_bgk_collision.py:35-71: f = 1 with a 2.0 spike, tau = 2, one step"""
    lt = _lt()
    from oracle import lbm_oracle as orc
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)

    def init(f):
        f[:, :, :] = 1.0
        f[:, 2, 2] = 2.0
    flow = _dummy_flow(lt, ctx, init)
    f0 = flow.f.cpu().clone()
    lt.Simulation(flow, lt.BGKCollision(2.0), [])(1)
    ref = orc.run(orc.OracleLattice("D2Q9", torch.float32), f0, 1, "bgk", dict(tau=2.0))
    assert flow.f.cpu().numpy() == pytest.approx(ref.numpy())      # rel 1e-6, pytest.approx's default


def test_native_bounce_back():
    """test_native_bounce_back.py:33-74: everything is wall, two NoCollision steps"""
    lt = _lt()
    from oracle import lbm_oracle as orc
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    mask = torch.ones([16, 16], dtype=torch.bool, device="cuda")

    def init(f):
        for i in range(9):
            f[i, :, :] = float(i + 1)
        f[:, 1, 1] = 0.5
    flow = _dummy_flow(lt, ctx, init, boundaries=[lt.BounceBackBoundary(mask)])
    f0 = flow.f.cpu().clone()
    sim = lt.Simulation(flow, lt.NoCollision(), [])
    lat = orc.OracleLattice("D2Q9", torch.float32)
    bnds = [orc.OracleBoundary("bounce_back", mask.cpu())]
    ncm, nsm = orc.build_masks(lat, [16, 16], bnds)
    ref = f0
    for _ in range(2):
        sim(1)
        ref = orc.run(lat, ref, 1, "none", {}, ncm=ncm, nsm=nsm, boundaries=bnds)
        assert torch.equal(flow.f.cpu(), ref)


def test_native_no_streaming_mask():
    """test_native_no_streaming_mask.py:4-22: 64 NoCollision steps leave a uniform field unchanged"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    flow = _test_flow(lt, ctx, "D2Q9", 16, 1, 0.01)
    sim = lt.Simulation(flow, lt.NoCollision(), [])
    sim.no_streaming_mask = ctx.zero_tensor(flow.resolution, dtype=bool)
    f0 = flow.f.clone()
    sim(64)
    assert torch.isclose(f0, flow.f).all()


@pytest.mark.parametrize("dtype", DTYPES)
def test_native_equilibrium_pu(dtype):
    """test_native_equilibrium_pu.py:30-74 / test_equilibrium_bc_pu.py:37-64: one NoCollision step with an
    equilibrium boundary on the column x = 0 sets it to feq(rho(p), u) (rel 1e-6)"""
    lt = _lt()
    from oracle import lbm_oracle as orc
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = _test_flow(lt, ctx, "D2Q9", 16)
    mask = torch.zeros([16, 16], dtype=torch.bool, device="cuda")
    mask[:1, :] = True
    flow.boundaries = [lt.EquilibriumBoundaryPU(ctx, mask, [0.5, -0.25], 0.02)]
    sim = lt.Simulation(flow, lt.NoCollision(), [])
    sim.no_streaming_mask = torch.ones([9, 16, 16], dtype=torch.bool, device="cuda")   # keep: look at the boundary only
    sim(1)
    lat = orc.OracleLattice("D2Q9", dtype)
    rho = flow.units.convert_pressure_pu_to_density_lu(ctx.convert_to_tensor(0.02)).cpu()
    u = flow.units.convert_velocity_to_lu(ctx.convert_to_tensor([0.5, -0.25])).cpu()
    feq = orc.quadratic_equilibrium(lat, rho.reshape(1), u)
    got = flow.f[:, 0, :].cpu().numpy()
    assert got == pytest.approx(np.broadcast_to(feq.numpy()[:, None], got.shape), rel=1e-6)


# ---- tests/collision/* -----------------------------------------------------------------------------
@pytest.mark.parametrize("stencil", STENCILS)
@pytest.mark.parametrize("dtype", DTYPES)
def test_collision_conserves_mass_and_momentum(stencil, dtype):
    """test_collision_conserves_mass.py:4-20, ..._momentum.py:4-22 (abs 1e-5)"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = _test_flow(lt, ctx, stencil)
    for coll in _conserving_collisions(lt, flow):
        rho0, j0 = flow.rho().cpu().numpy(), flow.j().cpu().numpy()
        flow.f = coll(flow)
        assert flow.rho().cpu().numpy() == pytest.approx(rho0, abs=1e-5), type(coll).__name__
        assert flow.j().cpu().numpy() == pytest.approx(j0, abs=1e-5), type(coll).__name__


@pytest.mark.parametrize("stencil", STENCILS)
@pytest.mark.parametrize("dtype", DTYPES)
def test_collision_fixpoint_2x(stencil, dtype):
    """test_collision_fixpoint_2x.py:4-21: BGK with tau = 0.5 applied twice is the identity (abs 1e-5)"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = _test_flow(lt, ctx, stencil)
    g = torch.Generator(device="cuda").manual_seed(1)
    flow.f = flow.f * (1 + 0.1 * torch.rand(flow.f.shape, device="cuda", dtype=dtype, generator=g))
    f_old = flow.f.clone()
    coll = lt.BGKCollision(0.5)
    flow.f = coll(flow)
    flow.f = coll(flow)
    assert f_old.cpu().numpy() == pytest.approx(flow.f.cpu().numpy(), abs=1e-5)


@pytest.mark.parametrize("stencil", STENCILS)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("name", ["BGKCollision", "TRTCollision", "KBCCollision", "RegularizedCollision"])
def test_collision_relaxes_shear_moments(name, stencil, dtype):
    """test_collision_relaxes_shear_moments.py:4-33: Pi_post = Pi - (Pi - Pi_eq) / tau (abs 1e-5).
    (KBC and Regularized read tau from the units, as in the reference; the test's tau = 0.6 only
    parameterises BGK and TRT, and for the uniform TestFlow Pi == Pi_eq, so every tau passes.)"""
    lt = _lt()
    if name == "KBCCollision" and stencil == "D3Q19":
        pytest.skip("KBCCollision only implemented for D2Q9 and D3Q27")
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = _test_flow(lt, ctx, stencil)
    feq = flow.equilibrium(flow)
    shear_pre, shear_eq_pre = flow.shear_tensor(), flow.shear_tensor(feq)
    tau = 0.6
    f_post = getattr(lt, name)(tau)(flow)
    shear_post = flow.shear_tensor(f_post)
    assert shear_post.cpu().numpy() == pytest.approx(
        (shear_pre - 1 / tau * (shear_pre - shear_eq_pre)).cpu().numpy(), abs=1e-5)


@pytest.mark.parametrize("stencil", ["D2Q9", "D3Q27"])
def test_collision_optimizes_pseudo_entropy(stencil):
    """test_collision_optimizes_pseudo_entropy.py:4-29: KBC's pseudo-entropy is above BGK's on the
    np.random.seed(1) field, tau = 0.5003"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    flow = _test_flow(lt, ctx, stencil, resolution=3)
    np.random.seed(1)
    flow.f = ctx.convert_to_tensor(np.random.random([flow.stencil.q] + [3] * flow.stencil.d))
    tau = 0.5003
    kbc = lt.KBCCollision(tau)
    # the reference's KBC ignores its argument and reads tau from the units: give it the same tau
    kbc.tau, kbc.beta = tau, 1. / (2 * tau)
    f_kbc, f_bgk = kbc(flow), lt.BGKCollision(tau)(flow)
    assert (flow.pseudo_entropy_local(f_bgk) < flow.pseudo_entropy_local(f_kbc)).all()


@pytest.mark.parametrize("stencil", STENCILS)
@pytest.mark.parametrize("dtype", DTYPES)
def test_equilibrium_conserves_mass_and_momentum(stencil, dtype):
    """tests/test_equilibrium.py:5-39: sum feq = rho, sum e feq = j (abs 1e-6)"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = _test_flow(lt, ctx, stencil)
    feq = flow.equilibrium(flow)
    assert flow.rho(feq).cpu().numpy() == pytest.approx(flow.rho().cpu().numpy(), abs=1e-6)
    assert flow.j(feq).cpu().numpy() == pytest.approx(flow.j().cpu().numpy(), abs=1e-6)


# ---- boundaries, flows, reporters, checkpoint --------------------------------------------------------
@pytest.mark.parametrize("stencil", STENCILS)
def test_bounce_back_boundary(stencil):
    """test_bounceback_bc.py:6-36: f -> f[opposite] on the mask after one NoCollision collide; an empty
    mask changes nothing"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    flow = _test_flow(lt, ctx, stencil)
    g = torch.Generator(device="cuda").manual_seed(2)
    flow.f = torch.rand(flow.f.shape, device="cuda", dtype=torch.float64, generator=g)
    f_old = flow.f.clone()
    mask = torch.zeros(flow.resolution, dtype=torch.bool, device="cuda")
    flow.boundaries = [lt.BounceBackBoundary(mask)]
    sim = lt.Simulation(flow, lt.NoCollision(), [])
    sim._collide()
    assert torch.equal(flow.f, f_old)
    mask[...] = True
    flow.boundaries = [lt.BounceBackBoundary(mask)]
    sim = lt.Simulation(flow, lt.NoCollision(), [])
    sim._collide()
    assert torch.equal(flow.f, f_old[flow.stencil.opposite])


def test_checkpoint(tmp_path):
    """tests/test_checkpoint.py:5-20: flow.dump -> 10 steps -> flow.load restores f exactly"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=[64, 64], reynolds_number=1, mach_number=0.02, stencil=lt.D2Q9)
    f0 = flow.f.clone()
    filename = str(tmp_path / "tgv")
    flow.dump(filename)
    sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
    sim(10)
    assert not torch.equal(f0, flow.f)
    flow.load(filename)
    assert torch.eq(f0, flow.f).all()
    sim(1)    # and the simulation goes on from the restored state


@pytest.mark.parametrize("observable", ["Enstrophy", "MaximumVelocity", "IncompressibleKineticEnergy", "Mass"])
@pytest.mark.parametrize("case", [[32] * 2, [32] * 3])
@pytest.mark.parametrize("dtype", DTYPES)
def test_generic_reporters(observable, case, dtype):
    """tests/reporter/test_generic_reporters.py:4-25: value after one step within 5 % of the initial value"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=dtype, use_native=True)
    flow = lt.TaylorGreenVortex(ctx, case, 10000, 0.05)
    rep = lt.ObservableReporter(getattr(lt, observable)(flow), interval=1, out=None)
    lt.Simulation(flow, lt.BGKCollision(tau=flow.units.relaxation_parameter_lu), [rep])(2)
    values = np.asarray(rep.out)
    assert values[1, 2] == pytest.approx(values[0, 2], rel=0.05)


@pytest.mark.parametrize("stencil", ["D2Q9", "D3Q27"])
def test_obstacle(stencil):
    """tests/flow/test_obstacle.py:4-32: the stock obstacle (equilibrium inlet, anti-bounce-back outlet,
    bounce-back solid) runs two steps and stays finite"""
    lt = _lt()
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    d = 2 if stencil == "D2Q9" else 3
    res = [32, 16] if d == 2 else [32, 16, 16]
    flow = lt.Obstacle(ctx, res, reynolds_number=100, mach_number=0.1, domain_length_x=2, stencil=getattr(lt, stencil))
    mask = np.zeros(res, dtype=bool)
    mask[tuple([slice(8, 12)] + [slice(6, 10)] * (d - 1))] = True
    flow.mask = mask
    sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
    sim(2)
    assert torch.isfinite(flow.f).all()
    assert sorted(torch.unique(sim.no_collision_mask).tolist()) == [0, 1, 2, 3]
