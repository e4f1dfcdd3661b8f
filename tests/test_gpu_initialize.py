"""lb_initialize: `Flow.initialize` + `initialize_f_neq` on the device
(reference lettuce/_flow.py:106-122, 309-336; gradients lettuce/util/utility.py:37-99).
  * against the initial populations the UNMODIFIED reference produced (f0 of the tgv_* golden files),
  * against this package's torch path (what a CPU context runs), incl. an equilibrium-only flow,
  * slab by slab with ghost planes == the matching slice of the undivided initial state, bit for bit."""
import pytest
import torch

from golden_util import DT, TOL, golden_names, load_case, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("tgv_") and "_bgk_" in n])
def test_initial_populations_vs_reference(name):
    import lettuce_b200 as lt
    case = load_case(name)
    meta = case["meta"]
    ctx = lt.Context(device="cuda", dtype=DT[meta["dtype"]], use_native=True)
    flow = lt.TaylorGreenVortex(ctx, resolution=list(meta["res"]), reynolds_number=100.0, mach_number=0.05,
                                stencil=getattr(lt, meta["stencil"]))     # tests/golden/make_golden.py tgv_field
    assert flow.f.shape == case["f0"].shape
    assert rel_err(flow.f.cpu().numpy(), case["f0"]) <= TOL[meta["dtype"]]


@pytest.mark.parametrize("stencil,res", [("D2Q9", [20, 12]), ("D3Q19", [8, 6, 10]), ("D3Q27", [5, 9, 7])])
@pytest.mark.parametrize("dtype", ["f32", "f64"])
def test_native_initialise_equals_torch_path(stencil, res, dtype):
    import lettuce_b200 as lt
    dev = lt.TaylorGreenVortex(lt.Context(device="cuda", dtype=DT[dtype], use_native=True), resolution=res,
                               reynolds_number=400, mach_number=0.1, stencil=getattr(lt, stencil))
    cpu = lt.TaylorGreenVortex(lt.Context(device="cpu", dtype=DT[dtype], use_native=False), resolution=res,
                               reynolds_number=400, mach_number=0.1, stencil=getattr(lt, stencil))
    err = rel_err(dev.f.cpu().numpy(), cpu.f.numpy())
    print(f"initialise {stencil} {dtype} {res}: {err:.3e}")
    assert err <= (2e-6 if dtype == "f32" else 1e-13)


def test_equilibrium_only_initialisation_obstacle():
    import lettuce_b200 as lt
    outs = []
    for device, native in (("cuda", True), ("cpu", False)):
        ctx = lt.Context(device=device, dtype=torch.float64, use_native=native)
        flow = lt.Obstacle(ctx, [40, 24], reynolds_number=100, mach_number=0.05, domain_length_x=1)
        outs.append(flow.f.cpu().numpy())
    assert rel_err(outs[0], outs[1]) <= 1e-13


@pytest.mark.parametrize("stencil,res", [("D3Q19", [12, 10, 16]), ("D2Q9", [14, 24])])
@pytest.mark.parametrize("world", [2, 4])
def test_slab_initialisation_is_the_global_slice(stencil, res, world):
    import lettuce_b200 as lt
    from lettuce_b200.slab import SlabTaylorGreenVortex
    ctx = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    g = lt.TaylorGreenVortex(ctx, resolution=res, reynolds_number=100, mach_number=0.05, stencil=getattr(lt, stencil))
    parts = [SlabTaylorGreenVortex(ctx, res, 100, 0.05, getattr(lt, stencil), rank=r, world=world).f
             for r in range(world)]
    assert torch.equal(torch.cat(parts, dim=-1), g.f)
