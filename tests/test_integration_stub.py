"""The reference-side binding (integration/b200.py, the file INTEGRATION.md tells a lettuce maintainer to
drop in as lettuce/cuda_native/b200.py): its host logic turns a lettuce `Simulation` into the C structures
of include/lettuce_b200.h.  Checked without a GPU
  * against the ctypes layout of lettuce_b200's own binding,
  * with lettuce_b200's mirror objects standing in for lettuce's (same attribute names), and
  * with the UNMODIFIED reference's objects (build container only: /root/reference), for which the
    descriptor, the boundary records and the mask pointers must equal what lettuce_b200's own
    `Simulation` plumbing builds for the same configuration."""
import ctypes as C
import importlib.util
import os
import types

import numpy as np
import pytest
import torch

import lettuce_b200 as lt
from lettuce_b200 import _native as nat
from oracle.reference_shim import import_reference, reference_available

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200_stub", os.path.join(ROOT, "integration", "b200.py"))
b200 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b200)


def test_struct_layouts_match_the_binding():
    for mine, theirs in ((b200.Desc, nat.LbDesc), (b200.Boundary, nat.LbBoundary), (b200.Fields, nat.LbFields)):
        assert C.sizeof(mine) == C.sizeof(theirs)
        for (name, _), (name2, _) in zip(mine._fields_, theirs._fields_):
            assert name == name2 and getattr(mine, name).offset == getattr(theirs, name2).offset


def _fake_simulation(L, flow, collision):
    """the attributes of a lettuce Simulation the stub reads, built as _simulation.py:57-86 does"""
    bnds = [None] + sorted(flow.boundaries, key=lambda b: str(b))
    ncm, nsm = lt.build_masks(flow, bnds, flow.context) if L is lt else (None, None)
    return types.SimpleNamespace(flow=flow, collision=collision, boundaries=bnds, context=flow.context,
                                 no_collision_mask=ncm, no_streaming_mask=nsm)


def _records(arr, n):
    return [(arr[k].kind, arr[k].direction, arr[k].value, list(arr[k].rho_stride), list(arr[k].u_stride),
             arr[k].u_comp_stride) for k in range(n)]


def _cases(L):
    ctx = L.Context(device="cpu", dtype=torch.float64, use_native=False)
    ctx32 = L.Context(device="cpu", dtype=torch.float32, use_native=False)
    cavity_cls = L.Cavity2D if hasattr(L, "Cavity2D") else None
    if cavity_cls is None:                      # the reference's module has no __all__ (liddrivencavity.py)
        from lettuce.ext._flows.liddrivencavity import Cavity2D as cavity_cls
    out = []
    f = L.TaylorGreenVortex(ctx32, [16, 16, 16], 1600, 0.1, stencil=L.D3Q19)
    out.append(("tgv_bgk", f, L.BGKCollision(tau=f.units.relaxation_parameter_lu)))
    f = L.TaylorGreenVortex(ctx, [12, 12], 100, 0.05, stencil=L.D2Q9)
    out.append(("tgv_trt", f, L.TRTCollision(tau=0.7, tau_minus=0.9)))
    out.append(("tgv_kbc", f, L.KBCCollision()))
    out.append(("tgv_none", f, L.NoCollision()))
    out.append(("tgv_reg", f, L.RegularizedCollision(tau=123.0)))
    out.append(("tgv_smag", f, L.SmagorinskyCollision(tau=0.6, smagorinsky_constant=0.2)))
    f = L.PoiseuilleFlow2D(ctx, 16, 10, 0.05)
    tau = f.units.relaxation_parameter_lu
    out.append(("poiseuille_guo", f, L.BGKCollision(tau=tau, force=L.Guo(f, tau, f.acceleration))))
    out.append(("poiseuille_shanchen", f, L.BGKCollision(tau=tau, force=L.ShanChen(f, 0.8, [1e-3, 2e-3]))))
    f = cavity_cls(ctx, [20, 16], 100, 0.05)
    out.append(("cavity", f, L.BGKCollision(tau=f.units.relaxation_parameter_lu)))
    f = L.Obstacle(ctx32, [24, 12, 8], reynolds_number=50, mach_number=0.05, domain_length_x=2, stencil=L.D3Q27)
    m = np.zeros([24, 12, 8], dtype=bool)
    m[6:9, 4:8, 2:6] = True
    f.mask = m
    out.append(("obstacle_stock", f, L.KBCCollision()))
    return out


def _check_against_mirror(name, sim, mirror_sim):
    """stub(sim) == lettuce_b200's own plumbing(mirror_sim)"""
    d, fl, keep = b200.build_call(sim, pack=False)
    want = lt.Simulation._desc(mirror_sim)
    assert bytes(d) == bytes(want), name
    assert fl.f == sim.flow.f.data_ptr() and fl.f_next == sim.flow.f_next.data_ptr()
    if mirror_sim.no_collision_mask is None:
        assert sim.no_collision_mask is None and fl.n_boundaries == 0 and not fl.no_collision_mask
        return
    assert fl.no_collision_mask == sim.no_collision_mask.data_ptr()
    arr, n, _ = lt.Simulation._boundary_array(mirror_sim)
    assert fl.n_boundaries == n
    assert _records(fl.boundaries, n) == _records(arr, n), name
    # equilibrium records point at the same lattice-unit values
    for k in range(n):
        if arr[k].kind == nat.LB_BND_EQUILIBRIUM:
            b = mirror_sim.boundaries[1 + k]
            rho, u = b._lattice_units(mirror_sim.flow)
            T = C.c_float if sim.flow.f.dtype == torch.float32 else C.c_double
            got_rho = C.cast(fl.boundaries[k].rho, C.POINTER(T))[0]
            got_u = [C.cast(fl.boundaries[k].u, C.POINTER(T))[i] for i in range(u.numel())]
            assert got_rho == pytest.approx(float(rho.reshape(-1)[0]), rel=1e-6)
            assert got_u == pytest.approx([float(v) for v in u.reshape(-1)], rel=1e-6, abs=1e-12)


def test_stub_on_mirror_objects():
    """lettuce_b200's classes carry the attribute names the stub reads from lettuce's"""
    for name, flow, coll in _cases(lt):
        sim = _fake_simulation(lt, flow, coll)
        _check_against_mirror(name, sim, sim)


@pytest.mark.skipif(not reference_available(), reason="needs the reference checkout (build container only)")
def test_stub_on_the_unmodified_reference():
    ref = import_reference()
    mirror = {name: (flow, coll) for name, flow, coll in _cases(lt)}
    for name, flow, coll in _cases(ref):
        sim = ref.Simulation(flow, coll, [])            # the reference builds its own masks and boundary order
        mflow, mcoll = mirror[name]
        msim = _fake_simulation(lt, mflow, mcoll)
        if sim.no_collision_mask is not None:
            assert torch.equal(sim.no_collision_mask, msim.no_collision_mask), name
            assert torch.equal(sim.no_streaming_mask, msim.no_streaming_mask), name
        _check_against_mirror(name, sim, msim)


def test_outlet_direction_from_plane_index():
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, [8, 8, 8], 100, 0.05, stencil=lt.D3Q19)
    for direction, want in (([1, 0, 0], 1), ([0, -1, 0], -2), ([0, 0, 1], 3), ([-1, 0, 0], -1)):
        b = lt.AntiBounceBackOutlet(direction, flow)
        assert b200._outlet_direction(b) == want == b.native_descriptor(flow)["direction"]


def _gpu_case(name):
    """(flow, collision) on the GPU from lettuce_b200's stock classes (the stub reads lettuce's class names)"""
    c32 = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
    c64 = lt.Context(device="cuda", dtype=torch.float64, use_native=True)
    if name == "obstacle2d_bgk":          # stock list: equilibrium inlet, anti-bounce-back outlet, bounce-back
        f = lt.Obstacle(c64, [48, 24], reynolds_number=100, mach_number=0.05, domain_length_x=2)
        m = np.zeros([48, 24], dtype=bool)
        m[10:14, 9:15] = True
        f.mask = m
        return f, lt.BGKCollision(tau=f.units.relaxation_parameter_lu)
    if name == "obstacle3d_kbc":
        f = lt.Obstacle(c32, [24, 12, 8], reynolds_number=50, mach_number=0.05, domain_length_x=2, stencil=lt.D3Q27)
        m = np.zeros([24, 12, 8], dtype=bool)
        m[6:9, 4:8, 2:6] = True
        f.mask = m
        return f, lt.KBCCollision()
    if name == "cavity_bgk":
        f = lt.Cavity2D(c32, [24, 20], 100, 0.05)
        return f, lt.BGKCollision(tau=f.units.relaxation_parameter_lu)
    if name == "poiseuille_guo":
        f = lt.PoiseuilleFlow2D(c64, 16, 10, 0.05)
        tau = f.units.relaxation_parameter_lu
        return f, lt.BGKCollision(tau=tau, force=lt.Guo(f, tau, f.acceleration))
    f = lt.TaylorGreenVortex(c32, [16, 12, 20], 400, 0.1, stencil=lt.D3Q19)
    return f, lt.TRTCollision(tau=f.units.relaxation_parameter_lu, tau_minus=0.8)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["obstacle2d_bgk", "obstacle3d_kbc", "cavity_bgk", "poiseuille_guo", "tgv_trt"])
def test_stub_invoke_equals_simulation(name):
    """b200.invoke(simulation), step by step through the stub's OWN ctypes structures, leaves bitwise the state
    lt.Simulation(n) leaves (same library, same kernels): boundary records, packed streaming masks, outlets,
    forces, f / f_next swapping over an odd number of steps"""
    n = 7
    fa, ca = _gpu_case(name)
    fb, cb = _gpu_case(name)
    assert torch.equal(fa.f, fb.f)
    a = lt.Simulation(fa, ca, [])
    b = lt.Simulation(fb, cb, [])            # only to borrow the masks / boundary order lettuce would build
    a(n)
    b200._lib = None
    os.environ[b200.LIB_ENV] = nat.LIB_PATH
    stub_sim = types.SimpleNamespace(flow=fb, collision=cb, boundaries=b.boundaries,
                                     no_collision_mask=b.no_collision_mask, no_streaming_mask=b.no_streaming_mask)
    with torch.cuda.device(fb.f.device):
        for _ in range(n):
            b200.invoke(stub_sim)
    torch.cuda.synchronize()
    assert torch.isfinite(fb.f).all()
    assert torch.equal(fa.f, fb.f)
