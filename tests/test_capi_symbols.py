"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/*.h declares,
and validates descriptors without touching a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            text = open(os.path.join(ROOT, "include", fn)).read()
            names += re.findall(r"^LB_API\s+[\w\s\*]+?\b(lb_\w+)\s*\(", text, flags=re.M)
    return sorted(set(names))


def test_header_declares_the_binding_table():
    from lettuce_b200 import _native as nat
    assert declared_symbols() == sorted(nat.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol():
    from lettuce_b200 import _native as nat
    assert os.path.exists(nat.LIB_PATH), "build the library first: python -c 'import __graft_entry__ as g; g.build()'"
    L = C.CDLL(nat.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), name
    assert nat.lib().lb_abi_version() == nat.ABI_VERSION


def test_descriptor_validation_needs_no_gpu():
    from lettuce_b200 import _native as nat
    import torch
    L = nat.lib()
    assert [L.lb_stencil_q(i) for i in range(4)] == [9, 19, 27, -1]
    d = nat.make_desc("D3Q19", torch.float32, [256, 256, 256], nat.LB_COLL_BGK, 0.6)
    assert L.lb_bytes_per_update(d) == 152            # 2 * Q * sizeof(float), BASELINE.md section 3
    assert L.lb_vector_width(d, 0) == 4 and L.lb_vector_width(d, 1) == 4   # masked streaming is vectorised too
    d64 = nat.make_desc("D2Q9", torch.float64, [256, 256], nat.LB_COLL_BGK, 0.6)
    assert L.lb_bytes_per_update(d64) == 144 and L.lb_vector_width(d64, 0) == 2
    odd = nat.make_desc("D3Q27", torch.float32, [5, 7, 9])
    assert L.lb_vector_width(odd, 0) == 1
    # KBC exists for D2Q9 / D3Q27 only (kbc_collision.py:101,115)
    bad = nat.make_desc("D3Q19", torch.float32, [8, 8, 8], nat.LB_COLL_KBC, 0.6)
    fl = nat.LbFields()
    assert L.lb_step(bad, fl, 1, None) == -3
    assert b"KBC" in L.lb_last_error()
    ok = nat.make_desc("D3Q19", torch.float32, [8, 8, 8], nat.LB_COLL_BGK, 0.6)
    assert L.lb_step(ok, fl, 1, None) == -1           # null population buffers
    neg = nat.make_desc("D3Q19", torch.float32, [8, 8, 8], nat.LB_COLL_BGK, -1.0)
    assert L.lb_step(neg, fl, 1, None) == -1
    with pytest.raises(nat.NativeError):
        nat.make_desc("D3Q15", torch.float32, [8, 8, 8])
    # rank-3 collisions: ids 4 and 5 are known, 6 is not; the Smagorinsky constant rides in tau_minus
    for cid in (nat.LB_COLL_REGULARIZED, nat.LB_COLL_SMAGORINSKY):
        assert L.lb_step(nat.make_desc("D3Q19", torch.float32, [8, 8, 8], cid, 0.6, 0.17), fl, 1, None) == -1
        assert b"null population buffer" in L.lb_last_error()
    assert L.lb_step(nat.make_desc("D3Q19", torch.float32, [8, 8, 8], 6, 0.6), fl, 1, None) == -1
    assert b"unknown collision" in L.lb_last_error()
    assert L.lb_step(nat.make_desc("D3Q19", torch.float32, [8, 8, 8], nat.LB_COLL_SMAGORINSKY, 0.6, -0.1), fl, 1, None) == -1
    assert b"Smagorinsky" in L.lb_last_error()


def test_output_fields_validation_needs_no_gpu():
    """lb_output_fields rejects bad descriptors / null buffers before any CUDA call"""
    from lettuce_b200 import _native as nat
    import torch
    L = nat.lib()
    d = nat.make_desc("D3Q19", torch.float32, [8, 8, 8])
    assert L.lb_output_fields(d, None, 1.0, 1.0, 1.0, None, None) == -1
    assert b"null pointer" in L.lb_last_error()
    bad = nat.make_desc("D3Q19", torch.float32, [8, 8, 8])
    bad.ndim = 2
    assert L.lb_output_fields(bad, 1, 1.0, 1.0, 1.0, 1, None) == -1
    assert b"ndim" in L.lb_last_error()


def test_rank3_collision_descriptors():
    """host mirror of RegularizedCollision / SmagorinskyCollision (regularized_collision.py:14-19,
    smagorinsky_collision.py:13-18): constructor signatures and what reaches the kernel"""
    import torch
    import lettuce_b200 as lt
    from lettuce_b200 import _native as nat
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=16, reynolds_number=10, mach_number=0.05, stencil=lt.D2Q9)
    reg = lt.RegularizedCollision(tau=99.0)                    # ignored, as in the reference
    d = reg.native_descriptor(flow)
    assert d == dict(collision=nat.LB_COLL_REGULARIZED, tau=float(flow.units.relaxation_parameter_lu))
    smag = lt.SmagorinskyCollision(0.6)
    assert smag.constant == 0.17 and smag.iterations == 2
    assert smag.native_descriptor(flow) == dict(collision=nat.LB_COLL_SMAGORINSKY, tau=0.6, tau_minus=0.17)
    with pytest.raises(NotImplementedError):
        lt.SmagorinskyCollision(0.6, force=object())
    # forcing schemes (ext/_force/guo.py, shan_chen.py): constructor (flow, tau, acceleration)
    guo = lt.Guo(flow, 0.7, [1e-4, 0.0])
    sc = lt.ShanChen(flow, 0.7, [1e-4, 0.0])
    assert guo.ueq_scaling_factor == 0.5 and sc.ueq_scaling_factor == 0.7
    bgk = lt.BGKCollision(0.6, force=guo)
    assert bgk.name() == "BGKCollision_Guo"
    assert bgk.native_descriptor(flow) == dict(collision=nat.LB_COLL_BGK, tau=0.6, force=nat.LB_FORCE_GUO,
                                               force_tau=0.7, acceleration=[1e-4, 0.0])
    assert lt.BGKCollision(0.6, force=sc).native_descriptor(flow)["force"] == nat.LB_FORCE_SHAN_CHEN
    with pytest.raises(NotImplementedError):
        lt.Guo(flow, 0.7, [1e-4, 0.0, 0.0])                    # wrong number of components
    with pytest.raises(NotImplementedError):
        lt.BGKCollision(0.6, force=object())
    # the C ABI refuses a force on any other collision, and bad force ids
    L = nat.lib()
    fl = nat.LbFields()
    d = nat.make_desc("D2Q9", torch.float64, [16, 16], nat.LB_COLL_TRT, 0.6, 1.0, nat.LB_FORCE_GUO, [1e-4, 0.0], 0.7)
    assert L.lb_step(d, fl, 1, None) == -3 and b"BGKCollision only" in L.lb_last_error()
    d = nat.make_desc("D2Q9", torch.float64, [16, 16], nat.LB_COLL_BGK, 0.6, 1.0, 7, [1e-4, 0.0], 0.7)
    assert L.lb_step(d, fl, 1, None) == -1 and b"unknown force" in L.lb_last_error()
    with pytest.raises(nat.NativeError):
        nat.make_desc("D2Q9", torch.float64, [16, 16], nat.LB_COLL_BGK, 0.6, 1.0, nat.LB_FORCE_GUO, [1e-4], 0.7)


def test_no_cpu_fallback():
    """the product refuses to step without the CUDA engine instead of falling back"""
    import torch
    import lettuce_b200 as lt
    ctx = lt.Context(device="cpu", dtype=torch.float64, use_native=False)
    flow = lt.TaylorGreenVortex(ctx, resolution=16, reynolds_number=10, mach_number=0.05, stencil=lt.D2Q9)
    with pytest.raises(RuntimeError, match="no torch / CPU fallback"):
        lt.Simulation(flow, lt.BGKCollision(0.6), [])
    with pytest.raises(lt.native.NativeError):
        lt.BGKCollision(0.6)(flow)


def test_product_does_not_import_the_oracle():
    """nothing under lettuce_b200/ imports, loads or executes anything under oracle/"""
    pkg = os.path.join(ROOT, "lettuce_b200")
    bad = re.compile(r"^\s*(from|import)\s+oracle\b|oracle[/.]lbm_oracle|oracle/_ref|import_module\(.oracle", re.M)
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, fn)).read()
                assert not bad.search(text), os.path.join(dirpath, fn)
