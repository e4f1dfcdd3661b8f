"""lb_observables_static: the allocation-free variant of lb_observables (reduction scratch in a static __device__
array inside the library, include/lettuce_b200.h).  Collected last on purpose (file name):
a failure here hides nothing behind it.

Called through the C ABI (ctypes) and compared with the existing entry point lb_observables on the same populations,
two calls in a row on different states (the array keeps the previous call's partials and is shared by all calls).
The reduction is deterministic and both entry points launch the same kernel over the same grid, so the values are
required to be EQUAL, which is inside the tolerance the other observable tests use (1e-5 fp32 / 1e-11 fp64 against
the oracle, also checked here)."""
import numpy as np
import pytest
import torch

import lettuce_b200 as lt

pytestmark = pytest.mark.gpu

DT = {"f32": torch.float32, "f64": torch.float64}


def _state(stencil, dtype, res, seed):
    from oracle import lbm_oracle as orc
    lat = orc.OracleLattice(stencil, DT[dtype])
    rng = np.random.default_rng(seed)
    w = np.asarray(orc.STENCILS[stencil]["w"]).reshape([lat.q] + [1] * lat.d)
    return lat, torch.tensor(w * (1 + 0.05 * rng.random([lat.q, *res])), dtype=DT[dtype])


@pytest.mark.parametrize("stencil,dtype,res", [
    ("D3Q19", "f32", [16, 16, 32]),     # vector kernel (4 nodes per thread), several groups of blocks
    ("D2Q9", "f64", [12, 40]),
    ("D3Q27", "f32", [7, 9, 11]),       # odd sizes: one node per thread
    ("D3Q19", "f64", [64, 64, 64]),     # > 256 blocks: the two-level ticket reduction
])
def test_static_scratch_equals_lb_observables_twice(stencil, dtype, res):
    from oracle import lbm_oracle as orc
    nat = lt.native
    L = nat.lib()
    desc = nat.make_desc(stencil, DT[dtype], res)
    scratch = torch.empty(L.lb_obs_scratch_bytes(desc), dtype=torch.uint8, device="cuda")
    rt = 1e-5 if dtype == "f32" else 1e-11
    for call, seed in enumerate((5, 6)):   # two calls in a row, different states
        lat, f0 = _state(stencil, dtype, res, seed)
        f = f0.cuda().contiguous()
        st = nat.current_stream_ptr(f.device)
        ref = torch.full((4,), -7.0, dtype=torch.float64, device="cuda")
        out = torch.full((4,), -7.0, dtype=torch.float64, device="cuda")
        nat.check(L.lb_observables(desc, f.data_ptr(), 0, None, scratch.data_ptr(), ref.data_ptr(), st))
        nat.check(L.lb_observables_static(desc, f.data_ptr(), out.data_ptr(), st))
        r, o = ref.cpu().numpy(), out.cpu().numpy()
        assert np.array_equal(o[:3], r[:3]), (call, o, r)
        assert o[3] == -7.0                      # the enstrophy slot is not touched
        assert o[0] == pytest.approx(float(orc.energy_lu(lat, f0)), rel=rt)
        assert np.sqrt(o[1]) == pytest.approx(float(orc.max_velocity_lu(lat, f0)), rel=rt)
        assert o[2] == pytest.approx(float(f0.double().sum()), rel=rt)


def test_static_scratch_rejects_lattices_beyond_its_maximum():
    nat = lt.native
    L = nat.lib()
    nmax = L.lb_static_obs_max_nodes()
    assert nmax == 256 ** 3
    desc = nat.make_desc("D2Q9", torch.float32, [4096, 4100])   # 16.79 M nodes > 16.78 M; never dereferenced
    dummy = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    rc = L.lb_observables_static(desc, dummy.data_ptr(), out.data_ptr(), nat.current_stream_ptr(dummy.device))
    assert rc != 0
    msg = L.lb_last_error().decode()
    assert "LB_STATIC_OBS_MAX_NODES" in msg and "lb_observables" in msg
    with pytest.raises(nat.NativeError):
        nat.check(rc)
    assert float(out.abs().sum()) == 0.0


def test_static_scratch_makes_no_allocation():
    """device memory in use (driver view) does not change across calls; torch's allocator is not involved at all"""
    nat = lt.native
    L = nat.lib()
    lat, f0 = _state("D3Q19", "f32", [32, 32, 32], 9)
    f = f0.cuda().contiguous()
    desc = nat.make_desc("D3Q19", torch.float32, [32, 32, 32])
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    st = nat.current_stream_ptr(f.device)
    nat.check(L.lb_observables_static(desc, f.data_ptr(), out.data_ptr(), st))   # set-up call (symbol address)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(5):
        nat.check(L.lb_observables_static(desc, f.data_ptr(), out.data_ptr(), st))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_reserved() == reserved0
