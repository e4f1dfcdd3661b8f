"""The single-purpose entry points of the C ABI called directly through ctypes on CUDA tensors:
lb_stream (Simulation._stream), lb_collide (Collision.__call__ / Simulation._collide), lb_moments
(Flow.rho / Flow.u), lb_pack_streaming_mask, lb_observables / lb_enstrophy -- against the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

from golden_util import DT, TOL, load_case, rel_err

pytestmark = pytest.mark.gpu


def _nat():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lettuce_b200 import _native as nat
    return nat


@pytest.mark.parametrize("name", ["stream_d2q9_rand_f32", "stream_d3q19_rand_f64", "stream_d3q27_rand_f32"])
def test_lb_stream_is_the_reference_streaming(name):
    nat = _nat()
    case = load_case(name)
    meta = case["meta"]
    desc = nat.make_desc(meta["stencil"], DT[meta["dtype"]], meta["res"])
    a = torch.from_numpy(case["f0"]).cuda().contiguous()
    b = torch.empty_like(a)
    st = nat.current_stream_ptr(a.device)
    for _ in range(meta["n_steps"]):
        nat.check(nat.lib().lb_stream(desc, a.data_ptr(), b.data_ptr(), None, st))
        a, b = b, a
    assert np.array_equal(a.cpu().numpy(), case["f_out"])
    # in place is refused, and so is a mask (that path is lb_step with LB_COLL_NONE)
    assert nat.lib().lb_stream(desc, a.data_ptr(), a.data_ptr(), None, st) == -1
    bits = torch.zeros(meta["res"], dtype=torch.int32, device="cuda")
    assert nat.lib().lb_stream(desc, a.data_ptr(), b.data_ptr(), bits.data_ptr(), st) == -1


@pytest.mark.parametrize("stencil,collision", [("D2Q9", "kbc"), ("D3Q19", "trt"), ("D3Q27", "smagorinsky")])
@pytest.mark.parametrize("dtype", ["f32", "f64"])
def test_lb_collide_with_boundaries_vs_oracle(stencil, collision, dtype):
    """collision where ncm == 0, bounce-back where ncm == 1, equilibrium where ncm == 2, in place"""
    if collision == "kbc" and dtype == "f32":
        pytest.skip("fp32 KBC on random fields is compared step-wise around the gamma guard in test_gpu_parity.py")
    nat = _nat()
    from oracle import lbm_oracle as orc
    lat = orc.OracleLattice(stencil, DT[dtype])
    res = [9, 20] if lat.d == 2 else [5, 6, 12]
    rng = np.random.default_rng(17)
    w = np.asarray(orc.STENCILS[stencil]["w"]).reshape([lat.q] + [1] * lat.d)
    f0 = torch.tensor(w * (1 + 0.02 * rng.random([lat.q, *res])), dtype=DT[dtype])
    solid = torch.tensor(rng.random(res) < 0.15)
    eqm = torch.tensor(rng.random(res) < 0.1)
    rho_b = torch.tensor(1.01, dtype=DT[dtype]).reshape([1] * (lat.d + 1))
    u_b = torch.tensor([0.03, -0.01, 0.02][:lat.d], dtype=DT[dtype]).reshape([lat.d] + [1] * lat.d)
    bnds = [orc.OracleBoundary("bounce_back", solid), orc.OracleBoundary("equilibrium", eqm, rho=rho_b, u=u_b)]
    ncm, _ = orc.build_masks(lat, res, bnds)
    params = dict(tau=0.7)
    second = 1.0
    if collision == "trt":
        params["tau_minus"] = second = 0.9
    if collision == "smagorinsky":
        params["smagorinsky_constant"] = second = 0.25
    ref = orc.collide(lat, f0, collision, params, ncm=ncm, boundaries=bnds)

    cid = {"kbc": nat.LB_COLL_KBC, "trt": nat.LB_COLL_TRT, "smagorinsky": nat.LB_COLL_SMAGORINSKY}[collision]
    desc = nat.make_desc(stencil, DT[dtype], res, cid, 0.7, second)
    f = f0.cuda().contiguous()
    ncm_d = ncm.to(torch.uint8).cuda().contiguous()
    rho_d, u_d = rho_b.cuda().contiguous(), u_b.cuda().contiguous()
    arr = (nat.LbBoundary * 2)()
    arr[0].kind = nat.LB_BND_BOUNCE_BACK
    arr[1].kind = nat.LB_BND_EQUILIBRIUM
    arr[1].rho, arr[1].u = rho_d.data_ptr(), u_d.data_ptr()
    arr[1].u_comp_stride = u_d.stride(0)
    nat.check(nat.lib().lb_collide(desc, f.data_ptr(), f.data_ptr(), ncm_d.data_ptr(),
                                   C.cast(arr, C.POINTER(nat.LbBoundary)), 2, nat.current_stream_ptr(f.device)))
    out = f.cpu().numpy()
    assert rel_err(out, ref.numpy()) <= TOL[dtype]
    # bounce-back nodes are pure data movement: bit-exact
    sel = (ncm == 1).numpy()
    assert np.array_equal(out[:, sel], ref.numpy()[:, sel])


@pytest.mark.parametrize("stencil", ["D2Q9", "D3Q19", "D3Q27"])
@pytest.mark.parametrize("dtype", ["f32", "f64"])
def test_lb_moments_and_observables_vs_oracle(stencil, dtype):
    nat = _nat()
    from oracle import lbm_oracle as orc
    lat = orc.OracleLattice(stencil, DT[dtype])
    res = [12, 40] if lat.d == 2 else [8, 8, 12]
    rng = np.random.default_rng(23)
    w = np.asarray(orc.STENCILS[stencil]["w"]).reshape([lat.q] + [1] * lat.d)
    f0 = torch.tensor(w * (1 + 0.05 * rng.random([lat.q, *res])), dtype=DT[dtype])
    desc = nat.make_desc(stencil, DT[dtype], res)
    f = f0.cuda().contiguous()
    rho = torch.empty(res, dtype=DT[dtype], device="cuda")
    u = torch.empty([lat.d, *res], dtype=DT[dtype], device="cuda")
    st = nat.current_stream_ptr(f.device)
    L = nat.lib()
    nat.check(L.lb_moments(desc, f.data_ptr(), rho.data_ptr(), u.data_ptr(), st))
    tol = 2e-6 if dtype == "f32" else 1e-13
    assert torch.allclose(rho.cpu(), orc.rho(f0)[0], rtol=tol, atol=0)
    assert torch.allclose(u.cpu(), orc.u(lat, f0), rtol=0, atol=tol)
    # rho only / u only
    rho2 = torch.empty_like(rho)
    nat.check(L.lb_moments(desc, f.data_ptr(), rho2.data_ptr(), None, st))
    assert torch.equal(rho2, rho)
    # reductions: energy, max |u|^2, mass, enstrophy (6th-order stencil needs >= 7 nodes per axis)
    scratch = torch.empty(L.lb_obs_scratch_bytes(desc), dtype=torch.uint8, device="cuda")
    obs = torch.zeros(4, dtype=torch.float64, device="cuda")
    us = torch.empty_like(u)
    nat.check(L.lb_observables(desc, f.data_ptr(), 1, us.data_ptr(), scratch.data_ptr(), obs.data_ptr(), st))
    o = obs.cpu().numpy()
    rt = 1e-5 if dtype == "f32" else 1e-11
    assert o[0] == pytest.approx(float(orc.energy_lu(lat, f0)), rel=rt)
    assert np.sqrt(o[1]) == pytest.approx(float(orc.max_velocity_lu(lat, f0)), rel=rt)
    assert o[2] == pytest.approx(float(f0.double().sum()), rel=rt)
    assert o[3] == pytest.approx(float(orc.enstrophy_lu(lat, f0)), rel=10 * rt)
    assert torch.allclose(us, u, rtol=0, atol=1e-7 if dtype == "f32" else 1e-15)
    # lb_enstrophy on a given velocity field
    obs2 = torch.zeros(4, dtype=torch.float64, device="cuda")
    nat.check(L.lb_enstrophy(desc, u.data_ptr(), scratch.data_ptr(), obs2.data_ptr(), st))
    assert obs2[3].item() == pytest.approx(o[3], rel=1e-12)


def test_lb_pack_streaming_mask_bits_and_hint():
    """bit q = (nsm[q, x] == 1); bit 31 = the node or one of its 26 periodic neighbours keeps something"""
    nat = _nat()
    res = [6, 5, 8]
    rng = np.random.default_rng(31)
    nsm = (rng.random([19, *res]) < 0.01)
    nsm[0] = False
    desc = nat.make_desc("D3Q19", torch.float32, res)
    nsm_d = torch.from_numpy(nsm.astype(np.uint8)).cuda().contiguous()
    bits = torch.empty(res, dtype=torch.int32, device="cuda")
    nat.check(nat.lib().lb_pack_streaming_mask(desc, nsm_d.data_ptr(), bits.data_ptr(),
                                               nat.current_stream_ptr(bits.device)))
    got = bits.cpu().numpy().astype(np.int64) & 0xffffffff
    want = np.zeros(res, dtype=np.int64)
    for q in range(19):
        want |= nsm[q].astype(np.int64) << q
    anyb = want != 0
    near = np.zeros(res, dtype=bool)
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                near |= np.roll(anyb, (dx, dy, dz), axis=(0, 1, 2))
    assert np.array_equal(got & 0x7fffffff, want)
    assert np.array_equal((got >> 31) != 0, near)
