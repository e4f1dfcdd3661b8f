"""CPU checks of the kernels' ARITHMETIC: lettuce_b200/csrc/lb_physics.cuh compiled for the host
(tests/host_emul, test tooling only -- the product has no CPU path) against the reference's golden
vectors.  Every fused multiply-add in those headers is explicit and the CUDA build uses
-fmad=false, so this is the same sequence of IEEE operations the GPU executes (fp64 bit for bit;
fp32 up to the 1-ulp MUFU.RCP inside KBC).  The GPU parity tests proper are tests/test_gpu_parity.py."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul"))
from golden_util import DT, TOL, collision_params, golden_names, load_case, rel_err  # noqa: E402
from host_emul import HostEmul  # noqa: E402
from oracle import lbm_oracle as orc  # noqa: E402

ILL_CONDITIONED_FIRST_STEP = {"tgv_d2q9_kbc_f32", "tgv_d2q9_kbc_f64"}   # see tests/test_gpu_parity.py


@pytest.fixture(scope="module")
def em():
    return HostEmul()


def _collision_cases():
    out = []
    for name in golden_names():
        meta = load_case(name)["meta"]
        if meta["collision"] != "none" and not meta["boundaries"] and name not in ILL_CONDITIONED_FIRST_STEP:
            out.append(name)
    return out


@pytest.mark.parametrize("name", _collision_cases())
def test_device_arithmetic_vs_golden(em, name):
    case = load_case(name)
    meta = case["meta"]
    e = orc.STENCILS[meta["stencil"]]["e"]
    p = collision_params(meta)
    second = p.get("tau_minus", p.get("smagorinsky_constant", 1.0))   # lb_desc.tau_minus slot
    out = em.run(meta["stencil"], e, case["f0"], meta["n_steps"], meta["collision"], p.get("tau"), second,
                 force=p.get("force"))
    assert rel_err(out, case["f_out"]) <= TOL[meta["dtype"]]
    if meta["dtype"] == "f32":
        # two nodes per packed F2 value: same bits as one node per float
        packed = em.run(meta["stencil"], e, case["f0"], meta["n_steps"], meta["collision"], p.get("tau"),
                        second, packed=True, force=p.get("force"))
        assert np.array_equal(packed, out)


@pytest.mark.parametrize("stencil", ["D2Q9", "D3Q19", "D3Q27"])
@pytest.mark.parametrize("dtype", ["f32", "f64"])
def test_device_moments_vs_oracle(em, stencil, dtype):
    lat = orc.OracleLattice(stencil, DT[dtype])
    rng = np.random.default_rng(5)
    res = [6, 7] if lat.d == 2 else [4, 5, 6]
    f = torch.tensor(rng.random([lat.q, *res]) + 0.5, dtype=DT[dtype])
    n = int(np.prod(res))
    fa = np.ascontiguousarray(f.numpy())
    rho = np.empty(n, fa.dtype)
    u = np.empty([3, n], fa.dtype)
    assert em.lib.emul_moments({"D2Q9": 0, "D3Q19": 1, "D3Q27": 2}[stencil], int(dtype == "f64"), fa.ctypes.data, n,
                               rho.ctypes.data, u.ctypes.data) == 0
    tol = 1e-6 if dtype == "f32" else 1e-14
    assert np.allclose(rho, orc.rho(f).numpy().reshape(-1), rtol=tol)
    uo = orc.u(lat, f).numpy().reshape(lat.d, -1)
    # kernel axes: a 2-D lattice runs as X = 1, so lattice (x, y) = kernel (Y, Z)
    uk = u[1:] if lat.d == 2 else u
    assert np.allclose(uk, uo, rtol=0, atol=tol)


def test_kbc_f32_stepwise_criterion_on_host(em):
    """the comparison rule the GPU test applies to fp32 KBC on random fields (guard discontinuity,
    see tests/test_gpu_parity.py), exercised here with the host-compiled arithmetic"""
    from test_gpu_parity import kbc_f32_stepwise_errors
    stencil, res, tau = "D2Q9", [24, 512], 0.62
    params = dict(tau=tau)
    lat, lat64 = orc.OracleLattice(stencil, torch.float32), orc.OracleLattice(stencil, torch.float64)
    rng = np.random.default_rng(300)
    w = np.asarray(orc.STENCILS[stencil]["w"]).reshape([9, 1, 1])
    f = torch.tensor(w * (1.0 + 0.02 * rng.random([9, *res])), dtype=torch.float32)
    n_special = 0
    for _ in range(4):
        out = em.run(stencil, lat.e_int, f.numpy(), 1, "kbc", tau, packed=True)
        e32, e64, n = kbc_f32_stepwise_errors(out, f, orc, lat, lat64, params, None, None, [], [])
        assert e32 <= TOL["f32"] and e64 <= TOL["f32"]
        n_special += n
        f = torch.from_numpy(out)
    assert 0 < n_special <= 0.01 * 4 * f.numel()


def _boundary_cases():
    out = []
    for name in golden_names():
        meta = load_case(name)["meta"]
        if meta["boundaries"] and all(b["kind"] in ("bounce_back", "equilibrium") for b in meta["boundaries"]) \
                and name not in ILL_CONDITIONED_FIRST_STEP:
            out.append(name)
    return out


@pytest.mark.parametrize("name", _boundary_cases())
def test_device_boundary_arithmetic_vs_golden(em, name):
    """bounce-back and equilibrium boundary nodes with the device headers' bounce_back_node() / equilibrium(),
    masks and destination-masked streaming in numpy, against the reference's runs with boundaries"""
    case = load_case(name)
    meta = case["meta"]
    e = orc.STENCILS[meta["stencil"]]["e"]
    p = collision_params(meta)
    second = p.get("tau_minus", p.get("smagorinsky_constant", 1.0))
    bnds = []
    for b in meta["boundaries"]:
        i = b["index"]
        bnds.append(dict(kind="equilibrium", rho=case[f"b{i}_rho"], u=case[f"b{i}_u"])
                    if b["kind"] == "equilibrium" else dict(kind="bounce_back"))
    out = em.run_masked(meta["stencil"], e, case["f0"], meta["n_steps"], meta["collision"], case["ncm"], case["nsm"],
                        bnds, tau=p.get("tau"), tau_minus=second, force=p.get("force"))
    if meta["collision"] == "none" and not any(b["kind"] == "equilibrium" for b in bnds):
        assert np.array_equal(out, case["f_out"])
    else:
        assert rel_err(out, case["f_out"]) <= TOL[meta["dtype"]]
