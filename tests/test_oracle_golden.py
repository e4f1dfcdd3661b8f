"""Pin the CPU oracle (oracle/lbm_oracle.py) against golden vectors produced by the
unmodified reference (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from golden_util import DT, TOL, collision_params, golden_names, load_case, rel_err
from oracle import lbm_oracle as orc


def oracle_inputs(case):
    meta = case["meta"]
    lat = orc.OracleLattice(meta["stencil"], DT[meta["dtype"]])
    f0 = torch.from_numpy(case["f0"])
    ncm = torch.from_numpy(case["ncm"]) if "ncm" in case else None
    nsm = torch.from_numpy(case["nsm"]) if "nsm" in case else None
    bnds = []
    for b in meta["boundaries"]:
        i = b["index"]
        if b["kind"] == "equilibrium":
            bnds.append(orc.OracleBoundary("equilibrium", None,
                                           rho=torch.from_numpy(case[f"b{i}_rho"]),
                                           u=torch.from_numpy(case[f"b{i}_u"])))
        elif b["kind"] in ("abb_outlet", "eq_outlet_p"):
            bnds.append(orc.OracleBoundary(b["kind"], None, direction=b["direction"],
                                           rho_outlet=b.get("rho_outlet", 1.0)))
        else:
            bnds.append(orc.OracleBoundary("bounce_back", None))
    return lat, f0, ncm, nsm, bnds


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference(name):
    case = load_case(name)
    meta = case["meta"]
    lat, f0, ncm, nsm, bnds = oracle_inputs(case)
    out = orc.run(lat, f0, meta["n_steps"], meta["collision"], collision_params(meta),
                  ncm=ncm, nsm=nsm, boundaries=bnds).numpy()
    ref = case["f_out"]
    assert out.dtype == ref.dtype and out.shape == ref.shape
    if meta["collision"] == "none" and not any(b["kind"] == "equilibrium" for b in meta["boundaries"]):
        assert np.array_equal(out, ref), "streaming / bounce-back must be bit-exact"
    else:
        # same torch kernels, same op order -> round-off level, far inside the budget
        assert rel_err(out, ref) <= TOL[meta["dtype"]] * 0.1


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("tgv_") and "_bgk_" in n])
def test_oracle_observables(name):
    case = load_case(name)
    meta = case["meta"]
    lat = orc.OracleLattice(meta["stencil"], DT[meta["dtype"]])
    c, dx, d = meta["u_lu_to_pu"], meta["dx"], lat.d
    for fkey, okey in (("f0", "obs0"), ("f_out", "obs_out")):
        f = torch.from_numpy(case[fkey])
        e = float(orc.energy_lu(lat, f)) * c * c * dx ** d
        m = float(orc.max_velocity_lu(lat, f)) * c
        w = float(orc.enstrophy_lu(lat, f)) * (c / dx) ** 2 * dx ** d
        tol = 2e-5 if meta["dtype"] == "f32" else 1e-11
        assert e == pytest.approx(case[okey][0], rel=tol)
        assert m == pytest.approx(case[okey][1], rel=tol)
        assert w == pytest.approx(case[okey][2], rel=10 * tol)


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("outlet_")])
def test_oracle_outlet_masks(name):
    """the oracle's restatement of make_no_collision_mask / make_no_streaming_mask of the outlet
    boundaries reproduces the masks the reference built"""
    case = load_case(name)
    meta = case["meta"]
    lat, f0, ncm, nsm, bnds = oracle_inputs(case)
    for i, (b, mb) in enumerate(zip(bnds, meta["boundaries"]), start=1):
        if mb["kind"] == "bounce_back":
            b.mask = torch.from_numpy(case[f"b{i}_mask"])
        elif mb["kind"] == "equilibrium":
            b.mask = ncm == i
    ncm2, nsm2 = orc.build_masks(lat, meta["res"], bnds)
    assert torch.equal(ncm2, ncm) and torch.equal(nsm2, nsm)
