"""lb_run_host: the host-buffer entry point of the C ABI (include/lettuce_b200.h) -- populations, masks
and boundary fields in HOST memory, one call = copy in, n steps, copy out.  Checked against golden files
the unmodified reference produced (tests/golden), through ctypes only: no torch CUDA tensor is involved."""
import ctypes as C

import numpy as np
import pytest
import torch

from golden_util import TOL, load_case, rel_err

pytestmark = pytest.mark.gpu

CASES = ["stream_d2q9_pulses_f32", "stream_d3q27_rand_f64", "maskstream_d3q19_f64", "tgv_d3q19_bgk_f32",
         "tgv_d3q27_kbc_f64", "rand_d2q9_trt_f64", "obstacle_d2q9_bgk_f64", "outlet_abb_d2q9_bgk_f32",
         "outlet_eqp_d3q19_bgk_f64", "rank3_tgv_d3q19_smagorinsky_f32", "rank3_rand_d3q27_regularized_f64",
         "rank3_force_d2q9_guo_f64", "rank3_force_d3q19_shan_chen_f32"]


def _desc(nat, meta):
    coll = {"none": nat.LB_COLL_NONE, "bgk": nat.LB_COLL_BGK, "trt": nat.LB_COLL_TRT, "kbc": nat.LB_COLL_KBC,
            "regularized": nat.LB_COLL_REGULARIZED, "smagorinsky": nat.LB_COLL_SMAGORINSKY}[meta["collision"]]
    fo = meta.get("force")
    kw = {}
    if fo:
        kw = dict(force=nat.LB_FORCE_GUO if fo["kind"] == "guo" else nat.LB_FORCE_SHAN_CHEN,
                  acceleration=fo["acceleration"], force_tau=fo["tau"])
    dt = torch.float32 if meta["dtype"] == "f32" else torch.float64
    return nat.make_desc(meta["stencil"], dt, meta["res"], coll, meta.get("tau") or 1.0, meta.get("tau_minus", 1.0), **kw)


@pytest.mark.parametrize("name", CASES)
def test_run_host_matches_golden(name):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lettuce_b200 import _native as nat
    case = load_case(name)
    meta = case["meta"]
    desc = _desc(nat, meta)
    f = np.ascontiguousarray(case["f0"]).copy()
    ncm = np.ascontiguousarray(case["ncm"]).astype(np.uint8) if "ncm" in case else None
    nsm = np.ascontiguousarray(case["nsm"]).astype(np.uint8) if "nsm" in case else None
    nb = len(meta["boundaries"])
    arr = (nat.LbBoundary * max(nb, 1))()
    keep = []
    for k, b in enumerate(meta["boundaries"]):
        i = b["index"]
        assert i == k + 1
        if b["kind"] == "equilibrium":
            rho = torch.from_numpy(np.ascontiguousarray(case[f"b{i}_rho"]))
            u = torch.from_numpy(np.ascontiguousarray(case[f"b{i}_u"]))
            arr[k].kind = nat.LB_BND_EQUILIBRIUM
            arr[k].rho, arr[k].u = rho.data_ptr(), u.data_ptr()
            rs = nat.broadcast_strides(rho, max(0, rho.dim() - len(meta["res"])), meta["res"])
            us = nat.broadcast_strides(u, 1, meta["res"])
            for a in range(3):
                arr[k].rho_stride[a], arr[k].u_stride[a] = rs[a], us[a]
            arr[k].u_comp_stride = u.stride(0)
            keep += [rho, u]
        elif b["kind"] in ("abb_outlet", "eq_outlet_p"):
            arr[k].kind = nat.LB_BND_ABB_OUTLET if b["kind"] == "abb_outlet" else nat.LB_BND_EQ_OUTLET_P
            axis = next(a for a, v in enumerate(b["direction"]) if v != 0)
            arr[k].direction = (axis + 1) * b["direction"][axis]
            arr[k].value = float(b.get("rho_outlet", 0.0))
        else:
            arr[k].kind = nat.LB_BND_BOUNCE_BACK
    obs = (C.c_double * 4)()
    rc = nat.lib().lb_run_host(desc, f.ctypes.data, None if ncm is None else ncm.ctypes.data,
                               None if nsm is None else nsm.ctypes.data,
                               C.cast(arr, C.POINTER(nat.LbBoundary)), nb, meta["n_steps"], obs)
    assert rc == 0, nat.lib().lb_last_error()
    ref = case["f_out"]
    exact = meta["collision"] == "none" and all(b["kind"] == "bounce_back" for b in meta["boundaries"])
    if exact:
        assert np.array_equal(f, ref)
    else:
        assert rel_err(f, ref) <= TOL[meta["dtype"]]
    # the observables of the final state: mass = sum of all populations
    assert obs[2] == pytest.approx(float(ref.astype(np.float64).sum()), rel=1e-6 if meta["dtype"] == "f32" else 1e-12)


def test_run_host_rejects_bad_arguments():
    from lettuce_b200 import _native as nat
    L = nat.lib()
    d = nat.make_desc("D2Q9", torch.float32, [8, 8], nat.LB_COLL_BGK, 0.6)
    assert L.lb_run_host(d, None, None, None, None, 0, 1, None) == -1
    assert b"null host buffer" in L.lb_last_error()
