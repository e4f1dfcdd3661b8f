"""Developer tool: rel. error of the host-compiled kernel arithmetic against the reference's
golden vectors and against the oracle on the seeded cases of tests/test_gpu_parity.py.
    python tests/host_emul/check_variants.py [-DFLAG ...] [name filter ...]"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from golden_util import DT, TOL, collision_params, golden_names, load_case, rel_err  # noqa: E402
from host_emul import HostEmul  # noqa: E402
from oracle import lbm_oracle as orc  # noqa: E402

em = HostEmul(tuple(a for a in sys.argv[1:] if a.startswith("-D")))
only = [a for a in sys.argv[1:] if not a.startswith("-D")]


def report(tag, name, dtype, err):
    flag = "" if err <= TOL[dtype] else "   <-- over the bar"
    print(f"{tag:8s} {name:44s} {err:.3e}{flag}")


for name in golden_names():
    case = load_case(name)
    meta = case["meta"]
    if meta["collision"] == "none" or meta["boundaries"] or (only and not any(o in name for o in only)):
        continue
    e = orc.STENCILS[meta["stencil"]]["e"]
    p = collision_params(meta)
    for packed in ([False, True] if meta["dtype"] == "f32" else [False]):
        out = em.run(meta["stencil"], e, case["f0"], meta["n_steps"], meta["collision"], p.get("tau"),
                     p.get("tau_minus", 1.0), packed)
        report("golden", name + (" packed" if packed else ""), meta["dtype"], rel_err(out, case["f_out"]))

SHAPES = {"D2Q9": [[24, 512], [7, 36], [5, 7]], "D3Q19": [[4, 6, 256], [3, 5, 20], [4, 3, 7]],
          "D3Q27": [[3, 4, 132], [2, 3, 9], [1, 2, 4]]}
for stencil, collision in [("D2Q9", "bgk"), ("D2Q9", "trt"), ("D2Q9", "kbc"), ("D3Q19", "bgk"), ("D3Q19", "trt"),
                           ("D3Q27", "bgk"), ("D3Q27", "trt"), ("D3Q27", "kbc")]:
    if only and not any(o in f"{stencil.lower()}_{collision}" for o in only):
        continue
    for dtype in ("f32", "f64"):
        params = dict(tau=0.62)
        if collision == "trt":
            params["tau_minus"] = 0.9
        for k, res in enumerate(SHAPES[stencil]):
            rng = np.random.default_rng(300 + k)
            lat = orc.OracleLattice(stencil, DT[dtype])
            w = np.asarray(orc.STENCILS[stencil]["w"]).reshape([lat.q] + [1] * len(res))
            f0 = torch.tensor(w * (1.0 + 0.02 * rng.random([lat.q, *res])), dtype=DT[dtype])
            ref = orc.run(lat, f0, 10, collision, params).numpy()
            out = em.run(stencil, lat.e_int, f0.numpy(), 10, collision, params["tau"], params.get("tau_minus", 1.0),
                         packed=(dtype == "f32"))
            report("seeded", f"{stencil} {collision} {dtype} {res}", dtype, rel_err(out, ref))
