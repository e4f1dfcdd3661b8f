"""Helpers shared by the oracle and GPU parity tests: load a golden case written by
tests/golden/make_golden.py and turn it into oracle inputs."""
import glob
import json
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DT = {"f32": torch.float32, "f64": torch.float64}


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0]
                  for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_case(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    case = dict(meta=meta)
    for k in z.files:
        if k != "meta":
            case[k] = z[k]
    return case


def collision_params(meta):
    c = meta["collision"]
    if c == "none":
        return {}
    if c == "trt":
        return dict(tau=meta["tau"], tau_minus=meta["tau_minus"])
    if c == "smagorinsky":   # make_golden_rank3.py stores the Smagorinsky constant in the tau_minus slot
        return dict(tau=meta["tau"], smagorinsky_constant=meta["tau_minus"])
    if c == "bgk" and meta.get("force"):   # BGKCollision(tau, force=Guo | ShanChen(flow, tau, acceleration))
        return dict(tau=meta["tau"], force=meta["force"])
    return dict(tau=meta["tau"])


def rel_err(a, b):
    """max |a - b| / max |b|  (SURVEY section 8c parity procedure)"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


# tolerances north_star states for N-step population parity
TOL = {"f32": 1e-5, "f64": 1e-12}
