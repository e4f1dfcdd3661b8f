"""Import the UNMODIFIED reference (PhiSpel/lettuce).

TEST / BENCH INFRASTRUCTURE ONLY.  Two places are searched, in this order:
  1. baseline/_ref/lettuce  -- the git-ignored install made by baseline/install_ref.sh (a verbatim copy of the
     reference's pure-Python package; pip's offline install fails on `setup_requires=['pytest-runner']`).  It
     travels to the GPU box with `gpurun`, so `bench.py --impl reference` and bench.py's `cpu_baseline` time the
     reference's OWN CPU PyTorch path there (kind "reference").
  2. /root/reference        -- the read-only checkout of the build container (golden-vector generation,
     `tests/golden/make_golden*.py`, and the live cross-checks of `tests/test_oracle_golden.py`).

The reference imports three optional I/O packages that are not installed here (h5py, pyevtk, pyevtk.hl); none
is touched by the stream-collide path, so empty stub modules are registered before the import (SURVEY.md
section 8c).  Nothing in lettuce_b200/ imports this module.
"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INSTALLED_ROOT = os.path.join(ROOT, "baseline", "_ref")
REFERENCE_ROOT = "/root/reference"


def reference_root():
    for r in (INSTALLED_ROOT, REFERENCE_ROOT):
        if os.path.isfile(os.path.join(r, "lettuce", "_simulation.py")):
            return r
    return None


def reference_available() -> bool:
    return reference_root() is not None


def import_reference():
    root = reference_root()
    if root is None:
        raise ImportError(f"reference not found (neither {INSTALLED_ROOT} nor {REFERENCE_ROOT})")
    for name in ("h5py", "pyevtk", "pyevtk.hl"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pyevtk"].hl = sys.modules["pyevtk.hl"]
    if root not in sys.path:
        sys.path.insert(0, root)
    import lettuce  # noqa: E402
    return lettuce
