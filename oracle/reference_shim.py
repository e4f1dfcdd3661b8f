"""Import the UNMODIFIED reference (PhiSpel/lettuce) from /root/reference.

TEST INFRASTRUCTURE ONLY, and only usable in the build container: /root/reference does
not exist on the GPU box.  Used by `tests/golden/make_golden.py` to generate the golden
vectors that pin `oracle/lbm_oracle.py`, and by the optional live cross-check in
`tests/test_oracle_golden.py` (skipped when the reference is absent).

The reference imports three optional I/O packages that are not installed here (h5py,
pyevtk, pyevtk.hl); none is touched by the stream-collide path, so empty stub modules are
registered before the import.
"""
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "lettuce"))


def import_reference():
    if not reference_available():
        raise ImportError(f"reference checkout not found at {REFERENCE_ROOT}")
    for name in ("h5py", "pyevtk", "pyevtk.hl"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pyevtk"].hl = sys.modules["pyevtk.hl"]
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import lettuce  # noqa: E402
    return lettuce
