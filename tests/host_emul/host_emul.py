"""TEST TOOLING ONLY: host build of the kernels' per-node arithmetic (see emul.cpp).

`HostEmul` compiles tests/host_emul/emul.cpp with g++ into a temporary shared object and steps a
lattice with it (collision = the device headers compiled for the host, streaming = numpy roll),
so that formula variants can be compared with the reference's golden vectors without a GPU.
The product never imports this module."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
STENCIL_ID = {"D2Q9": 0, "D3Q19": 1, "D3Q27": 2}
COLL_ID = {"none": 0, "bgk": 1, "trt": 2, "kbc": 3, "regularized": 4, "smagorinsky": 5, "bgk_force": 6}


class HostEmul:
    def __init__(self, extra_flags=()):
        self._dir = tempfile.mkdtemp(prefix="lb_host_emul_")
        so = os.path.join(self._dir, "emul.so")
        cmd = ["g++", "-O2", "-std=c++17", "-mfma", "-ffp-contract=off", "-shared", "-fPIC",
               *extra_flags, "-o", so, os.path.join(HERE, "emul.cpp")]
        subprocess.run(cmd, check=True, capture_output=True)
        self.lib = C.CDLL(so)
        self.lib.emul_collide.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_long,
                                          C.POINTER(C.c_double)]
        self.lib.emul_moments.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p]

    @staticmethod
    def params(collision, tau=None, tau_minus=1.0, force=None):
        """the scalars of lb::CollisionParams, as lb_capi.cu make_cp derives them (for 'smagorinsky'
        tau_minus carries the Smagorinsky constant, as in lb_desc)"""
        cp = (C.c_double * 16)()
        if collision == "none":
            return cp
        cp[0] = 1.0 / tau
        cp[1] = 1.0 / (2.0 * tau)
        cp[2] = 1.0 / (2.0 * tau_minus)
        beta = 1.0 / (2.0 * tau)
        cp[3] = beta
        cp[4] = 1.0 / beta
        cp[5] = 1.0 - 1.0 / tau
        cp[6] = tau
        cp[7] = (tau - 0.5) / 3.0
        cp[8] = tau_minus * tau_minus
        if force is not None:   # dict(kind='guo'|'shan_chen', tau=..., acceleration=[...]) in LATTICE axes
            acc = [float(a) for a in force["acceleration"]]
            acc = [0.0] + acc if len(acc) == 2 else acc          # kernel axes: a 2-D lattice runs as X = 1
            scale = 0.5 if force["kind"] == "guo" else force["tau"] * 1
            for c in range(3):
                cp[9 + c] = acc[c]
                cp[12 + c] = scale * acc[c]
            cp[15] = 1 - 1 / (2 * force["tau"]) if force["kind"] == "guo" else 0.0
        return cp

    def collide(self, stencil, f, collision, cp, packed=False):
        """f: numpy [Q, *res] (C-contiguous) -> collided copy"""
        if collision == "none":
            return f.copy()
        out = np.ascontiguousarray(f).copy()
        n = int(np.prod(f.shape[1:]))
        rc = self.lib.emul_collide(STENCIL_ID[stencil], 1 if f.dtype == np.float64 else 0, int(packed),
                                   COLL_ID[collision], out.ctypes.data, n, cp)
        assert rc == 0
        return out

    def run(self, stencil, e, f0, n_steps, collision, tau=None, tau_minus=1.0, packed=False, force=None):
        cp = self.params(collision, tau, tau_minus, force)
        if force is not None:
            assert collision == "bgk"
            collision = "bgk_force"
        f = np.ascontiguousarray(f0).copy()
        axes = tuple(range(f.ndim - 1))
        for _ in range(n_steps):
            f = self.collide(stencil, f, collision, cp, packed)
            for q in range(1, f.shape[0]):
                f[q] = np.roll(f[q], shift=tuple(e[q]), axis=axes)
        return f
