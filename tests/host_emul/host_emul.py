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
        self.lib.emul_equilibrium.argtypes = [C.c_int, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.emul_bounce_back.argtypes = [C.c_int, C.c_int, C.c_long, C.c_void_p]

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

    def equilibrium(self, stencil, rho, u, res, dtype):
        """rho / u: arrays in lattice units that broadcast against [1, *res] / [d, *res] (the shapes an
        EquilibriumBoundaryPU accepts) -> feq [Q, *res] from the device header's equilibrium()"""
        d = len(res)
        n = int(np.prod(res))
        rho = np.asarray(rho, dtype=dtype)
        u = np.asarray(u, dtype=dtype)
        rho_f = np.ascontiguousarray(np.broadcast_to(rho.reshape((1,) * (d - rho.ndim) + rho.shape if rho.ndim <= d
                                                                 else rho.shape[-d:]), res)).reshape(n)
        u = u.reshape([u.shape[0]] + [1] * (d + 1 - u.ndim) + list(u.shape[1:]))
        u_l = np.broadcast_to(u, [d, *res]).reshape(d, n)
        u_k = np.zeros([3, n], dtype=dtype)          # kernel axes: a 2-D lattice runs as X = 1
        u_k[3 - d:] = u_l
        q = {"D2Q9": 9, "D3Q19": 19, "D3Q27": 27}[stencil]
        out = np.empty([q, n], dtype=dtype)
        rc = self.lib.emul_equilibrium(STENCIL_ID[stencil], 1 if dtype == np.float64 else 0, n,
                                       rho_f.ctypes.data, np.ascontiguousarray(u_k).ctypes.data, out.ctypes.data)
        assert rc == 0
        return out.reshape([q, *res])

    def bounce_back(self, stencil, f):
        out = np.ascontiguousarray(f).copy()
        n = int(np.prod(f.shape[1:]))
        assert self.lib.emul_bounce_back(STENCIL_ID[stencil], 1 if f.dtype == np.float64 else 0, n,
                                         out.ctypes.data) == 0
        return out

    def run_masked(self, stencil, e, f0, n_steps, collision, ncm, nsm, boundaries, tau=None, tau_minus=1.0,
                   force=None):
        """collide where ncm == 0, boundary i where ncm == i (bounce-back or equilibrium, device arithmetic),
        destination-masked streaming (reference/lettuce/_simulation.py:160-189); boundaries: list of
        dict(kind='bounce_back') | dict(kind='equilibrium', rho=..., u=...) in index order"""
        cp = self.params(collision, tau, tau_minus, force)
        if force is not None:
            collision = "bgk_force"
        f = np.ascontiguousarray(f0).copy()
        res = list(f.shape[1:])
        axes = tuple(range(f.ndim - 1))
        feq = {i: self.equilibrium(stencil, b["rho"], b["u"], res, f.dtype.type)
               for i, b in enumerate(boundaries, start=1) if b["kind"] == "equilibrium"}
        for _ in range(n_steps):
            new = np.where(ncm[None] == 0, self.collide(stencil, f, collision, cp), f)
            bb = None
            for i, b in enumerate(boundaries, start=1):
                if b["kind"] == "equilibrium":
                    new = np.where(ncm[None] == i, feq[i], new)
                else:
                    bb = self.bounce_back(stencil, f) if bb is None else bb
                    new = np.where(ncm[None] == i, bb, new)
            f = new
            for q in range(1, f.shape[0]):
                f[q] = np.where(nsm[q] == 1, f[q], np.roll(f[q], shift=tuple(e[q]), axis=axes))
        return f
