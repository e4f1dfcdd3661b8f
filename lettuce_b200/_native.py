"""ctypes binding of the C ABI declared in include/lettuce_b200.h.

The shared library `liblettuce_b200.so` is built in-tree by `lettuce_b200/csrc/Makefile`
(`python __graft_entry__.py` / `__graft_entry__.build()`).  There is no fallback: if the
library is missing, every native call raises `NativeLibraryMissing`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

# LETTUCE_B200_LIB: another build of the same library (A/B of build variants in scripts/; never a fallback)
LIB_PATH = os.environ.get("LETTUCE_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "liblettuce_b200.so")

LB_D2Q9, LB_D3Q19, LB_D3Q27 = 0, 1, 2
LB_F32, LB_F64 = 0, 1
LB_COLL_NONE, LB_COLL_BGK, LB_COLL_TRT, LB_COLL_KBC = 0, 1, 2, 3
LB_COLL_REGULARIZED, LB_COLL_SMAGORINSKY = 4, 5
LB_FORCE_NONE, LB_FORCE_GUO, LB_FORCE_SHAN_CHEN = 0, 1, 2
LB_BND_BOUNCE_BACK, LB_BND_EQUILIBRIUM, LB_BND_ABB_OUTLET, LB_BND_EQ_OUTLET_P = 1, 2, 3, 4
LB_OBS_ENERGY, LB_OBS_MAXU2, LB_OBS_MASS, LB_OBS_ENSTROPHY, LB_OBS_COUNT = 0, 1, 2, 3, 4
LB_MAX_BOUNDARIES = 8
ABI_VERSION = 6

STENCIL_IDS = {"D2Q9": LB_D2Q9, "D3Q19": LB_D3Q19, "D3Q27": LB_D3Q27}
DTYPE_IDS = {torch.float32: LB_F32, torch.float64: LB_F64}

# every symbol include/lettuce_b200.h declares (checked by tests/test_capi_symbols.py)
EXPORTED_SYMBOLS = (
    "lb_abi_version", "lb_last_error", "lb_step", "lb_collide", "lb_stream",
    "lb_pack_streaming_mask", "lb_moments", "lb_initialize", "lb_obs_scratch_bytes", "lb_observables",
    "lb_enstrophy", "lb_run_host", "lb_stencil_q", "lb_bytes_per_update", "lb_vector_width",
    "lb_unpack_halo", "lb_zface_count", "lb_ipc_alloc", "lb_ipc_free", "lb_ipc_export",
    "lb_ipc_open", "lb_ipc_close", "lb_output_fields", "lb_unpack_halo_ex", "lb_halo_ll_bytes",
    "lb_enstrophy_slab", "lb_static_obs_max_nodes", "lb_observables_static",
    "lb_plan_resident",
)


class NativeLibraryMissing(RuntimeError):
    pass


class NativeError(RuntimeError):
    pass


class LbDesc(C.Structure):
    _fields_ = [("stencil", C.c_int32), ("dtype", C.c_int32), ("collision", C.c_int32),
                ("ndim", C.c_int32), ("dims", C.c_int64 * 3), ("tau", C.c_double),
                ("tau_minus", C.c_double), ("acceleration", C.c_double * 3), ("force_tau", C.c_double),
                ("force", C.c_int32), ("kernel_flags", C.c_int32)]


LB_KF_NO_TMA, LB_KF_FORCE_TMA, LB_KF_NO_LOOP = 1, 2, 4
LB_KF_FORCE_PREFETCH, LB_KF_NO_PREFETCH, LB_KF_NO_RESIDENT = 8, 16, 32


class LbBoundary(C.Structure):
    _fields_ = [("kind", C.c_int32), ("direction", C.c_int32), ("rho", C.c_void_p), ("u", C.c_void_p),
                ("rho_stride", C.c_int64 * 3), ("u_stride", C.c_int64 * 3),
                ("u_comp_stride", C.c_int64), ("value", C.c_double)]


LB_HALO_PACKED, LB_HALO_PUSH, LB_HALO_DIRECT, LB_HALO_REMOTE, LB_HALO_LL = 1, 2, 3, 4, 5


class LbHalo(C.Structure):
    _fields_ = [("mode", C.c_int32), ("epoch", C.c_uint32),
                ("send_hi", C.c_void_p * 2), ("send_lo", C.c_void_p * 2),
                ("recv_lo", C.c_void_p * 3), ("recv_hi", C.c_void_p * 3),
                ("peer_up_recv_lo", C.c_void_p * 3), ("peer_down_recv_hi", C.c_void_p * 3),
                ("peer_up_f", C.c_void_p * 2), ("peer_down_f", C.c_void_p * 2),
                ("signal_up", C.c_void_p), ("signal_down", C.c_void_p), ("wait_slots", C.c_void_p),
                ("flags", C.c_uint32), ("_pad", C.c_uint32)]


LB_HALO_F_NO_UNPACK, LB_HALO_F_CONSUME_FIRST = 1, 2


class LbFields(C.Structure):
    _fields_ = [("f", C.c_void_p), ("f_next", C.c_void_p), ("no_collision_mask", C.c_void_p),
                ("streaming_bits", C.c_void_p), ("boundaries", C.POINTER(LbBoundary)),
                ("n_boundaries", C.c_int32), ("_pad", C.c_int32), ("obs_out", C.c_void_p),
                ("obs_scratch", C.c_void_p), ("u_out", C.c_void_p), ("halo", C.POINTER(LbHalo))]


_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raise loudly when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeLibraryMissing(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(or `make -C lettuce_b200/csrc`). lettuce_b200 has no CPU / torch fallback.")
    L = C.CDLL(LIB_PATH)
    L.lb_abi_version.restype = C.c_int
    L.lb_last_error.restype = C.c_char_p
    L.lb_step.argtypes = [C.POINTER(LbDesc), C.POINTER(LbFields), C.c_int, C.c_void_p]
    L.lb_collide.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p,
                             C.POINTER(LbBoundary), C.c_int, C.c_void_p]
    L.lb_stream.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_pack_streaming_mask.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_moments.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_output_fields.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_double, C.c_double, C.c_double,
                                   C.c_void_p, C.c_void_p]
    L.lb_output_fields.restype = C.c_int
    L.lb_initialize.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                C.c_void_p]
    L.lb_obs_scratch_bytes.argtypes = [C.POINTER(LbDesc)]
    L.lb_obs_scratch_bytes.restype = C.c_size_t
    L.lb_observables.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p]
    L.lb_enstrophy.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_run_host.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p,
                              C.POINTER(LbBoundary), C.c_int, C.c_int, C.c_void_p]
    L.lb_unpack_halo.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_unpack_halo_ex.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                    C.c_void_p, C.c_void_p]
    L.lb_halo_ll_bytes.argtypes = [C.POINTER(LbDesc)]
    L.lb_halo_ll_bytes.restype = C.c_size_t
    L.lb_enstrophy_slab.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_zface_count.argtypes = [C.c_int]
    L.lb_static_obs_max_nodes.argtypes = []
    L.lb_static_obs_max_nodes.restype = C.c_size_t
    L.lb_observables_static.argtypes = [C.POINTER(LbDesc), C.c_void_p, C.c_void_p, C.c_void_p]
    L.lb_observables_static.restype = C.c_int
    L.lb_plan_resident.argtypes = [C.POINTER(LbDesc), C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_int)]
    L.lb_plan_resident.restype = C.c_int
    L.lb_ipc_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    L.lb_ipc_free.argtypes = [C.c_void_p]
    L.lb_ipc_export.argtypes = [C.c_void_p, C.c_char_p]
    L.lb_ipc_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.lb_ipc_close.argtypes = [C.c_void_p]
    L.lb_stencil_q.argtypes = [C.c_int]
    L.lb_bytes_per_update.argtypes = [C.POINTER(LbDesc)]
    L.lb_bytes_per_update.restype = C.c_size_t
    L.lb_vector_width.argtypes = [C.POINTER(LbDesc), C.c_int]
    for name in ("lb_step", "lb_collide", "lb_stream", "lb_pack_streaming_mask", "lb_moments", "lb_initialize",
                 "lb_observables", "lb_enstrophy", "lb_run_host", "lb_stencil_q", "lb_vector_width",
                 "lb_unpack_halo", "lb_unpack_halo_ex", "lb_enstrophy_slab", "lb_zface_count", "lb_ipc_alloc", "lb_ipc_free", "lb_ipc_export",
                 "lb_ipc_open", "lb_ipc_close"):
        getattr(L, name).restype = C.c_int
    if L.lb_abi_version() != ABI_VERSION:
        raise NativeError(f"ABI mismatch: library {L.lb_abi_version()} vs binding {ABI_VERSION}")
    _lib = L
    return L


def check(rc: int):
    if rc != 0:
        raise NativeError(f"lettuce_b200 native call failed ({rc}): {lib().lb_last_error().decode()}")


def make_desc(stencil_name: str, dtype: torch.dtype, resolution: Sequence[int], collision: int = LB_COLL_NONE,
              tau: float = 1.0, tau_minus: float = 1.0, force: int = LB_FORCE_NONE,
              acceleration: Sequence[float] = (), force_tau: float = 1.0, kernel_flags: int = 0) -> LbDesc:
    if stencil_name not in STENCIL_IDS:
        raise NativeError(f"stencil {stencil_name} has no native kernel (D2Q9, D3Q19, D3Q27 only)")
    if dtype not in DTYPE_IDS:
        raise NativeError(f"dtype {dtype} has no native kernel (float32 / float64 only)")
    d = LbDesc()
    d.stencil, d.dtype, d.collision = STENCIL_IDS[stencil_name], DTYPE_IDS[dtype], collision
    d.ndim = len(resolution)
    for i in range(3):
        d.dims[i] = int(resolution[i]) if i < len(resolution) else 1
    d.tau, d.tau_minus = float(tau), float(tau_minus)
    d.force, d.force_tau = int(force), float(force_tau)
    d.kernel_flags = int(kernel_flags)
    if force != LB_FORCE_NONE and len(acceleration) != len(resolution):
        raise NativeError(f"acceleration has {len(acceleration)} components on a {len(resolution)}-D lattice")
    for i, a in enumerate(acceleration):
        d.acceleration[i] = float(a)
    return d


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def current_stream_ptr(device) -> Optional[int]:
    if torch.device(device).type != "cuda":
        return None
    return torch.cuda.current_stream(device).cuda_stream


def broadcast_strides(t: torch.Tensor, lead: int, resolution: Sequence[int]):
    """Element strides of `t` viewed against the lattice: t has shape [*lead_dims, *dims] with
    each of the trailing dims either 1 or the resolution (or fewer trailing dims = broadcast)."""
    d = len(resolution)
    shape = list(t.shape[lead:])
    strides = list(t.stride()[lead:])
    if len(shape) > d:
        raise NativeError(f"boundary field of shape {tuple(t.shape)} does not broadcast to {resolution}")
    # torch broadcasting aligns trailing dims
    shape = [1] * (d - len(shape)) + shape
    strides = [0] * (d - len(strides)) + strides
    out = []
    for n, s, r in zip(shape, strides, resolution):
        if n == 1:
            out.append(0)
        elif n == r:
            out.append(s)
        else:
            raise NativeError(f"boundary field of shape {tuple(t.shape)} does not broadcast to {resolution}")
    return out + [0] * (3 - d)
