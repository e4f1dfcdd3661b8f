"""Reference-side binding of liblettuce_b200.so: the file a lettuce maintainer drops in as
`lettuce/cuda_native/b200.py`.  It depends on lettuce's OWN objects only
(`Simulation`, `Flow`, `Collision`, `Boundary` as found in PhiSpel/lettuce) plus ctypes and torch --
not on the `lettuce_b200` Python package -- and replaces the per-step seam
`Simulation._collide_and_stream(simulation)` (lettuce/_simulation.py:92-96,148,202; generated
counterpart: lettuce/cuda_native/_template.py:31-44).

    if self.context.use_native:                      # lettuce/_simulation.py:98-148 becomes
        from .cuda_native import b200
        self._collide_and_stream = b200.invoke

`build_call(simulation)` turns a lettuce Simulation into the C structures of include/lettuce_b200.h
(pure host logic; tests/test_integration_stub.py runs it against the unmodified reference on CPU
tensors and compares the structures with the ones lettuce_b200's own host mirror builds);
`invoke(simulation)` performs the call on the current CUDA stream and swaps f / f_next.
"""
import ctypes as C
import os

import torch

LIB_ENV = "LETTUCE_B200_LIB"          # path of liblettuce_b200.so (default: next to this file)
ABI_VERSION = 6                       # LB_ABI_VERSION of the include/lettuce_b200.h these structures mirror
_lib = None

LB_BND_BOUNCE_BACK, LB_BND_EQUILIBRIUM, LB_BND_ABB_OUTLET, LB_BND_EQ_OUTLET_P = 1, 2, 3, 4
STENCIL = {"D2Q9": 0, "D3Q19": 1, "D3Q27": 2}
COLLISION = {"NoCollision": 0, "BGKCollision": 1, "TRTCollision": 2, "KBCCollision": 3,
             "RegularizedCollision": 4, "SmagorinskyCollision": 5}
FORCE = {"Guo": 1, "ShanChen": 2}


class Desc(C.Structure):            # lb_desc
    _fields_ = [("stencil", C.c_int32), ("dtype", C.c_int32), ("collision", C.c_int32),
                ("ndim", C.c_int32), ("dims", C.c_int64 * 3), ("tau", C.c_double), ("tau_minus", C.c_double),
                ("acceleration", C.c_double * 3), ("force_tau", C.c_double), ("force", C.c_int32),
                ("kernel_flags", C.c_int32)]


class Boundary(C.Structure):        # lb_boundary
    _fields_ = [("kind", C.c_int32), ("direction", C.c_int32), ("rho", C.c_void_p), ("u", C.c_void_p),
                ("rho_stride", C.c_int64 * 3), ("u_stride", C.c_int64 * 3), ("u_comp_stride", C.c_int64),
                ("value", C.c_double)]


class Fields(C.Structure):          # lb_fields
    _fields_ = [("f", C.c_void_p), ("f_next", C.c_void_p), ("no_collision_mask", C.c_void_p),
                ("streaming_bits", C.c_void_p), ("boundaries", C.POINTER(Boundary)),
                ("n_boundaries", C.c_int32), ("_pad", C.c_int32), ("obs_out", C.c_void_p),
                ("obs_scratch", C.c_void_p), ("u_out", C.c_void_p), ("halo", C.c_void_p)]


def lib():
    global _lib
    if _lib is None:
        path = os.environ.get(LIB_ENV) or os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                       "liblettuce_b200.so")
        L = C.CDLL(path)
        L.lb_step.argtypes = [C.POINTER(Desc), C.POINTER(Fields), C.c_int, C.c_void_p]
        L.lb_pack_streaming_mask.argtypes = [C.POINTER(Desc), C.c_void_p, C.c_void_p, C.c_void_p]
        L.lb_last_error.restype = C.c_char_p
        L.lb_abi_version.restype = C.c_int
        if L.lb_abi_version() != ABI_VERSION:      # struct drift would corrupt the C structures silently
            raise RuntimeError(f"liblettuce_b200.so has ABI version {L.lb_abi_version()}, this binding was written "
                               f"for {ABI_VERSION}")
        _lib = L
    return _lib


def check_supported(simulation):
    """what lettuce_b200.Simulation refuses, refused here as well (instead of a silently different result)"""
    flow = simulation.flow
    eq = type(flow.equilibrium).__name__ if getattr(flow, "equilibrium", None) is not None else "QuadraticEquilibrium"
    if eq not in ("QuadraticEquilibrium", "QuadraticEquilibriumLessMemory"):
        raise NotImplementedError(f"equilibrium '{eq}' has no native kernel (the kernels evaluate the quadratic one)")
    if not flow.f.is_contiguous():
        raise ValueError("flow.f must be contiguous")
    outlets = [(i, b) for i, b in enumerate(simulation.boundaries[1:], start=1)
               if type(b).__name__ in ("AntiBounceBackOutlet", "EquilibriumOutletP")]
    if outlets and getattr(simulation.collision, "force", None) is not None:
        raise NotImplementedError("outlet boundaries together with a forcing scheme are not supported natively")
    for i, b in outlets:
        if min(flow.resolution) < 2:
            raise ValueError("outlet boundaries need at least 2 nodes along every axis")
        axis = abs(_outlet_direction(b))
        for j, other in outlets:
            # a lower-index outlet on another axis crosses this outlet's neighbour plane: the kernel takes the
            # neighbour's velocity from its incoming populations and cannot redo that outlet on it
            if j < i and abs(_outlet_direction(other)) != axis:
                raise NotImplementedError(f"outlet boundary {i}: the lower-index outlet {j} crosses its neighbour "
                                          f"plane; this ordering is not supported natively")


def make_desc(simulation) -> Desc:
    flow, coll = simulation.flow, simulation.collision
    name = type(coll).__name__
    if name in ("KBCCollision2D", "KBCCollision3D"):        # deprecated aliases (kbc_collision.py:169-180)
        name = "KBCCollision"
    d = Desc()
    d.stencil = STENCIL[type(flow.stencil).__name__]
    d.dtype = {torch.float32: 0, torch.float64: 1}[flow.f.dtype]
    d.collision = COLLISION[name]
    d.ndim = flow.stencil.d
    for i in range(3):
        d.dims[i] = int(flow.resolution[i]) if i < len(flow.resolution) else 1
    if name in ("KBCCollision", "RegularizedCollision"):
        tau = flow.units.relaxation_parameter_lu            # both read it from the units at their first call
    elif name == "TRTCollision":                            # (kbc_collision.py:97-99, regularized_collision.py:19)
        tau = coll.tau_plus
    else:
        tau = getattr(coll, "tau", 1.0)
    d.tau = float(tau)
    d.tau_minus = float(getattr(coll, "tau_minus", getattr(coll, "constant", 1.0)))   # Smagorinsky: its constant
    d.force_tau = 1.0
    force = getattr(coll, "force", None)
    if force is not None:                                   # BGKCollision(tau, force=Guo | ShanChen)
        if name != "BGKCollision":
            raise NotImplementedError("forcing schemes are native for BGKCollision only")
        d.force, d.force_tau = FORCE[type(force).__name__], float(force.tau)
        for i, a in enumerate(force.acceleration.tolist()):
            d.acceleration[i] = float(a)
    return d


def _strides(t, lead, resolution):
    """element strides of `t` (shape [*lead_dims, *dims], trailing dims 1 or the resolution) per lattice axis;
    0 = broadcast (equilibrium_boundary_pu.py:21-32; tests/boundary/test_equilibrium_bc_pu.py:125-148)"""
    d = len(resolution)
    shape, strides = list(t.shape[lead:]), list(t.stride()[lead:])
    if len(shape) > d:
        raise ValueError(f"boundary field of shape {tuple(t.shape)} does not broadcast to {resolution}")
    shape = [1] * (d - len(shape)) + shape
    strides = [0] * (d - len(strides)) + strides
    out = []
    for n, s, r in zip(shape, strides, resolution):
        if n != 1 and n != r:
            raise ValueError(f"boundary field of shape {tuple(t.shape)} does not broadcast to {resolution}")
        out.append(0 if n == 1 else s)
    return out + [0] * (3 - d)


def _outlet_direction(boundary) -> int:
    """+-(axis + 1) from the plane index the reference stores (anti_bounce_back_outlet.py:46-57)"""
    for axis, idx in enumerate(boundary.index):
        if not isinstance(idx, slice):
            return (axis + 1) if idx == -1 else -(axis + 1)
    raise ValueError("outlet boundary without a plane index")


def make_boundaries(simulation):
    """lb_boundary[] in the order of simulation.boundaries[1:] (= sorted by str, _simulation.py:57-58)"""
    flow = simulation.flow
    res = list(flow.resolution)
    bnds = simulation.boundaries[1:]
    arr = (Boundary * max(len(bnds), 1))()
    keep = []
    for k, b in enumerate(bnds):
        name = type(b).__name__
        if name == "BounceBackBoundary":
            arr[k].kind = LB_BND_BOUNCE_BACK
        elif name == "EquilibriumBoundaryPU":
            arr[k].kind = LB_BND_EQUILIBRIUM
            rho = flow.units.convert_pressure_pu_to_density_lu(b.pressure)       # converted every step (:28-29)
            u = flow.units.convert_velocity_to_lu(b.velocity)
            rho = rho.to(device=flow.f.device, dtype=flow.f.dtype).contiguous()
            u = u.to(device=flow.f.device, dtype=flow.f.dtype).contiguous()
            lead = max(0, rho.dim() - len(res))
            rs, us = _strides(rho, lead, res), _strides(u, 1, res)
            arr[k].rho, arr[k].u = rho.data_ptr(), u.data_ptr()
            for a in range(3):
                arr[k].rho_stride[a], arr[k].u_stride[a] = rs[a], us[a]
            arr[k].u_comp_stride = u.stride(0)
            keep += [rho, u]
        elif name == "AntiBounceBackOutlet":
            arr[k].kind, arr[k].direction = LB_BND_ABB_OUTLET, _outlet_direction(b)
        elif name == "EquilibriumOutletP":
            arr[k].kind, arr[k].direction = LB_BND_EQ_OUTLET_P, _outlet_direction(b)
            arr[k].value = float(b.rho_outlet)
        else:
            raise NotImplementedError(f"boundary '{name}' has no native kernel")
    return arr, len(bnds), keep


def build_call(simulation, pack=True):
    """(Desc, Fields, keepalive) for one lb_step call.  `pack=False` leaves a non-zero
    no_streaming_mask unpacked (host-side tests without a GPU)."""
    flow = simulation.flow
    check_supported(simulation)
    d = make_desc(simulation)
    fl = Fields()
    fl.f, fl.f_next = flow.f.data_ptr(), flow.f_next.data_ptr()
    keep = []
    if simulation.no_collision_mask is not None:
        fl.no_collision_mask = simulation.no_collision_mask.data_ptr()
        arr, n, keep = make_boundaries(simulation)
        fl.boundaries, fl.n_boundaries = C.cast(arr, C.POINTER(Boundary)), n
        keep.append(arr)
        nsm = simulation.no_streaming_mask
        if pack and nsm is not None and bool((nsm == 1).any()):
            bits = getattr(simulation, "_b200_bits", None)       # packed once; rebuild if the mask object changes
            if bits is None or bits[0] != (nsm.data_ptr(), nsm._version):
                words = torch.empty(list(flow.resolution), dtype=torch.int32, device=nsm.device)
                with torch.cuda.device(nsm.device):
                    rc = lib().lb_pack_streaming_mask(d, nsm.contiguous().data_ptr(), words.data_ptr(),
                                                      torch.cuda.current_stream(nsm.device).cuda_stream)
                if rc:
                    raise RuntimeError(lib().lb_last_error().decode())
                bits = simulation._b200_bits = ((nsm.data_ptr(), nsm._version), words)
            fl.streaming_bits = bits[1].data_ptr()
    return d, fl, keep


def invoke(simulation):
    """same contract as the generated module's invoke(simulation): one time step on the current stream,
    flow.f and flow.f_next swapped; no device-wide synchronisation (cf. _template.py:366)"""
    flow = simulation.flow
    d, fl, keep = build_call(simulation)
    simulation._b200_keep = keep        # the boundary tensors stay alive while the launch is in flight
    with torch.cuda.device(flow.f.device):      # the flow's device, whatever the current one is
        rc = lib().lb_step(d, fl, 1, torch.cuda.current_stream(flow.f.device).cuda_stream)
    if rc:
        raise RuntimeError(lib().lb_last_error().decode())
    flow.f, flow.f_next = flow.f_next, flow.f
