"""Build lettuce_b200 flows / simulations from golden cases or raw fields (GPU tests, smoke)."""
import torch

import lettuce_b200 as lt
from golden_util import DT


class FieldFlow(lt.ExtFlow):
    """A flow whose initial populations are given explicitly (cf. the reference tests'
    DummyFlow, reference/tests/conftest.py:243-266)."""

    def __init__(self, context, resolution, stencil, f0, re=100.0, ma=0.05):
        self._f0 = f0
        self._boundaries = []
        lt.ExtFlow.__init__(self, context, list(resolution), re, ma, stencil)

    def make_resolution(self, resolution, stencil=None):
        return resolution

    def make_units(self, reynolds_number, mach_number, resolution):
        return lt.UnitConversion(reynolds_number, mach_number, characteristic_length_lu=resolution[0])

    def initial_pu(self):
        ...

    def initialize(self):
        self.f = self.context.convert_to_tensor(self._f0).clone().contiguous()

    @property
    def boundaries(self):
        return self._boundaries


class LUEquilibriumBoundary(lt.EquilibriumBoundaryPU):
    """Equilibrium boundary whose rho / u are given directly in lattice units (the golden files
    store the values the reference converted)."""

    def __init__(self, context, mask, rho_lu, u_lu, nsm=None, tag=0):
        self._mask = mask
        self._rho = context.convert_to_tensor(rho_lu)
        self._u = context.convert_to_tensor(u_lu)
        self._nsm = nsm
        self._tag = tag

    def _lattice_units(self, flow):
        return self._rho, self._u

    def make_no_streaming_mask(self, shape, context):
        return self._nsm

    def __str__(self):          # keeps the golden index order under sorted(key=str)
        return f"boundary_{self._tag:03d}"


class TaggedBounceBack(lt.BounceBackBoundary):
    def __init__(self, mask, nsm=None, tag=0):
        super().__init__(mask)
        self._nsm = nsm
        self._tag = tag

    def make_no_streaming_mask(self, shape, context):
        return self._nsm

    def __str__(self):
        return f"boundary_{self._tag:03d}"


class TaggedABBOutlet(lt.AntiBounceBackOutlet):
    def __init__(self, direction, flow, tag=0):
        super().__init__(direction, flow)
        self._tag = tag

    def __str__(self):
        return f"boundary_{self._tag:03d}"


class TaggedEqOutletP(lt.EquilibriumOutletP):
    def __init__(self, direction, flow, rho_outlet=1.0, tag=0):
        super().__init__(direction, flow, rho_outlet=rho_outlet)
        self._tag = tag

    def __str__(self):
        return f"boundary_{self._tag:03d}"


def make_collision(meta, flow=None):
    c = meta["collision"]
    if c == "none":
        return lt.NoCollision()
    if c == "bgk":
        force = None
        if meta.get("force"):
            fo = meta["force"]
            force = (lt.Guo if fo["kind"] == "guo" else lt.ShanChen)(flow, fo["tau"], fo["acceleration"])
        return lt.BGKCollision(tau=meta["tau"], force=force)
    if c == "trt":
        return lt.TRTCollision(tau=meta["tau"], tau_minus=meta["tau_minus"])
    if c == "kbc":
        k = lt.KBCCollision()
        k.tau = meta["tau"]
        k.beta = 1. / (2 * meta["tau"])      # the golden meta stores the tau the reference used
        return k
    if c == "regularized":
        r = lt.RegularizedCollision()
        r.tau = meta["tau"]                  # as above: the tau the reference read from its units
        r._bound = True
        return r
    if c == "smagorinsky":
        return lt.SmagorinskyCollision(tau=meta["tau"], smagorinsky_constant=meta["tau_minus"])
    raise ValueError(c)


def simulation_from_case(case, reporters=(), device="cuda"):
    meta = case["meta"]
    dtype = DT[meta["dtype"]]
    ctx = lt.Context(device=device, dtype=dtype, use_native=True)
    stencil = getattr(lt, meta["stencil"])
    flow = FieldFlow(ctx, meta["res"], stencil, torch.from_numpy(case["f0"]))
    if "ncm" in case:
        ncm = torch.from_numpy(case["ncm"])
        nsm = torch.from_numpy(case["nsm"]).to(torch.bool)
        for k, b in enumerate(meta["boundaries"]):
            i = b["index"]
            mask = ncm == i
            this_nsm = nsm if k == 0 else None         # the OR of all masks, attached once
            if b["kind"] == "equilibrium":
                flow.boundaries.append(LUEquilibriumBoundary(
                    ctx, mask, torch.from_numpy(case[f"b{i}_rho"]), torch.from_numpy(case[f"b{i}_u"]),
                    nsm=this_nsm, tag=i))
            elif b["kind"] in ("abb_outlet", "eq_outlet_p"):
                # the real classes: their own masks must reproduce the reference's (checked below)
                cls = TaggedABBOutlet if b["kind"] == "abb_outlet" else TaggedEqOutletP
                kw = dict(rho_outlet=b["rho_outlet"]) if b["kind"] == "eq_outlet_p" else {}
                flow.boundaries.append(cls(b["direction"], flow, tag=i, **kw))
            else:
                if f"b{i}_mask" in case:          # outlet files store the bounce-back mask itself
                    mask = torch.from_numpy(case[f"b{i}_mask"])
                    this_nsm = None
                flow.boundaries.append(TaggedBounceBack(mask, nsm=this_nsm, tag=i))
    sim = lt.Simulation(flow, make_collision(meta, flow), list(reporters))
    return sim
