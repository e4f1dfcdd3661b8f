"""Where does the stock lt.Obstacle step spend its time?  The same 3-D channel with boundary lists that
add one ingredient at a time (A/B timing, device resident, CUDA events).

    python scripts/obstacle_variants.py [D3Q19|D3Q27|D2Q9]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lettuce_b200 as lt  # noqa: E402
from scripts.bench_configs import timed, PEAK  # noqa: E402

st_name = sys.argv[1] if len(sys.argv) > 1 else "D3Q19"
ST = getattr(lt, st_name)
RES = [2048, 1024] if st_name == "D2Q9" else [384, 192, 192]
STEPS = int(os.environ.get("STEPS", 40))


class NsmBounceBack(lt.BounceBackBoundary):
    """bounce-back that ALSO hands in a (here: empty or sparse) no-streaming mask -> masked-streaming kernel"""

    def __init__(self, mask, nsm):
        super().__init__(mask)
        self._nsm = nsm

    def make_no_streaming_mask(self, shape, context):
        return self._nsm


class Variant(lt.Obstacle):
    which = "stock"

    @property
    def boundaries(self):
        stock = lt.Obstacle.boundaries.fget(self)   # [inlet, ABB outlet, bounce-back]
        inlet, outlet, bb = stock
        w = self.which
        if w == "stock":
            return stock
        if w == "no_outlet":
            return [inlet, bb]
        if w == "empty_nsm":
            nsm = torch.zeros([self.stencil.q, *self.resolution], dtype=torch.bool, device=self.context.device)
            return [inlet, NsmBounceBack(self.mask, nsm)]
        if w == "plane_nsm":   # the outlet's no-streaming mask without the outlet itself
            nsm = outlet.make_no_streaming_mask([self.stencil.q, *self.resolution], self.context)
            return [inlet, NsmBounceBack(self.mask, nsm)]
        if w == "eq_outlet_p":
            return [inlet, lt.EquilibriumOutletP(outlet.direction, self), bb]
        if w == "outlet_only":
            return [outlet]
        raise ValueError(w)


def main():
    for which in ("no_outlet", "empty_nsm", "plane_nsm", "stock", "eq_outlet_p", "outlet_only"):
        ctx = lt.Context(device="cuda", dtype=torch.float32, use_native=True)
        Variant.which = which
        flow = Variant(ctx, RES, reynolds_number=300, mach_number=0.01, domain_length_x=2, stencil=ST)
        g = flow.grid
        r2 = sum((c - (0.3 if i == 0 else 0.5) * c.max()) ** 2 for i, c in enumerate(g))
        flow.mask = r2 < (0.1 * g[1].max()) ** 2
        sim = lt.Simulation(flow, lt.BGKCollision(flow.units.relaxation_parameter_lu), [])
        t = timed(sim, STEPS)
        n = flow.f[0].numel()
        gbs = n * STEPS * 2 * flow.stencil.q * 4 / t / 1e9
        print(json.dumps({"variant": which, "stencil": st_name, "us_per_step": round(t / STEPS * 1e6, 1),
                          "frac": round(gbs / PEAK, 3)}), flush=True)
        del flow, sim


if __name__ == "__main__":
    main()
