"""lettuce_b200 -- a B200-native engine for lettuce's per-timestep stream-collide hot path,
behind lettuce's own API (`Context`, `Stencil`, `Flow`, `Collision`, `Boundary`, `Reporter`,
`Simulation`).  Use it as `import lettuce_b200 as lt`.

The stepping itself is hand-written sm_100a CUDA behind a C ABI (include/lettuce_b200.h,
lettuce_b200/csrc); this package is the host-side mirror of the reference interface.
Nothing in here falls back to torch or the CPU for the time step."""
from ._context import *
from ._stencil import *
from ._unit import *
from ._flow import *
from ._equilibrium import *
from ._collision import *
from ._force import *
from ._boundary import *
from ._flows import *
from ._reporter import *
from ._output import *
from ._simulation import *
from .util import *
from . import _native as native

__version__ = "0.1.0"
