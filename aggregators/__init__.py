"""Reference-compatible top-level `aggregators` package (alias of garfield_b200.aggregators).

`import aggregators; aggregators.gars["krum"](gradients=[...], f=1)` works exactly as with
/root/reference/pytorch_impl/libs/aggregators.
"""

from garfield_b200.aggregators import *  # noqa: F401,F403
from garfield_b200.aggregators import gars, register, make_gar  # noqa: F401
import garfield_b200.aggregators as _impl

for _name, _rule in gars.items():
    globals()[_name] = _rule
    globals()[_name.replace("-", "_")] = _rule
