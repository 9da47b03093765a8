"""Reference-compatible top-level `tools` package (see /root/reference/pytorch_impl/libs/tools).

Thin alias of `garfield_b200.utils`; sub-modules `tools.misc`, `tools.pytorch`, `tools.cluster`
and `tools.access` resolve to the same objects.
"""

import sys as _sys

from garfield_b200.utils import *  # noqa: F401,F403
from garfield_b200.utils import misc, cluster, access, flat as pytorch  # noqa: F401
from garfield_b200.utils import __all__  # noqa: F401

for _name, _mod in (("misc", misc), ("cluster", cluster), ("access", access), ("pytorch", pytorch)):
    _sys.modules[__name__ + "." + _name] = _mod
