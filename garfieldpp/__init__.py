"""Reference-compatible top-level `garfieldpp` package (alias of garfield_b200.garfieldpp).

    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.byzServer import ByzServer
    from garfieldpp import tools, datasets, models
"""

import sys as _sys

import garfield_b200.garfieldpp as _impl
from garfield_b200.garfieldpp import tools, datasets, worker, server, byzWorker, byzServer  # noqa: F401
import garfield_b200.models as models  # noqa: F401

for _name in ("tools", "datasets", "worker", "server", "byzWorker", "byzServer"):
    _sys.modules[__name__ + "." + _name] = getattr(_impl, _name)
_sys.modules[__name__ + ".models"] = models
