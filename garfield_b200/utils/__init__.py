"""garfield_b200.utils — the `tools` toolbox (console contexts, registries, flat-buffer helpers,
cluster parsing) plus framework services (checkpointing, metrics, tracing, config)."""

from .console import *  # noqa: F401,F403
from .console import __all__ as _console_all
from . import misc, flat, cluster, access  # noqa: F401
from .misc import *  # noqa: F401,F403
from .flat import *  # noqa: F401,F403
from .cluster import *  # noqa: F401,F403
from .access import *  # noqa: F401,F403

__all__ = list(_console_all) + misc.__all__ + flat.__all__ + cluster.__all__ + access.__all__
