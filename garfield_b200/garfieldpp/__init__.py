"""garfield_b200.garfieldpp — the role objects and their helpers (reference: libs/garfieldpp)."""

from . import tools, datasets, worker, server, byzWorker, byzServer  # noqa: F401
from .worker import Worker  # noqa: F401
from .server import Server  # noqa: F401
from .byzWorker import ByzWorker  # noqa: F401
from .byzServer import ByzServer  # noqa: F401
