"""Framework-neutral counterparts of the reference's TensorFlow-side library (tensorflow_impl/):
numpy aggregator façade, numpy attacks, cluster-JSON parser and the gRPC pull service. TensorFlow
itself is not required (and not installed here): models/optimizers come from the torch side."""

from .aggregator_tf import Aggregator_tf  # noqa: F401
from .attacker import Attacker  # noqa: F401
from .network import Network  # noqa: F401
from .model import ModelManager  # noqa: F401
