"""garfield_b200.models — the model zoo (reference: garfieldpp/models)."""

from .small import *  # noqa: F401,F403
from .cifar_zoo import *  # noqa: F401,F403
from .native_nets import *  # noqa: F401,F403
from . import small as _small, cifar_zoo as _zoo, native_nets as _native

__all__ = list(_small.__all__) + list(_zoo.__all__) + list(_native.__all__)
