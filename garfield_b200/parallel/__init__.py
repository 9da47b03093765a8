"""garfield_b200.parallel — device-resident data plane: flat arenas, peer-memory transport and
the topology step engines (SSMW / MSMW / decentralized)."""

from .arena import FlatModel, GradMatrix, flat_size  # noqa: F401
