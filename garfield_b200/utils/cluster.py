"""Cluster specification parsing: JSON `{"ps": [...], "workers": [...]}` or a named special parser.

Parity with /root/reference/pytorch_impl/libs/tools/cluster.py:27-73 (Grid5000 `OAR_FILE_NODES`
node-file parser + JSON), plus a `torchrun` parser that derives the cluster from the
`WORLD_SIZE`/`MASTER_ADDR` environment of a one-process-per-GPU launch.
"""

from __future__ import annotations

import json
import os
import pathlib

from .console import UserException

__all__ = ["cluster_parse", "cluster_parsers"]

_G5K_KEY = "OAR_FILE_NODES"
_cache = {}


def _g5k_parser():
    """First distinct node of the OAR node file is the ps, the others are workers (port 7000)."""
    if "g5k" not in _cache:
        if _G5K_KEY not in os.environ:
            raise UserException("Key %r not found in environment; are you running on Grid5000?" % _G5K_KEY)
        lines = pathlib.Path(os.environ[_G5K_KEY]).read_text().split()
        hosts = list(dict.fromkeys(lines))  # order-preserving de-duplication
        nodes = [host + ":7000" for host in hosts]
        _cache["g5k"] = {"ps": nodes[:1], "workers": nodes[1:]}
    return _cache["g5k"]


def _torchrun_parser():
    """One ps on rank 0's address, every rank a worker (roles are co-located per GPU)."""
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = int(os.environ.get("MASTER_PORT", "29500"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return {"ps": ["%s:%d" % (addr, port)], "workers": ["%s:%d" % (addr, port + 1 + r) for r in range(world)]}


_PARSERS = {"G5k": _g5k_parser, "torchrun": _torchrun_parser}
cluster_parsers = ", ".join(repr(name) for name in _PARSERS)


def cluster_parse(text):
    """Return the cluster dict for a special parser name or a JSON string."""
    if text in _PARSERS:
        return _PARSERS[text]()
    return json.loads(text)
