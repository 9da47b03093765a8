"""Shared pytest fixtures.

* `gpu` marker: tests that need a real B200 (`pytest -m gpu`); everything else runs on CPU.
* `ref`: the reference's pure-torch aggregators imported read-only from /root/reference as an
  oracle (skipped when the checkout is absent, e.g. on the GPU box).
"""

import importlib
import os
import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

REFERENCE_LIBS = pathlib.Path("/root/reference/pytorch_impl/libs")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running CPU test")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def _load_reference_aggregators():
    """Import the reference `aggregators` (and its `tools`) under private names, leaving this
    repo's same-named top-level packages untouched in sys.modules."""
    if not (REFERENCE_LIBS / "aggregators" / "__init__.py").exists():
        return None
    names = ("tools", "aggregators", "native")
    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in names}
    for k in saved:
        del sys.modules[k]
    sys.modules["native"] = None  # the reference natives need a torch-header JIT build: skip them
    sys.path.insert(0, str(REFERENCE_LIBS))
    stdout, stderr, hook = sys.stdout, sys.stderr, sys.excepthook
    try:
        ref = importlib.import_module("aggregators")
        ref_tools = sys.modules.get("tools")
    finally:
        sys.path.remove(str(REFERENCE_LIBS))
        sys.stdout, sys.stderr, sys.excepthook = stdout, stderr, hook  # the reference wraps them on import
        for k in [k for k in sys.modules if k.split(".")[0] in names]:
            del sys.modules[k]
        sys.modules.update(saved)
    ref._tools = ref_tools
    return ref


_ref_cache = {}


@pytest.fixture(scope="session")
def ref():
    if "v" not in _ref_cache:
        try:
            _ref_cache["v"] = _load_reference_aggregators()
        except Exception as err:  # pragma: no cover
            _ref_cache["v"] = None
            _ref_cache["err"] = err
    if _ref_cache["v"] is None:
        pytest.skip("reference checkout not importable: %s" % _ref_cache.get("err", "absent"))
    return _ref_cache["v"]
