"""Bulyan over Multi-Krum.

* `bulyan`        — python-reference semantics (static scores, lower median), reference
                    aggregators/bulyan.py:31-84 incl. its quirk that scores are never updated
                    after a gradient is removed.
* `native-bulyan` — native semantics with pruned-distance score updates and upper median,
                    reference native/py_bulyan/bulyan.cpp:53-193.

Both: θ = n−2f−2 selection rounds produce a θ×n averaging-weight matrix W; phase 2 takes, per
coordinate, the mean of the β = θ−2f round-averages closest to their median. The CUDA path
never materialises the θ×d intermediate in HBM: one kernel forms the θ round averages of a
coordinate in registers (W·x) and runs the sorting network on them.
"""

import math

import torch

from . import register
from . import _host
from ._common import basic_list_error, byz_error, on_device, device_ops, native_cpu, host_distances


def _host_bulyan(gradients, f, m, native):
    dist = host_distances(gradients, squared=native)
    rounds = (_host.bulyan_rounds_native if native else _host.bulyan_rounds_python)(dist, f, m)
    weights = _host.weights_from_rounds(rounds, len(gradients), dtype=gradients[0].dtype)
    selected = _host.apply_weights(gradients, weights)
    fast = None
    backend = native_cpu()
    if not native and backend is not None and selected.device.type == "cpu" and selected.dtype in (torch.float32, torch.float64):
        fast = backend.coordwise(list(selected.unbind(0)), "averaged_median", selected.shape[0] - 2 * f)   # lower median, β = θ − 2f
    return fast if fast is not None else _host.bulyan_phase2(selected, f, upper=native)


def aggregate(gradients, f, m=None, **kwargs):
    """Bulyan, python-reference semantics."""
    if on_device(gradients):
        return device_ops().bulyan(gradients, f=f, m=m, native=False)
    return _host_bulyan(gradients, f, m, native=False)


def aggregate_native(gradients, f, m=None, **kwargs):
    """Bulyan, native semantics."""
    if on_device(gradients):
        return device_ops().bulyan(gradients, f=f, m=m, native=True)
    native = native_cpu()
    if native is not None and native.supports(gradients):
        return native.bulyan.aggregate(gradients, f, len(gradients) - f - 2 if m is None else m)
    return _host_bulyan(gradients, f, m, native=True)


def check(gradients, f, m=None, **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    err = byz_error(gradients, f, lambda f: 4 * f + 3, lambda n: (n - 3) // 4)
    if err is not None:
        return err
    if m is not None and (not isinstance(m, int) or m < 1 or m > len(gradients) - f - 2):
        return "Invalid number of selected gradients, got m = %r, expected 1 ≤ m ≤ %d" % (m, len(gradients) - f - 2)
    return None


def upper_bound(n, f, d):
    return 1 / math.sqrt(2 * (n - f + f * (n + f * (n - f - 2) - 2) / (n - 2 * f - 2)))


register("bulyan", aggregate, check, upper_bound=upper_bound)
register("native-bulyan", aggregate_native, check, upper_bound=upper_bound)
