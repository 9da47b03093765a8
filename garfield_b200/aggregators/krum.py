"""Multi-Krum.

* `krum`        — python-reference semantics: non-squared distances, score = Σ of the n−f−1
                  smallest, mean of the m best (default m = n−f−2); reference
                  aggregators/krum.py:30-80.
* `native-krum` — paper/native semantics: squared distances, score over the n−f−2 smallest;
                  reference native/py_krum/krum.cpp:50-118.

CUDA path: pairwise squared distances (TMA-bulk streamed, one pass over the n×d data) →
single-warp on-device scoring/selection → weighted average, no host round trip.
"""

import math

from . import register
from . import _host
from ._common import basic_list_error, byz_error, on_device, device_ops, native_cpu, host_distances


def _selection(gradients, f, m, native):
    n = len(gradients)
    if m is None:
        m = n - f - 2
    dist = host_distances(gradients, squared=native)
    return _host.krum_selection(dist, f, m, n - f - 2 if native else n - f - 1), m


def aggregate(gradients, f, m=None, **kwargs):
    """Multi-Krum, python-reference scoring."""
    n = len(gradients)
    if m is None:
        m = n - f - 2
    if on_device(gradients):
        return device_ops().krum(gradients, f=f, m=m, neighbors=n - f - 1, squared=False)
    chosen, m = _selection(gradients, f, m, native=False)
    out = gradients[chosen[0]].clone()
    for i in chosen[1:]:
        out.add_(gradients[i])
    return out.div_(m)


def aggregate_native(gradients, f, m=None, **kwargs):
    """Multi-Krum, native scoring (squared distances, n−f−2 neighbours)."""
    n = len(gradients)
    if m is None:
        m = n - f - 2
    if on_device(gradients):
        return device_ops().krum(gradients, f=f, m=m, neighbors=n - f - 2, squared=True)
    native = native_cpu()
    if native is not None and native.supports(gradients):
        return native.krum.aggregate(gradients, f, m)
    chosen, m = _selection(gradients, f, m, native=True)
    out = gradients[chosen[0]].clone()
    for i in chosen[1:]:
        out.add_(gradients[i])
    return out.div_(m)


def check(gradients, f, m=None, **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    err = byz_error(gradients, f, lambda f: 2 * f + 3, lambda n: (n - 3) // 2)
    if err is not None:
        return err
    if m is not None and (not isinstance(m, int) or m < 1 or m > len(gradients) - f - 2):
        return "Invalid number of selected gradients, got m = %r, expected 1 ≤ m ≤ %d" % (m, len(gradients) - f - 2)
    return None


def upper_bound(n, f, d):
    return 1 / math.sqrt(2 * (n - f + f * (n + f * (n - f - 2) - 2) / (n - 2 * f - 2)))


def influence(honests, attacks, f, m=None, **kwargs):
    """Fraction of the m selected gradients that are attack gradients (identity comparison)."""
    gradients = honests + attacks
    chosen, m = _selection(gradients, f, m, native=False)
    hits = sum(1 for i in chosen if any(gradients[i] is a for a in attacks))
    return hits / m


register("krum", aggregate, check, upper_bound, influence)
register("native-krum", aggregate_native, check, upper_bound)
