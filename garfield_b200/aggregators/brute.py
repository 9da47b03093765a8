"""Brute / MDA (minimum-diameter averaging).

Average of the subset of n−f gradients with the smallest diameter (largest pairwise
distance). `brute` = python-reference semantics (aggregators/brute.py:32-80: non-squared
distances, subsets with a non-finite distance skipped, first minimal subset in lexicographic
order wins); `native-brute` = native semantics (native/py_brute/brute.cpp:47-113: squared
distances, non-finite → max). `mda` is an alias of `brute`.

CUDA path: the C(n, n−f) subsets are enumerated inside one CTA by combinadic unranking (one
subset per thread) with a block-wide lexicographic arg-min — no host combination search.
"""

from . import register
from . import _host
from ._common import basic_list_error, byz_error, on_device, device_ops, native_cpu, host_distances


def _selection(gradients, f, native):
    dist = host_distances(gradients, squared=native)
    if native:  # non-finite distances count as the largest representable value, never skipped
        big = float("1.79e308")
        dist = [[big if v == float("inf") else v for v in row] for row in dist]
    subset = _host.mda_selection(dist, f)
    assert subset is not None, "Too many non-finite gradients: a non-Byzantine gradient must only contain finite coordinates"
    return subset


def _mean_of(gradients, subset):
    out = gradients[subset[0]].clone()
    for i in subset[1:]:
        out.add_(gradients[i])
    return out.div_(len(subset))


def aggregate(gradients, f, **kwargs):
    """MDA, python-reference semantics."""
    if on_device(gradients):
        return device_ops().mda(gradients, f=f, squared=False)
    return _mean_of(gradients, _selection(gradients, f, native=False))


def aggregate_native(gradients, f, **kwargs):
    """MDA, native semantics."""
    if on_device(gradients):
        return device_ops().mda(gradients, f=f, squared=True)
    native = native_cpu()
    if native is not None and native.supports(gradients):
        return native.brute.aggregate(gradients, f)
    return _mean_of(gradients, _selection(gradients, f, native=True))


def check(gradients, f, **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    return byz_error(gradients, f, lambda f: 2 * f + 1, lambda n: (n - 1) // 2)


def upper_bound(n, f, d):
    return (n - f) / (2 * f)


def influence(honests, attacks, f, **kwargs):
    gradients = honests + attacks
    subset = _selection(gradients, f, native=False)
    hits = sum(1 for i in subset if any(gradients[i] is a for a in attacks))
    return hits / (len(gradients) - f)


register("brute", aggregate, check, upper_bound, influence)
register("mda", aggregate, check, upper_bound, influence)
register("native-brute", aggregate_native, check, upper_bound)
