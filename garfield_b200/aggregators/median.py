"""Coordinate-wise median, trimmed mean and mean-around-median.

* `median`          — `torch.median` semantics (lower median, NaN-propagating), reference
                      aggregators/median.py:31-39.
* `native-median`   — NaN/inf-resilient: per coordinate the ⌊k/2⌋-th of the k finite values, 0 if
                      none (reference native/py_median/median.cpp:55-72, median.cu:60-83).
* `trimmed_mean`    — drop the f smallest and f largest per coordinate, average the rest (new;
                      SURVEY.md §2.7 note).
* `averaged_median` — mean of the β = n−f values closest to the coordinate median (TF-only in
                      the reference: tensorflow_impl/rsrcs/aggregators/averaged-median.py:30-67).

On CUDA tensors all four are one launch of the `coordwise_select` sorting-network kernel.
"""

import math

from . import register
from . import _host
from ._common import basic_list_error, on_device, device_ops, native_cpu


def _host_fast(gradients, rule, arg=0):
    """The vectorised C++ kernel for CPU fp32/fp64 lists (None when unavailable or when it declines the input)."""
    native = native_cpu()
    if native is None or not isinstance(gradients, (list, tuple)) or not native.supports(gradients):
        return None
    if any(g.dim() != 1 or not g.is_contiguous() for g in gradients):
        return None
    return native.coordwise(list(gradients), rule, arg)


def aggregate(gradients, **kwargs):
    """Coordinate-wise (lower) median."""
    if on_device(gradients):
        return device_ops().coordwise(gradients, "median_lower")
    fast = _host_fast(gradients, "median_lower")
    return fast if fast is not None else _host.median_lower(_host.as_matrix(gradients))


def aggregate_native(gradients, **kwargs):
    """Non-finite-resilient coordinate-wise (upper) median."""
    if on_device(gradients):
        return device_ops().coordwise(gradients, "median_finite_upper")
    native = native_cpu()
    if native is not None and native.supports(gradients):
        return native.median.aggregate(gradients)
    return _host.median_finite_upper(_host.as_matrix(gradients))


def aggregate_trimmed(gradients, f, **kwargs):
    """Coordinate-wise f-trimmed mean."""
    if on_device(gradients):
        return device_ops().coordwise(gradients, "trimmed_mean", f=f)
    fast = _host_fast(gradients, "trimmed_mean", f)
    return fast if fast is not None else _host.trimmed_mean(_host.as_matrix(gradients), f)


def aggregate_avgmed(gradients, f, beta=None, **kwargs):
    """Mean of the β (default n−f) coordinates closest to the coordinate-wise median."""
    if beta is None:
        beta = len(gradients) - f
    if on_device(gradients):
        return device_ops().coordwise(gradients, "averaged_median", beta=beta)
    fast = _host_fast(gradients, "averaged_median", beta)
    return fast if fast is not None else _host.averaged_median(_host.as_matrix(gradients), beta)


def check(gradients, **kwargs):
    return basic_list_error(gradients)


def check_trimmed(gradients, f, **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    if not isinstance(f, int) or f < 0 or len(gradients) < 2 * f + 1:
        return "Invalid number of Byzantine gradients to tolerate, got f = %r, expected 0 ≤ f ≤ %d" % (
            f, (len(gradients) - 1) // 2)
    return None


def check_avgmed(gradients, f, beta=None, **kwargs):
    err = check_trimmed(gradients, f)
    if err is not None:
        return err
    if beta is not None and (not isinstance(beta, int) or beta < 1 or beta > len(gradients)):
        return "Invalid number of averaged coordinates, got beta = %r, expected 1 ≤ beta ≤ %d" % (beta, len(gradients))
    return None


def upper_bound(n, f, d):
    """Bound on (honest std-dev / gradient norm) for the median to be usable: 1/√(n−f)."""
    return 1 / math.sqrt(n - f)


register("median", aggregate, check, upper_bound=upper_bound)
register("native-median", aggregate_native, check, upper_bound=upper_bound)
register("trimmed_mean", aggregate_trimmed, check_trimmed, upper_bound=upper_bound)
register("averaged_median", aggregate_avgmed, check_avgmed, upper_bound=upper_bound)
