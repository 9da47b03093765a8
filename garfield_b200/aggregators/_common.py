"""Shared argument checks and backend dispatch for the rule modules."""

from __future__ import annotations

import torch

from . import _host

__all__ = ["basic_list_error", "byz_error", "on_device", "device_ops", "native_cpu"]


def basic_list_error(gradients):
    if not isinstance(gradients, list) or len(gradients) < 1:
        return "Expected a list of at least one gradient to aggregate, got %r" % (gradients,)
    return None


def byz_error(gradients, f, min_n, max_f):
    """Common `1 ≤ f` and `n ≥ min_n(f)` validation; `max_f(n)` is only used for the message."""
    if not isinstance(f, int) or isinstance(f, bool) or f < 1 or len(gradients) < min_n(f):
        return "Invalid number of Byzantine gradients to tolerate, got f = %r, expected 1 ≤ f ≤ %d" % (
            f, max_f(len(gradients)))
    return None


def device_ops():
    """The CUDA backend module, or None when it cannot be used in this process."""
    try:
        from ..ops import gar as _gar
    except Exception:  # pragma: no cover - import problems are reported by ops itself
        return None
    return _gar if _gar.available() else None


def on_device(gradients):
    """True when the inputs are CUDA float tensors and the sm_100a library is loaded."""
    g0 = gradients[0]
    if not (isinstance(g0, torch.Tensor) and g0.is_cuda):
        return False
    if g0.dtype not in (torch.float32, torch.float64) or len(gradients) > 64:
        return False  # more than 64 rows (and other dtypes) take the (device-resident) torch path
    if any(g.dtype != g0.dtype for g in gradients):
        return False
    return device_ops() is not None


def native_cpu():
    """The C++ host library bindings (garfield_b200.native), or None."""
    try:
        from .. import native as _native
    except Exception:
        return None
    return _native if _native.available() else None


def host_distances(gradients, squared):
    """n×n list-of-lists of (squared) L2 distances with +inf on the diagonal and for non-finite values — from the C++ host
    library (blocked, multi-threaded, one pass) when it can take the inputs, else the direct torch formulation."""
    import math
    from . import _host
    native = native_cpu()
    if (native is not None and isinstance(gradients, (list, tuple)) and native.supports(gradients)
            and all(g.dim() == 1 and g.is_contiguous() for g in gradients)):
        n = len(gradients)
        matrix = native.pairwise_sqdist(list(gradients)).tolist()
        for i in range(n):
            for j in range(n):
                v = matrix[i][j]
                if i == j or not math.isfinite(v):
                    matrix[i][j] = math.inf
                elif not squared:
                    matrix[i][j] = math.sqrt(v)
        return matrix
    return _host.pairwise_distances(gradients, squared=squared)
