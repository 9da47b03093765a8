"""Condense: per coordinate take the median with probability p, else the first gradient's value.

Parity: reference aggregators/condense.py:25-76. The CUDA path draws the Bernoulli mask from a
counter-based Philox stream inside the median kernel (seed from `torch.initial_seed()` and a
call counter, or the explicit `seed=` keyword) so no mask tensor is ever written to HBM.
"""

import math

import torch

from . import register
from . import _host
from ._common import basic_list_error, byz_error, on_device, device_ops

_calls = [0]


def aggregate(gradients, f, p=0.9, seed=None, **kwargs):
    """Bernoulli(p) mix of the coordinate-wise median and gradients[0]."""
    if on_device(gradients) and gradients[0].dtype == __import__("torch").float32:
        if seed is None:
            _calls[0] += 1
            seed = (torch.initial_seed() * 0x9E3779B97F4A7C15 + _calls[0]) & 0x7FFFFFFFFFFFFFFF
        return device_ops().coordwise(gradients, "condense", p=float(p), seed=int(seed))
    generator = None
    if seed is not None:
        generator = torch.Generator(device=gradients[0].device).manual_seed(int(seed))
    keep = torch.rand(gradients[0].shape, dtype=gradients[0].dtype, device=gradients[0].device,
                      generator=generator) < p
    med = _host.median_lower(_host.as_matrix(gradients))
    return torch.where(keep, med, gradients[0])


def check(gradients, f, p=0.9, **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    err = byz_error(gradients, f, lambda f: 2 * f + 2, lambda n: (n - 2) // 2)
    if err is not None:
        return err
    if p <= 0 or p > 1:
        return "Expected positive selection probability, got %s" % (p,)
    return None


def upper_bound(n, f, d):
    return 1 / math.sqrt(n - f)


register("condense", aggregate, check, upper_bound=upper_bound)
