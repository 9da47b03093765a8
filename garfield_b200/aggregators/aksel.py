"""Aksel: average of the c gradients closest (squared L2) to the coordinate-wise median.

c = (n+1)//2 in mode 'mid', n−f in mode 'n-f'. Parity: reference aggregators/aksel.py:24-111.
CUDA path: median kernel → fused distance-to-vector reduction → on-device top-c → weighted
average.
"""

from . import register
from . import _host
from ._common import basic_list_error, byz_error, on_device, device_ops


def _count(n, f, mode):
    if mode == "mid":
        return (n + 1) // 2
    if mode == "n-f":
        return n - f
    raise NotImplementedError(mode)


def _ranking(gradients, f, mode):
    n = len(gradients)
    med = _host.median_lower(_host.as_matrix(gradients))
    sq = [(i, (g - med).pow_(2).sum().item()) for i, g in enumerate(gradients)]
    sq.sort(key=lambda pair: pair[1])
    return sq, _count(n, f, mode)


def aggregate(gradients, f, mode="mid", **kwargs):
    """Aksel rule."""
    if on_device(gradients):
        return device_ops().aksel(gradients, c=_count(len(gradients), f, mode))
    ranking, c = _ranking(gradients, f, mode)
    out = gradients[ranking[0][0]].clone()
    for i, _ in ranking[1:c]:
        out.add_(gradients[i])
    return out.div_(c)


def check(gradients, f, mode="mid", **kwargs):
    err = basic_list_error(gradients)
    if err is not None:
        return err
    err = byz_error(gradients, f, lambda f: 2 * f + 1, lambda n: (n - 1) // 2)
    if err is not None:
        return err
    if mode not in ("mid", "n-f"):
        return "Invalid operation mode %r" % (mode,)
    return None


def influence(honests, attacks, f, mode="mid", **kwargs):
    gradients = honests + attacks
    ranking, c = _ranking(gradients, f, mode)
    hits = sum(1 for i, _ in ranking[:c] if any(gradients[i] is a for a in attacks))
    return hits / c


register("aksel", aggregate, check, influence=influence)
