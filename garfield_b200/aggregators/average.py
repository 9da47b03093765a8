"""Plain averaging (not Byzantine-resilient). Parity: reference aggregators/average.py:21-55."""

from . import register
from ._common import basic_list_error, on_device, device_ops


def aggregate(gradients, **kwargs):
    """Arithmetic mean of the gradients: (1/n) Σ g_i."""
    if on_device(gradients):
        return device_ops().average(gradients)
    total = gradients[0].clone()
    for g in gradients[1:]:
        total.add_(g)
    return total.div_(len(gradients))


def check(gradients, **kwargs):
    return basic_list_error(gradients)


def influence(honests, attacks, **kwargs):
    """Every input is accepted, so the attack ratio is f/n."""
    return len(attacks) / (len(honests) + len(attacks))


register("average", aggregate, check, influence=influence)
