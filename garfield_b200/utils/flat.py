"""Flat-buffer helpers: make a set of tensors views of one contiguous vector, gradient access,
gradient statistics and a PGM/PPM image dump.

Parity with /root/reference/pytorch_impl/libs/tools/pytorch.py:27-158. `flatten`/`relink` are the
seed of this framework's arena design (garfield_b200.parallel.arena): parameters, gradients and
optimizer state are views into flat device buffers so that aggregation and the fused optimizer
operate on one d-length vector with no torch.cat / narrow copies.
"""

from __future__ import annotations

import math
import types

import torch

from .console import UserException

__all__ = ["relink", "flatten", "grad_of", "grads_of", "compute_avg_dev_max", "pnm"]


def relink(tensors, common):
    """Re-point every tensor's `.data` at consecutive segments of the flat tensor `common`."""
    tensors = tuple(tensors) if isinstance(tensors, types.GeneratorType) else tensors
    offset = 0
    for t in tensors:
        n = t.numel()
        t.data = common[offset:offset + n].view(t.shape)
        offset += n
    common.linked_tensors = tensors
    return common


def flatten(tensors):
    """Concatenate the tensors into one new flat tensor and relink them onto it."""
    tensors = tuple(tensors) if isinstance(tensors, types.GeneratorType) else tuple(tensors)
    flat = torch.cat([t.detach().reshape(-1) for t in tensors])
    return relink(tensors, flat)


def grad_of(tensor):
    """`tensor.grad`, creating (and attaching) a zero gradient when missing."""
    if tensor.grad is None:
        tensor.grad = torch.zeros_like(tensor)
    return tensor.grad


def grads_of(tensors):
    return (grad_of(t) for t in tensors)


def compute_avg_dev_max(samples):
    """(mean gradient, ‖mean‖, sqrt(Σ‖g−mean‖²/(k−1)), max|mean|) over k gradient samples."""
    k = len(samples)
    if k == 0:
        return None, math.nan, math.nan, math.nan
    stacked = torch.stack([s.detach() for s in samples])
    mean = stacked.mean(dim=0)
    dev = math.nan
    if k >= 2:
        dev = math.sqrt((stacked - mean).pow(2).sum().item() / (k - 1))
    return mean, mean.norm().item(), dev, mean.abs().max().item()


def pnm(fd, tn):
    """Write a 2-D tensor as binary PGM (1 = black) or a 3×H×W tensor as binary PPM."""
    original = tuple(tn.shape)
    while tn.dim() > 3 and tn.shape[0] == 1:
        tn = tn[0]
    if tn.dim() == 3 and tn.shape[0] == 1:
        tn = tn[0]
    if tn.dim() == 3:
        if tn.shape[0] != 3:
            raise UserException("Expected 3 color channels for the first dimension of a 3D tensor, got %d channels" % tn.shape[0])
        fd.write(("P6\n%d %d 255\n" % (tn.shape[1], tn.shape[2])).encode())
        pixels = tn.permute(1, 2, 0).mul(256).clamp(0., 255.).to(torch.uint8).contiguous()
        fd.write(pixels.cpu().numpy().tobytes())
        return
    if tn.dim() == 2:
        fd.write(("P5\n%d %d 255\n" % (tn.shape[0], tn.shape[1])).encode())
        pixels = (1.0 - tn).mul(256).clamp(0., 255.).to(torch.uint8).contiguous()
        fd.write(pixels.cpu().numpy().tobytes())
        return
    raise UserException("Expected a 2D or 3D tensor, got %d dimensions %r" % (len(original), original))
