"""Byzantine attack generation on flat vectors (device kernels, torch fallback on CPU).

Formulas of SURVEY.md §2.9 (reference: garfieldpp/byzWorker.py:78-143, byzServer.py:86-107,
tensorflow_impl/libs/attacker.py:35-128). Constants are the reference's defaults but are
parameters here (Q10). `drop` implements the *intended* attack; `strict_reference=True`
reproduces the reference's no-op (Q5).
"""

from __future__ import annotations

import torch

from . import _lib

__all__ = ["random_like", "reverse", "drop", "lie", "empire", "WORKER_ATTACKS", "SERVER_ATTACKS"]

WORKER_ATTACKS = ("random", "reverse", "drop", "lie", "empire")
SERVER_ATTACKS = ("random", "reverse", "drop")

_counter = [0]


def _seed(seed):
    if seed is not None:
        return int(seed) & 0xFFFFFFFFFFFFFFFF
    _counter[0] += 1
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + _counter[0]) & 0x7FFFFFFFFFFFFFFF


def _dev(t):
    return t.is_cuda and t.dtype == torch.float32 and _lib.available()


def random_like(vec, seed=None):
    """U[0,1) vector of the same shape/device (replaces the real gradient/model)."""
    if _dev(vec):
        out = torch.empty_like(vec)
        with torch.cuda.device(vec.device):
            _lib.check(_lib.require().gb_attack_random(out.data_ptr(), out.numel(), _seed(seed), _lib.stream_ptr(vec.device)))
        _lib.count()
        return out
    gen = None if seed is None else torch.Generator(device=vec.device).manual_seed(int(seed))
    return torch.rand(vec.shape, dtype=vec.dtype, device=vec.device, generator=gen)


def reverse(vec, coeff=100.0):
    """−coeff · vec."""
    if _dev(vec):
        vec = vec.contiguous()
        out = torch.empty_like(vec)
        with torch.cuda.device(vec.device):
            _lib.check(_lib.require().gb_scale(vec.data_ptr(), out.data_ptr(), vec.numel(), -float(coeff), _lib.stream_ptr(vec.device)))
        _lib.count()
        return out
    return vec * -coeff


def drop(vec, p=0.3, seed=None, strict_reference=False):
    """Zero each coordinate with probability p (no-op copy when `strict_reference`)."""
    if strict_reference:
        return vec.clone()
    if _dev(vec):
        vec = vec.contiguous()
        out = torch.empty_like(vec)
        with torch.cuda.device(vec.device):
            _lib.check(_lib.require().gb_attack_drop(vec.data_ptr(), out.data_ptr(), vec.numel(), float(p), _seed(seed), _lib.stream_ptr(vec.device)))
        _lib.count()
        return out
    gen = None if seed is None else torch.Generator(device=vec.device).manual_seed(int(seed))
    mask = torch.rand(vec.shape, device=vec.device, generator=gen) < p
    return vec.masked_fill(mask, 0)


def _collude(estimates, kind, coeff, nan_single):
    first = estimates[0]
    if _dev(first):
        rows = [e.contiguous() for e in estimates]
        out = torch.empty_like(rows[0])
        with torch.cuda.device(first.device):
            _lib.check(_lib.require().gb_attack_collude(_lib.row_table(rows), len(rows), out.numel(), kind, float(coeff),
                                                        int(nan_single), out.data_ptr(), _lib.stream_ptr(first.device)))
        _lib.count()
        return out
    stack = torch.stack(list(estimates))
    mu = stack.mean(dim=0)
    if kind == 1:
        return -coeff * mu
    if len(estimates) > 1:
        sigma = stack.std(dim=0)
    else:
        sigma = torch.full_like(mu, float("nan")) if nan_single else torch.zeros_like(mu)
    return mu + coeff * sigma


def lie(estimates, z=1.035, strict_reference=False):
    """"A little is enough": μ + z·σ over the colluders' honest estimates (σ unbiased).
    With a single estimate σ := 0 (the reference yields NaN there; `strict_reference=True`)."""
    return _collude(estimates, 0, z, strict_reference)


def empire(estimates, eps=10.0):
    """"Fall of empires": −eps·μ over the colluders' honest estimates."""
    return _collude(estimates, 1, eps, False)
