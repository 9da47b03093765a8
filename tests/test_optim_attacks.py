"""Flat optimizers vs torch.optim and attack formulas (CPU; the CUDA kernels are checked by
tests/test_gpu_kernels.py against the same oracles)."""

import pytest
import torch

from garfield_b200.ops import attacks
from garfield_b200.ops.optim import FlatOptimizer

CONFIGS = [
    ("sgd", dict(lr=0.1)),
    ("sgd", dict(lr=0.2, momentum=0.9, weight_decay=5e-4)),
    ("sgd", dict(lr=0.05, momentum=0.8, dampening=0.1, weight_decay=1e-3)),
    ("sgd", dict(lr=0.05, momentum=0.9, nesterov=True)),
    ("adam", dict(lr=1e-2, weight_decay=1e-2)),
    ("adamw", dict(lr=1e-2)),
    ("rmsprop", dict(lr=1e-2, momentum=0.5, weight_decay=1e-3)),
    ("rmsprop", dict(lr=1e-2)),
    ("adagrad", dict(lr=0.1, lr_decay=0.01, weight_decay=1e-3)),
]
TORCH = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW,
         "rmsprop": torch.optim.RMSprop, "adagrad": torch.optim.Adagrad}


def run_pair(name, kwargs, device, steps=5, d=1003):
    gen = torch.Generator().manual_seed(7)
    init = torch.randn(d, generator=gen)
    grads = [torch.randn(d, generator=gen) for _ in range(steps)]
    ref_p = torch.nn.Parameter(init.clone().double())
    ref = TORCH[name]([ref_p], **kwargs)
    flat = init.clone().to(device)
    ours = FlatOptimizer(flat, name, **kwargs)
    for g in grads:
        ref_p.grad = g.double()
        ref.step()
        ours.step(g.to(device))
    return flat.cpu().double(), ref_p.detach()


@pytest.mark.parametrize("name,kwargs", CONFIGS)
def test_flat_optimizer_matches_torch_cpu(name, kwargs):
    got, want = run_pair(name, kwargs, "cpu")
    torch.testing.assert_close(got, want, rtol=2e-5, atol=2e-6)


def test_optimizer_state_dict_roundtrip():
    a = FlatOptimizer(torch.zeros(8), "sgd", lr=0.1, momentum=0.9)
    a.step(torch.ones(8))
    b = FlatOptimizer(a.param.clone(), "sgd", lr=0.5, momentum=0.1)
    b.load_state_dict(a.state_dict())
    a.step(torch.ones(8)); b.step(torch.ones(8))
    assert torch.equal(a.param, b.param)
    assert b.hyper["lr"] == 0.1


def test_attack_formulas_cpu():
    torch.manual_seed(0)
    g = torch.randn(1000)
    r = attacks.random_like(g, seed=1)
    assert r.min() >= 0 and r.max() < 1 and abs(r.mean().item() - 0.5) < 0.05
    assert torch.equal(attacks.reverse(g), g * -100)
    dropped = attacks.drop(g, p=0.3, seed=3)
    zeroed = (dropped == 0)
    assert 0.2 < zeroed.float().mean().item() < 0.4
    assert torch.equal(dropped[~zeroed], g[~zeroed])
    assert torch.equal(attacks.drop(g, strict_reference=True), g)
    est = [torch.randn(1000) for _ in range(4)]
    stack = torch.stack(est)
    torch.testing.assert_close(attacks.lie(est), stack.mean(0) + 1.035 * stack.std(0))
    torch.testing.assert_close(attacks.empire(est), -10 * stack.mean(0))
    torch.testing.assert_close(attacks.lie(est[:1]), est[0])
    assert torch.isnan(attacks.lie(est[:1], strict_reference=True)).all()
