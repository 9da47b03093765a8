"""Property-based checks of the aggregation rules on CPU (hypothesis): invariances every rule must have and the
resilience property the robust ones are built for. Complements the example-based oracle comparisons."""
import math

import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

import aggregators  # noqa: E402

ROBUST = ("median", "krum", "bulyan", "brute", "aksel", "trimmed_mean", "averaged_median", "native-median", "native-krum", "native-bulyan",
          "native-brute")
ALL = ("average",) + ROBUST


# MDA ("brute") is excluded: its ties are structural, not measure-zero — every subset containing the same farthest pair has
# the same diameter, and the reference breaks such ties by lexicographic order of the subset, i.e. by input order.
ORDER_FREE = tuple(name for name in ALL if "brute" not in name)


def _inputs(seed, n, d, dtype=torch.float32):
    gen = torch.Generator().manual_seed(seed)
    return [torch.randn(d, generator=gen, dtype=dtype) for _ in range(n)]


def _n_for(name, f):
    return 4 * f + 3 if "bulyan" in name else 2 * f + 3


@settings(max_examples=25, deadline=None, derandomize=True)
@given(seed=st.integers(0, 10_000), f=st.integers(1, 2), extra=st.integers(0, 3), d=st.integers(1, 40), name=st.sampled_from(ORDER_FREE))
def test_permutation_invariance(seed, f, extra, d, name):
    """The aggregate does not depend on the order in which the gradients are listed (ties aside: continuous random inputs)."""
    n = _n_for(name, f) + extra
    grads = _inputs(seed, n, d)
    rule = aggregators.gars[name]
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(seed + 1)).tolist()
    a = rule(gradients=grads, f=f)
    b = rule(gradients=[grads[i] for i in perm], f=f)
    assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(seed=st.integers(0, 10_000), f=st.integers(1, 2), d=st.integers(3, 40), name=st.sampled_from(ORDER_FREE),
       scale=st.floats(0.25, 8.0), shift=st.floats(-3.0, 3.0))
def test_translation_and_positive_scaling_equivariance(seed, f, d, name, scale, shift):
    """rule(a·g_i + b) = a·rule(g_i) + b for a > 0: distances scale uniformly, order statistics commute with monotone maps.
    (d >= 3: on a line, sums of neighbour distances tie exactly — |a−b| + |b−c| = |a−c| — and rounding then decides.)"""
    n = _n_for(name, f) + 1
    grads = _inputs(seed, n, d, torch.float64)
    rule = aggregators.gars[name]
    base = rule(gradients=grads, f=f)
    moved = rule(gradients=[g * scale + shift for g in grads], f=f)
    assert torch.allclose(moved, base * scale + shift, rtol=1e-9, atol=1e-9)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(seed=st.integers(0, 10_000), f=st.integers(1, 2), d=st.integers(2, 40), name=st.sampled_from(ROBUST), magnitude=st.floats(1e3, 1e9))
def test_f_arbitrary_inputs_cannot_drag_the_aggregate_away(seed, f, d, name, magnitude):
    """With f Byzantine inputs of arbitrary magnitude the aggregate stays within the box spanned by the honest gradients
    (coordinate-wise for the coordinate-wise rules; as a convex combination of honest inputs for the selection rules)."""
    n = _n_for(name, f) + 1
    honest = _inputs(seed, n - f, d, torch.float64)
    attack = [torch.full((d,), magnitude, dtype=torch.float64) * (1 if k % 2 == 0 else -1) for k in range(f)]
    out = aggregators.gars[name](gradients=honest + attack, f=f)
    stack = torch.stack(honest)
    lo, hi = stack.min(0).values, stack.max(0).values
    slack = 1e-9 * (1 + hi.abs().max().item())
    assert bool(((out >= lo - slack) & (out <= hi + slack)).all()), (name, float((out - hi).max()), float((lo - out).max()))


@settings(max_examples=20, deadline=None, derandomize=True)
@given(seed=st.integers(0, 10_000), f=st.integers(1, 2), d=st.integers(1, 30), name=st.sampled_from(ALL))
def test_unanimity(seed, f, d, name):
    """If every worker sends the same gradient, that gradient is the aggregate."""
    n = _n_for(name, f)
    g = _inputs(seed, 1, d)[0]
    out = aggregators.gars[name](gradients=[g.clone() for _ in range(n)], f=f)
    assert torch.allclose(out, g, rtol=1e-6, atol=1e-7)


def test_average_is_not_robust_but_declares_it():
    """The non-robust baseline moves arbitrarily far with one outlier — which is what `influence` reports (f/n)."""
    honest = _inputs(0, 7, 5)
    out = aggregators.gars["average"](gradients=honest + [torch.full((5,), 1e9)], f=1)
    assert out.abs().max().item() > 1e7
    assert math.isclose(aggregators.gars["average"].influence(honest, [torch.zeros(5)], f=1), 1 / 8)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(seed=st.integers(0, 10_000), f=st.integers(1, 2), extra=st.integers(0, 2), d=st.integers(1, 20), name=st.sampled_from(("brute", "native-brute")))
def test_mda_selects_a_minimum_diameter_subset_in_any_order(seed, f, extra, d, name):
    """MDA under permutation: the selected subset may change (ties) but it always has the minimum diameter, i.e. the
    aggregate is the mean of SOME minimum-diameter subset of n - f inputs."""
    import itertools
    n = 2 * f + 3 + extra
    grads = _inputs(seed, n, d, torch.float64)
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(seed + 1)).tolist()
    shuffled = [grads[i] for i in perm]
    out = aggregators.gars[name](gradients=shuffled, f=f)

    def diameter(subset):
        return max(((shuffled[i] - shuffled[j]) ** 2).sum().item() for i, j in itertools.combinations(subset, 2))
    subsets = list(itertools.combinations(range(n), n - f))
    best = min(diameter(s) for s in subsets)
    candidates = [sum(shuffled[i] for i in s) / len(s) for s in subsets if diameter(s) <= best * (1 + 1e-12)]
    assert any(torch.allclose(out, c, rtol=1e-9, atol=1e-12) for c in candidates)
