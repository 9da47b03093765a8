"""GAR semantics: every registry name against (1) the reference's pure-torch rules imported
read-only and (2) independent restatements of SURVEY.md §2.8, on CPU, fp32 and fp64."""

import itertools
import math

import pytest
import torch

import aggregators
from garfield_b200.aggregators import _host


def make(n, d, dtype=torch.float32, seed=0, spread=1.0):
    gen = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=gen, dtype=torch.float64)
    return [(base + spread * torch.randn(d, generator=gen, dtype=torch.float64)).to(dtype) for _ in range(n)]


CASES = [(5, 1), (7, 1), (8, 1), (11, 2), (15, 3), (23, 5)]


def test_registry_surface():
    expected = {"average", "median", "krum", "bulyan", "brute", "aksel", "condense", "native-median",
                "native-krum", "native-bulyan", "native-brute", "trimmed_mean", "averaged_median", "mda"}
    assert expected <= set(aggregators.gars)
    for name, rule in aggregators.gars.items():
        for attr in ("check", "checked", "unchecked", "upper_bound", "influence"):
            assert hasattr(rule, attr), (name, attr)
        assert getattr(aggregators, name.replace("-", "_")) is rule
    with pytest.raises(TypeError):
        aggregators.gars["average"]([torch.zeros(3)])  # keyword-only, like the reference


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("n,f", CASES)
@pytest.mark.parametrize("name", ["average", "median", "krum", "brute", "aksel", "bulyan"])
def test_matches_reference_python_rules(ref, name, n, f, dtype):
    grads = make(n, 257, dtype, seed=n * 31 + f)
    ours, theirs = aggregators.gars[name], ref.gars[name]
    if theirs.check(gradients=grads, f=f) is not None:
        assert ours.check(gradients=grads, f=f) is not None
        with pytest.raises(Exception):
            ours.checked(gradients=grads, f=f)
        return
    if name == "brute" and math.comb(n, f) > 5000:
        pytest.skip("python brute oracle too slow")
    want = theirs.unchecked(gradients=[g.clone() for g in grads], f=f)
    got = ours(gradients=grads, f=f)
    tol = 2e-5 if dtype == torch.float32 else 1e-12
    assert got.dtype == dtype and got.shape == want.shape
    assert got.data_ptr() not in {g.data_ptr() for g in grads}
    torch.testing.assert_close(got, want, rtol=tol, atol=tol)


@pytest.mark.parametrize("n,f,m", [(7, 1, 1), (7, 1, 3), (11, 2, 5), (11, 2, 2)])
def test_krum_m_parameter(ref, n, f, m):
    grads = make(n, 64, seed=5)
    torch.testing.assert_close(aggregators.krum(gradients=grads, f=f, m=m),
                               ref.gars["krum"].unchecked(gradients=grads, f=f, m=m), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("mode", ["mid", "n-f"])
def test_aksel_modes(ref, mode):
    grads = make(9, 100, seed=3)
    torch.testing.assert_close(aggregators.aksel(gradients=grads, f=2, mode=mode),
                               ref.gars["aksel"].unchecked(gradients=grads, f=2, mode=mode), rtol=1e-5, atol=1e-6)


def test_check_boundaries_match_reference(ref):
    for name in ("krum", "bulyan", "brute", "aksel", "condense", "median", "average"):
        for n in range(1, 14):
            grads = make(n, 8)
            for f in (0, 1, 2, 3, "x", 1.0):
                ours = aggregators.gars[name].check(gradients=grads, f=f)
                theirs = ref.gars[name].check(gradients=grads, f=f)
                assert (ours is None) == (theirs is None), (name, n, f, ours, theirs)
    assert aggregators.krum.check(gradients=[], f=1) is not None
    assert aggregators.krum.check(gradients=make(7, 4), f=1, m=5) is not None
    assert aggregators.krum.check(gradients=make(7, 4), f=1, m=4) is None


def test_upper_bounds_match_reference(ref):
    for name in ("krum", "bulyan", "brute", "median", "condense"):
        for n, f in ((11, 2), (23, 5), (9, 1)):
            assert aggregators.gars[name].upper_bound(n, f, 100) == pytest.approx(ref.gars[name].upper_bound(n, f, 100))


def test_influence(ref):
    honests = make(8, 50, seed=1, spread=0.1)
    attacks = [h * -100 for h in make(2, 50, seed=2)]
    for name in ("average", "krum", "brute", "aksel"):
        ours = aggregators.gars[name].influence(honests, attacks, f=2)
        theirs = ref.gars[name].influence(honests, attacks, f=2)
        assert ours == pytest.approx(theirs), name
    near = [honests[0] + 1e-4, honests[1] - 1e-4]
    assert aggregators.krum.influence(honests, near, f=2) > 0


# ---- independent restatements (no reference needed) -------------------------------------

def brute_force_krum(grads, f, m, neighbors, squared):
    n = len(grads)
    dist = [[math.inf] * n for _ in range(n)]
    for i, j in itertools.combinations(range(n), 2):
        v = (grads[i].double() - grads[j].double()).pow(2).sum().item()
        dist[i][j] = dist[j][i] = v if squared else math.sqrt(v)
    scores = [sum(sorted(dist[i][j] for j in range(n) if j != i)[:neighbors]) for i in range(n)]
    best = sorted(range(n), key=lambda i: (scores[i], i))[:m]
    return torch.stack([grads[i].double() for i in best]).mean(0)


@pytest.mark.parametrize("n,f", [(7, 1), (11, 2), (16, 3)])
def test_native_krum_semantics(n, f):
    grads = make(n, 300, torch.float64, seed=n)
    want = brute_force_krum(grads, f, n - f - 2, n - f - 2, squared=True)
    torch.testing.assert_close(aggregators.gars["native-krum"](gradients=grads, f=f), want, rtol=1e-10, atol=1e-10)
    want_py = brute_force_krum(grads, f, n - f - 2, n - f - 1, squared=False)
    torch.testing.assert_close(aggregators.krum(gradients=grads, f=f), want_py, rtol=1e-10, atol=1e-10)


def test_median_variants():
    grads = make(6, 40, torch.float64, seed=9)
    mat = torch.stack(grads)
    srt = mat.sort(0).values
    torch.testing.assert_close(aggregators.median(gradients=grads), srt[2])           # lower
    torch.testing.assert_close(aggregators.gars["native-median"](gradients=grads), srt[3])  # upper
    torch.testing.assert_close(aggregators.trimmed_mean(gradients=grads, f=1), srt[1:5].mean(0))
    torch.testing.assert_close(aggregators.trimmed_mean(gradients=grads, f=0), mat.mean(0))
    # non-finite handling of native-median
    bad = [g.clone() for g in grads]
    bad[0][0] = float("nan"); bad[1][0] = float("inf"); bad[2][0] = -float("inf")
    for g in bad:
        g[1] = float("nan")
    out = aggregators.gars["native-median"](gradients=bad)
    finite0 = sorted(g[0].item() for g in bad[3:])
    assert out[0].item() == finite0[len(finite0) // 2]
    assert out[1].item() == 0.0
    assert torch.isnan(aggregators.median(gradients=bad)[1])


def test_averaged_median_definition():
    grads = make(9, 64, torch.float64, seed=4)
    mat = torch.stack(grads)
    med = mat.median(0).values
    beta = 6
    order = (mat - med).abs().argsort(0)[:beta]
    want = mat.gather(0, order).mean(0)
    torch.testing.assert_close(aggregators.averaged_median(gradients=grads, f=3), want)
    torch.testing.assert_close(_host.averaged_median(mat, beta), want)


def test_condense_is_mix_of_median_and_first():
    grads = make(8, 5000, seed=6)
    out = aggregators.condense(gradients=grads, f=1, p=0.9)
    med = torch.stack(grads).median(0).values
    from_med = out == med
    from_first = out == grads[0]
    assert bool((from_med | from_first).all())
    assert 0.85 < from_med.float().mean().item() <= 1.0
    assert torch.equal(aggregators.condense(gradients=grads, f=1, p=1.0), med)


def test_byzantine_inputs_do_not_leak():
    honests = make(9, 128, seed=11, spread=0.05)
    nan_attack = [torch.full((128,), float("nan")), torch.full((128,), float("inf"))]
    grads = honests + nan_attack
    for name in ("krum", "native-krum", "bulyan", "native-bulyan", "brute", "native-median"):
        out = aggregators.gars[name](gradients=grads, f=2)
        assert torch.isfinite(out).all(), name
    big = [h * -100 for h in honests[:2]]
    mean = torch.stack(honests).mean(0)
    for name in ("krum", "bulyan", "brute", "median", "trimmed_mean", "aksel"):
        out = aggregators.gars[name](gradients=honests + big, f=2)
        assert (out - mean).norm() < 1.0, name


def test_native_bulyan_against_restated_algorithm():
    n, f = 11, 2
    grads = make(n, 200, torch.float64, seed=21)
    dist = _host.pairwise_distances(grads, squared=True)
    rounds = _host.bulyan_rounds_native(dist, f, None)
    theta = n - 2 * f - 2
    assert len(rounds) == theta and [len(r) for r in rounds] == [n - f - 2 - k for k in range(theta)]
    rows = torch.stack([torch.stack([grads[i] for i in r]).mean(0) for r in rounds])
    srt = rows.sort(0).values
    med = srt[theta // 2]
    beta = theta - 2 * f
    order = (rows - med).abs().argsort(0)[:beta]
    want = rows.gather(0, order).mean(0)
    torch.testing.assert_close(aggregators.gars["native-bulyan"](gradients=grads, f=f), want, rtol=1e-9, atol=1e-9)
