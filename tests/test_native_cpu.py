"""C++ host library: native-* rules vs the pure-torch definitions, combinations vs itertools,
thread-pool coverage, multi-replica register."""

import ctypes
import itertools
import math
import threading

import pytest
import torch

import native
from garfield_b200.aggregators import _host
from garfield_b200.utils.register import MultiRegister, RegisterClosed, RegisterTimeout


def make(n, d, dtype, seed):
    gen = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=gen, dtype=torch.float64)
    return [(base + torch.randn(d, generator=gen, dtype=torch.float64)).to(dtype) for _ in range(n)]


def test_library_loads():
    assert native.available()
    assert native.num_threads() >= 1


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 3e-5), (torch.float64, 1e-11)])
@pytest.mark.parametrize("n,f,d", [(7, 1, 10), (11, 2, 5000), (15, 3, 4097), (23, 5, 333)])
def test_native_rules_match_host_semantics(n, f, d, dtype, tol):
    g = make(n, d, dtype, seed=n + d)
    dist = _host.pairwise_distances(g, squared=True)
    sel = _host.krum_selection(dist, f, n - f - 2, n - f - 2)
    torch.testing.assert_close(native.krum.aggregate(g, f, n - f - 2), torch.stack([g[i] for i in sel]).mean(0), rtol=tol, atol=tol)
    torch.testing.assert_close(native.krum.aggregate(g, f, 1), g[sel[0]], rtol=tol, atol=tol)
    torch.testing.assert_close(native.median.aggregate(g), _host.median_finite_upper(torch.stack(g)), rtol=0, atol=0)
    rounds = _host.bulyan_rounds_native(dist, f, None)
    rows = _host.apply_weights(g, _host.weights_from_rounds(rounds, n, dtype=dtype))
    torch.testing.assert_close(native.bulyan.aggregate(g, f, n - f - 2), _host.bulyan_phase2(rows, f, upper=True), rtol=tol, atol=tol)
    if math.comb(n, f) <= 5000:
        sub = _host.mda_selection(dist, f)
        torch.testing.assert_close(native.brute.aggregate(g, f), torch.stack([g[i] for i in sub]).mean(0), rtol=tol, atol=tol)


def test_pairwise_sqdist_and_nonfinite():
    g = make(6, 1000, torch.float32, 2)
    got = native.pairwise_sqdist(g)
    for i, j in itertools.combinations(range(6), 2):
        want = (g[i].double() - g[j].double()).pow(2).sum().item()
        assert got[i, j].item() == pytest.approx(want, rel=1e-5)
        assert got[i, j] == got[j, i]
    assert math.isinf(got[0, 0].item())
    g[2][5] = float("nan")
    out = native.krum.aggregate(g + make(3, 1000, torch.float32, 3), 1, 3)
    assert torch.isfinite(out).all()


def test_argument_errors():
    g = make(4, 8, torch.float32, 0)
    with pytest.raises(ValueError):
        native.krum.aggregate(g, 1, 1)  # n < 2f+3
    with pytest.raises(ValueError):
        native.bulyan.aggregate(make(6, 8, torch.float32, 0), 1, 1)
    with pytest.raises(ValueError):
        native.brute.aggregate(g, 2)
    with pytest.raises(ValueError):
        native.median.aggregate([g[0], g[1][:4]])
    with pytest.raises(TypeError):
        native.median.aggregate([t.half() for t in g])


@pytest.mark.parametrize("n,k", [(1, 1), (5, 0), (5, 2), (6, 3), (8, 7), (10, 5)])
def test_combinations_match_itertools(n, k):
    want = list(itertools.combinations(range(n), k))
    assert native.combinations.count(n, k) == len(want) == math.comb(n, k)
    if k == 0:
        return
    assert list(native.combinations.iterate(n, k)) == want
    for pos in range(0, len(want), max(1, len(want) // 7)):
        assert native.combinations.seek(n, k, pos) == want[pos]
        assert list(native.combinations.iterate(n, k, pos, pos + 3)) == want[pos:pos + 3]


def test_combinations_overflow():
    assert native.combinations.count(60, 30) == math.comb(60, 30)
    with pytest.raises(OverflowError):
        native.combinations.count(200, 100)


@pytest.mark.parametrize("total,grain", [(1, 0), (17, 0), (1000, 7), (100000, 0), (5, 100)])
def test_parallel_for_covers_range_exactly_once(total, grain):
    hits = torch.zeros(total, dtype=torch.int32)
    chunks = native.lib().gb_cpu_parallel_for_probe(total, grain, hits.data_ptr())
    assert chunks >= 1
    assert bool((hits == 1).all())


@pytest.mark.parametrize("shape,axis", [((7,), 0), ((5, 9), 0), ((5, 9), 1), ((4, 6, 3), 1), ((3, 4, 5, 6), 2), ((3, 4, 5, 6), 0)])
def test_strided_view_axis_nth_matches_sort(shape, axis):
    """std::nth_element over the strided iterator (csrc/cpu/strided_view.hpp) along any axis, in place."""
    import ctypes
    torch.manual_seed(sum(shape) + axis)
    data = torch.randn(*shape)
    k = shape[axis] // 2
    want = data.sort(dim=axis).values.select(axis, k).reshape(-1)
    work = data.clone()
    out = torch.empty(want.numel())
    dims = (ctypes.c_int64 * len(shape))(*shape)
    assert native.lib().gb_cpu_axis_nth_f32(work.data_ptr(), len(shape), dims, axis, k, out.data_ptr()) == 0
    assert torch.equal(out, want)
    assert torch.equal(work.sort(dim=axis).values, data.sort(dim=axis).values)   # a permutation along the axis only
    assert torch.equal(work.select(axis, k).reshape(-1), want)                   # k-th element in place
    assert native.lib().gb_cpu_axis_nth_f32(work.data_ptr(), len(shape), dims, len(shape), k, out.data_ptr()) != 0


def test_multiregister_latest_value_and_close():
    reg = MultiRegister(3, 16)
    out = torch.zeros(4)
    with pytest.raises(RegisterTimeout):
        reg.read(out, timeout=0.01)
    for v in range(5):
        reg.write(torch.full((4,), float(v)))
    seq = reg.read(out)
    assert seq == 5 and out.tolist() == [4.0] * 4
    assert reg.read(out) == 5  # non-consuming: the newest value stays readable
    reg.close()
    with pytest.raises(RegisterClosed):
        reg.write(torch.zeros(4))


def test_multiregister_consume_mode_is_a_queue():
    reg = MultiRegister(2, 8, consume=True)
    got = []

    def consumer():
        buf = torch.zeros(2)
        try:
            while True:
                reg.read(buf, timeout=2.0)
                got.append(buf[0].item())
        except (RegisterClosed, RegisterTimeout):
            pass

    th = threading.Thread(target=consumer)
    th.start()
    for v in range(20):
        reg.write(torch.full((2,), float(v)), timeout=2.0)
    import time
    deadline = time.time() + 2
    while len(got) < 20 and time.time() < deadline:
        time.sleep(0.01)
    reg.close()
    th.join()
    assert got == [float(v) for v in range(20)]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("n,d", [(1, 5), (2, 70), (7, 129), (8, 1000), (23, 333), (32, 64), (33, 50)])
def test_host_coordwise_kernel_matches_the_reference_formulations(n, d, dtype):
    """gb_cpu_coordwise (vectorised network) against aggregators/_host.py: lower median with NaN propagation and ±inf,
    trimmed mean, mean-around-median; inputs it declines (NaN / non-finite columns, n > 32) must return None."""
    from garfield_b200.aggregators import _host
    gen = torch.Generator().manual_seed(n * 1000 + d)
    rows = [torch.randn(d, generator=gen, dtype=dtype) for _ in range(n)]
    matrix = torch.stack(rows)
    got = native.coordwise(rows, "median_lower")
    if n > 32:
        assert got is None
        return
    assert torch.equal(got, _host.median_lower(matrix))
    f = (n - 1) // 2 if n < 5 else 2
    assert torch.allclose(native.coordwise(rows, "trimmed_mean", f), _host.trimmed_mean(matrix, f), rtol=1e-6, atol=1e-7)
    for beta in {1, max(1, n - f), n}:
        assert torch.allclose(native.coordwise(rows, "averaged_median", beta), _host.averaged_median(matrix, beta), rtol=1e-6, atol=1e-7), beta
    if n >= 3 and d >= 5:
        rows[0][1] = float("nan"); rows[1][2] = float("inf"); rows[2][3] = float("-inf"); rows[n - 1][4] = float("nan")
        matrix = torch.stack(rows)
        got, want = native.coordwise(rows, "median_lower"), _host.median_lower(matrix)
        assert torch.equal(torch.isnan(got), torch.isnan(want)) and torch.equal(got[~torch.isnan(got)], want[~torch.isnan(want)])
        assert native.coordwise(rows, "trimmed_mean", 1) is None and native.coordwise(rows, "averaged_median", 2) is None
    with pytest.raises(ValueError):
        native.coordwise(rows, "trimmed_mean", n)


def test_registry_cpu_rules_use_the_host_kernel_and_fall_back():
    import aggregators
    rows = [torch.randn(257) for _ in range(9)]
    matrix = torch.stack(rows)
    assert torch.equal(aggregators.gars["median"](gradients=rows, f=1), matrix.median(0).values)
    rows[3][5] = float("nan")
    out = aggregators.gars["trimmed_mean"](gradients=rows, f=1)           # NaN column: reference formulation takes over
    want = torch.stack(rows).sort(0).values[1:8].mean(0)
    assert torch.allclose(out, want, equal_nan=True)
