"""Host-side (pure torch / Python) definitions of every aggregation rule's semantics.

These are the *oracle* definitions of SURVEY.md §2.8, used (a) for CPU tensors, (b) as the
fallback when the sm_100a library is not loadable, and (c) by the tests as the reference the
CUDA kernels are compared against. The O(n²)/O(C(n,f)) selection logic runs on tiny host
matrices; everything that touches the d-length vectors is a vectorised torch op.

Two families of semantics coexist in the reference under one name family:

* "python" rules (`krum`, `bulyan`, `median`, `brute`): non-squared L2 distances, Python's
  stable sort, `torch.median` (lower median, NaN-propagating)
  — /root/reference/pytorch_impl/libs/aggregators/{krum,bulyan,median,brute}.py
* "native" rules (`native-*`): squared distances, n−f−2 neighbours, non-finite treated as
  +inf, upper median of the finite values, real Bulyan score updates
  — /root/reference/pytorch_impl/libs/native/py_{krum,bulyan,median,brute}/*.cpp
"""

from __future__ import annotations

import itertools
import math

import torch

__all__ = [
    "as_matrix", "pairwise_distances", "krum_scores", "krum_selection", "bulyan_rounds_python",
    "bulyan_rounds_native", "mda_selection", "weights_from_rounds", "apply_weights",
    "median_lower", "median_finite_upper", "trimmed_mean", "averaged_median", "bulyan_phase2",
]


def as_matrix(gradients):
    """Stack a list of n 1-D tensors into an [n, d] matrix (a view when given a matrix)."""
    if isinstance(gradients, torch.Tensor):
        return gradients
    return torch.stack(tuple(gradients))


def pairwise_distances(gradients, squared):
    """n×n list-of-lists of Python floats: (squared) L2 distances, +inf on the diagonal and for
    any non-finite value. Computed by direct differences (no Gram cancellation), accumulating
    in the input dtype exactly like `sub().norm()` / Σ(x−y)² in the reference."""
    n = len(gradients)
    dist = [[math.inf] * n for _ in range(n)]
    for i in range(n - 1):
        for j in range(i + 1, n):
            diff = gradients[i].sub(gradients[j])
            value = diff.dot(diff).item() if squared else diff.norm().item()
            if not math.isfinite(value):
                value = math.inf
            dist[i][j] = dist[j][i] = value
    return dist


def krum_scores(dist, k):
    """score_i = sum of the k smallest distances from i to the others."""
    n = len(dist)
    scores = []
    for i in range(n):
        row = sorted(dist[i][j] for j in range(n) if j != i)
        scores.append(sum(row[:k]))
    return scores


def _stable_order(scores):
    """Indices by increasing score, ties by increasing index, non-finite last."""
    return sorted(range(len(scores)), key=lambda i: (not math.isfinite(scores[i]), scores[i] if math.isfinite(scores[i]) else 0.0, i))


def krum_selection(dist, f, m, k):
    """Indices of the m best-scored gradients (Multi-Krum), best first."""
    return _stable_order(krum_scores(dist, k))[:m]


def bulyan_rounds_python(dist, f, m):
    """Selection rounds of the reference *python* Bulyan (static scores).

    Returns a list of θ = n−2f−2 index lists; round i averages `min(m, m_max−i)` survivors
    with the best (never updated) scores, then drops the best one.
    """
    n = len(dist)
    m_max = n - f - 2
    if m is None:
        m = m_max
    scores = krum_scores(dist, m)  # python bulyan sums the m nearest (self is inf, sorted last)
    alive = _stable_order(scores)
    rounds = []
    for i in range(n - 2 * f - 2):
        m = min(m, m_max - i)
        rounds.append(list(alive[:m]))
        alive = alive[1:]
    return rounds


def bulyan_rounds_native(dist2, f, m):
    """Selection rounds of the reference *native* Bulyan (py_bulyan/bulyan.cpp:102-165):
    scores over the n−f−2 nearest squared distances; each row keeps only those neighbours'
    distances ("pruning"), and removing gradient r lowers score_i by the kept dist2[i][r]."""
    n = len(dist2)
    keep = n - f - 2
    if m is None:
        m = keep
    kept = [[0.0] * n for _ in range(n)]
    scores = []
    for i in range(n):
        order = sorted((j for j in range(n) if j != i), key=lambda j: (dist2[i][j], j))
        near = order[:keep]
        scores.append(math.fsum(dist2[i][j] for j in near))
        for j in near:
            kept[i][j] = dist2[i][j]
    rounds = []
    theta = n - 2 * f - 2
    for k in range(theta):
        order = _stable_order(scores)
        rounds.append(order[:m - k])
        if k + 1 >= theta:
            break
        gone = order[0]
        scores[gone] = math.inf
        for i in range(n):
            if i != gone and math.isfinite(scores[i]):
                scores[i] -= kept[i][gone]
    return rounds


def mda_selection(dist, f):
    """Minimum-diameter subset of size n−f (first minimal in lexicographic order); subsets
    containing a non-finite distance are skipped. None when every subset is skipped."""
    n = len(dist)
    best, best_diam = None, None
    for subset in itertools.combinations(range(n), n - f):
        diam = 0.0
        for x, y in itertools.combinations(subset, 2):
            value = dist[x][y]
            if not math.isfinite(value):
                break
            if value > diam:
                diam = value
        else:
            if best is None or diam < best_diam:
                best, best_diam = subset, diam
    return best


def weights_from_rounds(rounds, n, dtype=torch.float32, device="cpu"):
    """θ×n averaging-weight matrix: row k holds 1/|S_k| on the indices selected in round k."""
    weights = torch.zeros(len(rounds), n, dtype=dtype)
    for k, chosen in enumerate(rounds):
        for i in chosen:
            weights[k, i] = 1.0 / len(chosen)
    return weights.to(device)


def apply_weights(gradients, weights):
    """rows_k = Σ_i W[k,i]·g_i, skipping zero weights so excluded (possibly NaN) inputs never
    contaminate the output. Returns a [θ, d] matrix."""
    first = gradients[0]
    out = torch.zeros(weights.shape[0], first.shape[0], dtype=first.dtype, device=first.device)
    host = weights.tolist()
    for k, row in enumerate(host):
        for i, w in enumerate(row):
            if w != 0.0:
                out[k].add_(gradients[i], alpha=w)
    return out


def median_lower(matrix):
    """`torch.median` semantics: lower median for even n, NaN-propagating."""
    return matrix.median(dim=0).values


def median_finite_upper(matrix):
    """Per coordinate: element ⌊len/2⌋ of the sorted *finite* values; 0 when none is finite
    (native median, py_median/median.cpp:55-72)."""
    finite = torch.isfinite(matrix)
    big = torch.finfo(matrix.dtype).max
    keyed = torch.where(finite, matrix, torch.full_like(matrix, big))
    ordered, _ = keyed.sort(dim=0)
    count = finite.sum(dim=0)
    pick = (count // 2).clamp_(max=matrix.shape[0] - 1)
    out = ordered.gather(0, pick.unsqueeze(0)).squeeze(0)
    return torch.where(count > 0, out, torch.zeros_like(out))


def trimmed_mean(matrix, f):
    """Per coordinate: drop the f smallest and f largest, average the remaining n−2f."""
    n = matrix.shape[0]
    ordered, _ = matrix.sort(dim=0)
    return ordered[f:n - f].mean(dim=0)


def averaged_median(matrix, beta, upper=False):
    """Per coordinate: mean of the β values closest to the median (ties: lower value first,
    which is what a sort by (|x−med|, x) gives)."""
    n = matrix.shape[0]
    ordered, _ = matrix.sort(dim=0)
    med = ordered[n // 2 if upper else (n - 1) // 2]
    # the β closest to the median form a contiguous window of the sorted column
    best = None
    best_cost = None
    for start in range(0, n - beta + 1):
        window = ordered[start:start + beta]
        cost = torch.maximum((window[0] - med).abs(), (window[-1] - med).abs())
        if best is None:
            best, best_cost = window.mean(dim=0), cost
        else:
            better = cost < best_cost
            best = torch.where(better, window.mean(dim=0), best)
            best_cost = torch.where(better, cost, best_cost)
    return best


def bulyan_phase2(selected, f, upper=False):
    """Bulyan's coordinate-wise step on the θ×d matrix of round averages: mean of the
    β = θ−2f values closest to the (lower | upper) median."""
    return averaged_median(selected, selected.shape[0] - 2 * f, upper=upper)
