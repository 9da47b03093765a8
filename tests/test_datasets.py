"""garfieldpp.datasets: seeded partitioner parity with the reference's algorithm, per-worker batch
lists, the global test set, the Pima table (real CSV when reachable, synthetic stand-in otherwise) and the synthetic
fall-back used on machines without the torchvision files."""
import os
import random

import pytest
import torch

os.environ.setdefault("GARFIELD_SYNTHETIC", "1")
os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1200")

from garfieldpp import datasets  # noqa: E402


def _reference_partitions(length, sizes, seed=1234):
    """The reference's algorithm restated (garfieldpp/datasets.py:121-144): consecutive chunks, each shuffled by ONE seeded stream."""
    rng = random.Random()
    rng.seed(seed)
    indexes, out = list(range(length)), []
    for frac in sizes:
        take = int(frac * length)
        chunk = indexes[:take]
        rng.shuffle(chunk)
        out.append(chunk)
        indexes = indexes[take:]
    return out


@pytest.mark.parametrize("length,parts", [(1000, 4), (997, 7), (60, 8)])
def test_partitioner_matches_the_reference_algorithm(length, parts):
    data = list(range(length))
    mine = datasets.DataPartitioner(data, [1.0 / parts] * parts)
    want = _reference_partitions(length, [1.0 / parts] * parts)
    assert mine.partitions == want
    seen = [i for p in mine.partitions for i in p]
    assert len(seen) == len(set(seen)) and all(len(p) == int(length / parts) for p in mine.partitions)   # disjoint, equal shares
    view = mine.use(1)
    assert len(view) == len(want[1]) and view[0] == data[want[1][0]]


def test_train_set_is_a_list_of_batches_per_worker_and_test_set_is_global():
    n, ps, batch = 5, 1, 16
    sets = [datasets.DatasetManager("cifar10", batch, n, n + ps, ps + k).get_train_set() for k in range(n)]
    for batches in sets:
        assert isinstance(batches, list) and len(batches) == len(sets[0])
        x, y = batches[0]
        assert x.shape == (batch, 3, 32, 32) and y.shape == (batch,) and y.dtype == torch.long
        assert sum(b[0].shape[0] for b in batches) == 1200 // n                        # this worker's share of the train set
    assert not torch.equal(sets[0][0][0], sets[1][0][0])                                 # different partitions
    again = datasets.DatasetManager("cifar10", batch, n, n + ps, ps).get_train_set()
    assert torch.equal(again[0][0], sets[0][0][0])                                       # and reproducible
    test = datasets.DatasetManager("cifar10", batch, n, n + ps, 0).get_test_set()
    assert all(b[0].shape[0] <= 100 for b in test) and test[0][0].shape[0] == 100
    mnist = datasets.DatasetManager("mnist", 8, 2, 3, 1).get_train_set()
    assert mnist[0][0].shape == (8, 1, 28, 28)
    with pytest.raises(Exception):
        datasets.DatasetManager("svhn", 8, 2, 3, 1)


def test_synthetic_data_is_learnable_and_deterministic():
    a = datasets.SyntheticDataset((3, 8, 8), 4, 256, seed=0)
    b = datasets.SyntheticDataset((3, 8, 8), 4, 256, seed=0)
    assert torch.equal(a.data, b.data) and torch.equal(a.targets, b.targets)
    centers = torch.stack([a.data[a.targets == c].mean(0) for c in range(4)])
    nearest = ((a.data.unsqueeze(1) - centers.unsqueeze(0)) ** 2).flatten(2).sum(-1).argmin(1)
    assert (nearest == a.targets).float().mean().item() > 0.9                            # class-dependent means: separable


def test_pima_table_shapes_and_standardisation():
    train, test = datasets.PimaDiabetesDataset(train=True), datasets.PimaDiabetesDataset(train=False)
    assert len(train) == 600 and len(test) == 168
    x, y = train[0]
    assert x.shape == (8,) and y.shape == (1,) and y.item() in (0.0, 1.0)
    feats, labels = train.tensors()
    assert torch.allclose(feats.mean(0), torch.zeros(8), atol=1e-5) and torch.allclose(feats.std(0), torch.ones(8), atol=1e-4)
    assert 0.2 < labels.mean().item() < 0.8
    assert len(datasets.PimaDiabetesDataset(train=True, train_size=100)) == 100
    batches = datasets.DatasetManager("pima", 16, 3, 4, 1).get_train_set()
    assert batches[0][0].shape == (16, 8) and batches[0][1].shape == (16, 1)
