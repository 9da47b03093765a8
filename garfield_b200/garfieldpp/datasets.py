"""Dataset management: `DatasetManager(dataset, minibatch, num_workers, size, rank, train_size)`.

Same API and partitioning rule as the reference (garfieldpp/datasets.py:121-250): datasets
`mnist`, `cifar10`, `pima`; the train set is split in `num_workers` equal shuffled partitions
(seed 1234) and worker `rank - num_ps` gets partition `rank - num_ps` as a list of batches; the
test set is global, batch 100.

Differences: (1) there is no network on the target machines, so when the torchvision files are
not already on disk (or when the name starts with `synthetic-`, or GARFIELD_SYNTHETIC=1) a
deterministic synthetic dataset of the same shape/classes is used; (2) batches are produced by
vectorised tensor ops (no DataLoader worker processes) and kept in pinned host memory when CUDA
is available so that the per-step H2D copy can be asynchronous.
"""

from __future__ import annotations

import logging
import os
import pathlib
import random

import torch

logger = logging.getLogger(__name__)

__all__ = ["DatasetManager", "DataPartitioner", "Partition", "PimaDiabetesDataset", "SyntheticDataset", "datasets_list"]

datasets_list = ["mnist", "cifar10", "pima"]
MNIST, CIFAR10, PIMA = 0, 1, 2

_SHAPES = {"mnist": ((1, 28, 28), 10, 60000, 10000), "cifar10": ((3, 32, 32), 10, 50000, 10000)}
# The Pima table is public data (UCI / Kaggle) that this repository does not redistribute: it is read from the
# first location that has it — next to this module, ~/data, $GARFIELD_PIMA_CSV, or a mounted reference checkout.
_PIMA_CANDIDATES = (
    pathlib.Path(__file__).with_name("pima_diabetes.csv"),
    pathlib.Path.home() / "data" / "pima_diabetes.csv",
    pathlib.Path(os.environ.get("GARFIELD_PIMA_CSV", "/nonexistent")),
    pathlib.Path("/root/reference/pytorch_impl/libs/garfieldpp/pima_diabetes.csv"),
    pathlib.Path(__file__).resolve().parents[2] / "baseline" / "_ref" / "pytorch_impl" / "libs" / "garfieldpp" / "pima_diabetes.csv",
)


class SyntheticDataset(torch.utils.data.Dataset):
    """Deterministic random images with class-dependent means (so that models can learn)."""

    def __init__(self, shape, num_classes, size, seed=0):
        gen = torch.Generator().manual_seed(seed)
        self.targets = torch.randint(0, num_classes, (size,), generator=gen)
        centers = torch.randn(num_classes, *shape, generator=torch.Generator().manual_seed(1234))
        self.data = centers[self.targets] * 0.5 + torch.randn(size, *shape, generator=gen)

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, idx):
        return self.data[idx], int(self.targets[idx])

    def tensors(self):
        return self.data, self.targets


class PimaDiabetesDataset(torch.utils.data.Dataset):
    """Pima Indians diabetes table: 8 z-scored features → binary label (600 train / 168 test).

    Reads `pima_diabetes.csv` (same format as the reference's bundled CSV, garfieldpp/datasets.py:52-94)
    from the first of `_PIMA_CANDIDATES` that exists; otherwise a synthetic stand-in with the same shape
    is generated from a fixed logistic model.
    """
    TRAIN_SPLIT = 600
    TEST_SPLIT = 168

    def __init__(self, train, train_size=None):
        rows = self._load()
        self.source = self._source          # path of the CSV that was read, or None for the synthetic stand-in
        split = self.TRAIN_SPLIT if train_size is None or train_size >= self.TRAIN_SPLIT else train_size
        rows = rows[:split] if train else rows[-self.TEST_SPLIT:]
        feats, labels = rows[:, :-1], rows[:, -1:]
        feats = (feats - feats.mean(dim=0)) / feats.std(dim=0)
        self.data, self.targets = feats.float(), labels.float()

    _source = None
    _warned = False

    @classmethod
    def _load(cls):
        for candidate in _PIMA_CANDIDATES:
            if candidate.is_file():
                import pandas as pd
                cls._source = str(candidate)
                return torch.tensor(pd.read_csv(candidate).values, dtype=torch.float64)
        cls._source = None
        if not cls._warned:      # never silently: the numbers of a run on the stand-in are not Pima numbers
            cls._warned = True
            logger.warning("pima_diabetes.csv not found (looked in %s): using a SYNTHETIC stand-in of the same shape",
                           ", ".join(str(c) for c in _PIMA_CANDIDATES[:3]))
        gen = torch.Generator().manual_seed(768)
        feats = torch.randn(768, 8, generator=gen, dtype=torch.float64) * 3 + 5
        weights = torch.tensor([0.4, 1.1, -0.2, 0.1, 0.3, 0.9, 0.5, 0.6], dtype=torch.float64)
        logits = ((feats - 5) / 3) @ weights + 0.3 * torch.randn(768, generator=gen, dtype=torch.float64) - 0.6
        return torch.cat([feats, (logits > 0).double().unsqueeze(1)], dim=1)

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, idx):
        return self.data[idx], self.targets[idx]

    def tensors(self):
        return self.data, self.targets


class Partition:
    """Dataset-like view over a subset of indices."""

    def __init__(self, data, index):
        self.data, self.index = data, index

    def __len__(self):
        return len(self.index)

    def __getitem__(self, i):
        return self.data[self.index[i]]


class DataPartitioner:
    """Split `data` into consecutive chunks of the given fractions, each shuffled with `seed`
    (one RNG stream across chunks, like the reference, so partitions are reproducible)."""

    def __init__(self, data, sizes=(0.7, 0.2, 0.1), seed=1234):
        self.data = data
        rng = random.Random()
        rng.seed(seed)
        remaining = list(range(len(data)))
        self.partitions = []
        total = len(data)
        for frac in sizes:
            take = int(frac * total)
            chunk = remaining[:take]
            rng.shuffle(chunk)
            self.partitions.append(chunk)
            remaining = remaining[take:]

    def use(self, partition):
        return Partition(self.data, self.partitions[partition])


def _want_synthetic(name):
    return name.startswith("synthetic-") or bool(os.environ.get("GARFIELD_SYNTHETIC"))


class DatasetManager:
    def __init__(self, dataset, minibatch, num_workers, size, rank, train_size=None):
        base = dataset[len("synthetic-"):] if dataset.startswith("synthetic-") else dataset
        if base not in datasets_list:
            raise Exception("Existing datasets are: ", datasets_list)
        self.name = base
        self.synthetic = _want_synthetic(dataset)
        self.dataset = datasets_list.index(base)
        self.batch = minibatch * num_workers
        self.num_workers = num_workers
        self.num_ps = size - num_workers
        self.rank = rank
        self.train_size = train_size

    # -- sources -----------------------------------------------------------------------------
    def _torchvision(self, train):
        import torchvision
        from torchvision import transforms
        root = str(pathlib.Path.home() / "data")
        if self.dataset == MNIST:
            tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
            return torchvision.datasets.MNIST(root, train=train, download=False, transform=tf)
        norm = transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
        steps = [transforms.RandomCrop(32, padding=4), transforms.RandomHorizontalFlip()] if train else []
        tf = transforms.Compose(steps + [transforms.ToTensor(), norm])
        return torchvision.datasets.CIFAR10(root, train=train, download=False, transform=tf)

    def fetch_dataset(self, train=True):
        """The train or test `Dataset` (real files when present on disk, synthetic otherwise)."""
        if self.dataset == PIMA:
            return PimaDiabetesDataset(train=train, train_size=self.train_size)
        if not self.synthetic:
            try:
                return self._torchvision(train)
            except Exception as err:
                logger.warning("dataset %s unavailable on disk (%s): using synthetic data", self.name, err)
                self.synthetic = True
        shape, classes, n_train, n_test = _SHAPES[self.name]
        size = n_train if train else n_test
        if train and self.train_size is not None:
            size = min(size, int(self.train_size))
        cap = os.environ.get("GARFIELD_SYNTHETIC_SIZE")
        if cap:
            size = min(size, int(cap) if train else max(100, int(cap) // 5))
        return SyntheticDataset(shape, classes, size, seed=0 if train else 1)

    # -- batching ----------------------------------------------------------------------------
    @staticmethod
    def _batches(dataset, indices, bsz, pin):
        """Materialise `indices` of `dataset` as a list of (data, target) batches."""
        if hasattr(dataset, "tensors"):
            data, targets = dataset.tensors()
            idx = torch.as_tensor(indices, dtype=torch.long)
            data, targets = data[idx], targets[idx]
        else:
            samples = [dataset[i] for i in indices]
            data = torch.stack([s[0] for s in samples])
            targets = torch.as_tensor([s[1] for s in samples])
        out = []
        for start in range(0, data.shape[0], bsz):
            x, y = data[start:start + bsz].contiguous(), targets[start:start + bsz].contiguous()
            if pin:
                x, y = x.pin_memory(), y.pin_memory()
            out.append((x, y))
        return out

    def get_train_set(self):
        """This worker's partition as a list of (data, target) batches of `minibatch` samples."""
        train = self.fetch_dataset(train=True)
        parts = self.num_workers
        bsz = int(self.batch / float(parts))
        partition = DataPartitioner(train, [1.0 / parts] * parts).use(self.rank - self.num_ps)
        return self._batches(train, partition.index, bsz, torch.cuda.is_available())

    def get_test_set(self):
        """The global test set as a list of batches of 100 samples."""
        test = self.fetch_dataset(train=False)
        return self._batches(test, list(range(len(test))), 100, torch.cuda.is_available())
