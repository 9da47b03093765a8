"""`Aggregator_tf(agg, nb_worker, nb_byz_worker, native)` — numpy-in / numpy-out aggregation façade.

API parity with tensorflow_impl/rsrcs/aggregator_tf/aggregator.py:10-28 (names Average, Median,
Krum, Brute, Aksel, Condense, Bulyan; `native=True` selects the native semantics, which in the
reference loads the `op_krum/op_bulyan/op_median` custom ops) plus the TF1 graph-mode names of
tensorflow_impl/rsrcs/aggregators (`average-nan`, `averaged-median`). Everything is backed by the
one registry of this framework, so numpy gradients on a GPU box go through the sm_100a kernels.
"""

from __future__ import annotations

import numpy as np
import torch

from .. import aggregators as _registry

__all__ = ["Aggregator_tf"]

_NAMES = {
    "Average": ("average", "average"), "Median": ("median", "native-median"), "Krum": ("krum", "native-krum"),
    "Brute": ("brute", "native-brute"), "Aksel": ("aksel", "aksel"), "Condense": ("condense", "condense"),
    "Bulyan": ("bulyan", "native-bulyan"), "TrimmedMean": ("trimmed_mean", "trimmed_mean"),
    "AveragedMedian": ("averaged_median", "averaged_median"),
}


class Aggregator_tf:
    def __init__(self, agg="Average", nb_worker=0, nb_byz_worker=0, native=False, device=None):
        assert agg in _NAMES, "Aggregation not implemented"
        self.name = _NAMES[agg][1 if native else 0]
        self.nb_worker, self.f = nb_worker, nb_byz_worker
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self._rule = _registry.gars[self.name]

    def aggregate(self, gradients):
        """`gradients`: sequence (or 2-D array) of n flat float32 vectors → aggregated float32 vector."""
        rows = [torch.from_numpy(np.array(g, dtype=np.float32, copy=True)).reshape(-1).to(self.device) for g in gradients]   # inputs may be read-only gRPC buffers
        out = self._rule(gradients=rows, f=self.f)
        return out.detach().cpu().numpy()


def average_nan(gradients):
    """Coordinate-wise mean ignoring non-finite entries (TF1 `average-nan`, rsrcs/aggregators/average-nan.py)."""
    stack = np.stack([np.asarray(g, dtype=np.float32).reshape(-1) for g in gradients])
    finite = np.isfinite(stack)
    total = np.where(finite, stack, 0.0).sum(axis=0)
    count = finite.sum(axis=0)
    return np.where(count > 0, total / np.maximum(count, 1), 0.0).astype(np.float32)
