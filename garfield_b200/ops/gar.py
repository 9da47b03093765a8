"""Device aggregation rules: thin Python drivers around the sm_100a kernels.

Every rule is at most two launches and never synchronises with the host:

    pairwise_select  one TMA-streamed pass → n×n distances → in-kernel selection → W (θ×n)
    coordwise        one TMA-streamed pass → per-coordinate rule on x or on W·x
                     (+ optional fused SGD update and publication to peer arenas)

Inputs are lists of n 1-D float32 CUDA tensors (any mix of local and peer-mapped memory).
Reference counterparts: /root/reference/pytorch_impl/libs/native/py_*/ *.cu (see the kernel
files for line-level citations).
"""

from __future__ import annotations

import ctypes

import torch

from . import _lib
from ._lib import FusedUpdate

__all__ = ["available", "pairwise_sqdist", "select", "coordwise", "krum", "bulyan", "mda", "aksel", "distance_rules",
           "average", "weighted", "SgdFusion", "MAX_ROWS"]

MAX_ROWS = 64

RULES = {"linear": 0, "median_lower": 1, "median_finite_upper": 2, "trimmed_mean": 3,
         "averaged_median": 4, "condense": 5}
MODES = {"none": 0, "krum": 1, "bulyan_python": 2, "bulyan_native": 3, "mda": 4, "closest": 5}


def available():
    return _lib.available()


class _Workspace:
    """Per-device cached scratch (partials + counter), distance / weight / info buffers."""
    _cache = {}

    @classmethod
    def get(cls, device, n, tag=0, dtype=torch.float32):
        """`tag` separates workspaces that are in flight at the same time (e.g. distances of one coordinate range
        computed on a side stream while another call runs)."""
        key = (device.index if device.index is not None else torch.cuda.current_device(), n, tag, dtype)
        ws = cls._cache.get(key)
        if ws is None:
            lib = _lib.require()
            with torch.cuda.device(device):
                nbytes = int(lib.gb_pairwise_scratch_bytes(n) if dtype == torch.float32 else lib.gb_pairwise_scratch_bytes_f64(n))
            ws = dict(
                scratch=torch.zeros(nbytes, dtype=torch.uint8, device=device),
                dist=torch.empty(n, n, dtype=dtype, device=device),
                W=torch.zeros(n, n, dtype=torch.float32, device=device),
                info=torch.zeros(1 + n, dtype=torch.int32, device=device),
            )
            cls._cache[key] = ws
        return ws


def _prep(gradients):
    n = len(gradients)
    if n < 1 or n > MAX_ROWS:
        raise ValueError("device rules support 1..%d gradients, got %d" % (MAX_ROWS, n))
    g0 = gradients[0]
    rows = []
    for g in gradients:
        if g.dtype not in (torch.float32, torch.float64) or g.dtype != g0.dtype or not g.is_cuda or g.dim() != 1 or g.shape != g0.shape:
            raise ValueError("expected 1-D float32 (or float64) CUDA tensors of equal length and dtype")
        rows.append(g if g.is_contiguous() else g.contiguous())
    return rows, n, g0.numel(), g0.device


class SgdFusion:
    """Description of an SGD step (torch.optim.SGD semantics) to fuse into the aggregation
    epilogue: `param`/`mom` are flat fp32 arenas of length d, `mirrors` are further arenas
    (typically peer-mapped) that receive the updated parameters."""

    def __init__(self, param, mom, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
                 first_step=False, mirrors=(), multicast=None):
        self.param, self.mom, self.lr, self.momentum = param, mom, lr, momentum
        self.dampening, self.weight_decay, self.nesterov = dampening, weight_decay, nesterov
        self.first_step, self.mirrors = first_step, tuple(mirrors)
        self.multicast = multicast          # NVLS multicast address of `param` (int) or None: publish by multimem.st

    def struct(self):
        u = FusedUpdate()
        u.mode, u.nesterov, u.first_step, u.n_mirrors = 1, int(self.nesterov), int(self.first_step), len(self.mirrors)
        u.lr, u.momentum, u.dampening, u.weight_decay = self.lr, self.momentum, self.dampening, self.weight_decay
        u.param = self.param.data_ptr()
        u.mom = self.mom.data_ptr() if self.mom is not None else None
        for k, t in enumerate(self.mirrors):
            u.mirror[k] = t.data_ptr() if isinstance(t, torch.Tensor) else int(t)
        u.multicast = int(self.multicast) if self.multicast else None
        return u


def coordwise(gradients, rule, f=0, beta=0, upper=False, p=0.9, seed=0, W=None, theta=0, out=None, update=None,
              want_out=True):
    """Run one per-coordinate rule. Returns the aggregated vector (None if only updating)."""
    rows, n, d, device = _prep(gradients)
    lib = _lib.require()
    code = RULES[rule]
    rank = f if rule == "trimmed_mean" else beta
    if rows[0].dtype == torch.float64:      # double-precision inputs: SIMT fp64 kernel (csrc/cuda/gar_f64.cu)
        if update is not None or rule == "condense":
            raise ValueError("float64 rows: no fused update / condense on the device path")
        if out is None:
            out = torch.empty(d, dtype=torch.float64, device=device)
        with torch.cuda.device(device):
            _lib.check(lib.gb_coordwise_f64(_lib.row_table(rows), n, d, code, int(rank), int(bool(upper)), W.data_ptr() if W is not None else None,
                                            int(theta), out.data_ptr(), _lib.stream_ptr(device)))
        _lib.count()
        return out
    if out is None and want_out:
        out = torch.empty(d, dtype=torch.float32, device=device)
    upd = update.struct() if update is not None else None
    with torch.cuda.device(device):
        _lib.check(lib.gb_coordwise(
            _lib.row_table(rows), n, d, code, int(rank), int(bool(upper)), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF,
            W.data_ptr() if W is not None else None, int(theta),
            out.data_ptr() if out is not None else None,
            ctypes.byref(upd) if upd is not None else None, _lib.stream_ptr(device)))
    _lib.count()
    return out


# `engine="auto"` is the EXACT direct-difference kernel for every n: the registry rules promise the reference's selections, and
# TF32 Gram distances (error ~1e-3·(‖x_i‖² + ‖x_j‖²) after the G_ii + G_jj − 2 G_ij cancellation) can reorder scores when honest
# gradients sit close together relative to their norm — or when an adversary aims a row into that error band. The tcgen05
# Gram kernel is opt-in: `engine="tc"` per call, or GARFIELD_GRAM_TC=<min rows> to route n >= that many rows (<= 32) to it.
import os as _os
TENSOR_CORE_MIN_ROWS = [int(_os.environ.get("GARFIELD_GRAM_TC", "0")) or 10 ** 9]


def select(gradients, mode, f=0, m=0, neighbors=0, squared=True, engine="auto", workspace_tag=0, max_ctas=0):
    """Distances + selection in one launch. Returns the device workspace dict with
    `dist` (n×n), `W` (n×n buffer, first θ rows valid) and `info`.

    engine: "simt" = exact direct-difference kernel, "tc" = tcgen05 TF32 Gram kernel (2..32
    aligned rows), "auto" = exact unless GARFIELD_GRAM_TC opts n >= TENSOR_CORE_MIN_ROWS rows into the Gram kernel."""
    rows, n, d, device = _prep(gradients)
    lib = _lib.require()
    if rows[0].dtype == torch.float64:
        ws = _Workspace.get(device, n, workspace_tag, torch.float64)
        with torch.cuda.device(device):
            _lib.check(lib.gb_pairwise_select_f64(
                _lib.row_table(rows), n, d, MODES[mode], int(f), int(m), int(neighbors), int(bool(squared)),
                ws["scratch"].data_ptr(), ws["scratch"].numel(), ws["dist"].data_ptr(), ws["W"].data_ptr(),
                ws["info"].data_ptr(), _lib.stream_ptr(device)))
        _lib.count()
        return ws
    ws = _Workspace.get(device, n, workspace_tag)
    aligned = all(r.data_ptr() % 16 == 0 for r in rows)
    use_tc = engine == "tc" or (engine == "auto" and TENSOR_CORE_MIN_ROWS[0] <= n <= 32 and aligned and d >= 4096)
    if use_tc:
        with torch.cuda.device(device):
            _lib.check(lib.gb_gram_select_tc(
                _lib.row_table(rows), n, d, MODES[mode], int(f), int(m), int(neighbors), int(bool(squared)),
                ws["scratch"].data_ptr(), ws["scratch"].numel(), ws["dist"].data_ptr(), ws["W"].data_ptr(),
                ws["info"].data_ptr(), _lib.stream_ptr(device)))
        _lib.count()
        return ws
    with torch.cuda.device(device):
        previous = lib.gb_pairwise_cta_limit(int(max_ctas)) if max_ctas else None      # background launch: a bounded share of the SMs
        try:
            _lib.check(lib.gb_pairwise_select(
                _lib.row_table(rows), n, d, MODES[mode], int(f), int(m), int(neighbors), int(bool(squared)),
                ws["scratch"].data_ptr(), ws["scratch"].numel(), ws["dist"].data_ptr(), ws["W"].data_ptr(),
                ws["info"].data_ptr(), _lib.stream_ptr(device)))
        finally:
            if previous is not None:
                lib.gb_pairwise_cta_limit(previous)
    _lib.count()
    return ws


def pairwise_sqdist(gradients, engine="simt"):
    """n×n matrix of squared L2 distances (device tensor; +inf diagonal and non-finite)."""
    return select(gradients, "none", engine=engine)["dist"].clone()


def weighted(gradients, W, theta=1, **kw):
    """out = W[0]·rows (θ = 1) — the selection-average pass."""
    return coordwise(gradients, "linear", W=W, theta=theta, **kw)


def krum(gradients, f, m, neighbors, squared, update=None, want_out=True, out=None):
    ws = select(gradients, "krum", f=f, m=m, neighbors=neighbors, squared=squared)
    return coordwise(gradients, "linear", W=ws["W"], theta=1, update=update, want_out=want_out, out=out)


def bulyan(gradients, f, m=None, native=False, update=None, want_out=True, out=None):
    n = len(gradients)
    m = n - f - 2 if m is None else m
    theta = n - 2 * f - 2
    ws = select(gradients, "bulyan_native" if native else "bulyan_python", f=f, m=m, squared=native)
    return coordwise(gradients, "averaged_median", beta=theta - 2 * f, upper=native, W=ws["W"], theta=theta,
                     update=update, want_out=want_out, out=out)


def mda(gradients, f, squared=False, update=None, want_out=True, out=None):
    ws = select(gradients, "mda", f=f, squared=squared)
    return coordwise(gradients, "linear", W=ws["W"], theta=1, update=update, want_out=want_out, out=out)


def aksel(gradients, c, update=None, want_out=True, out=None):
    med = coordwise(gradients, "median_lower")
    ws = select(list(gradients) + [med], "closest", m=c, squared=True)
    # W has n+1 columns (last = the median row, weight 0): reuse it on the extended row list
    return coordwise(list(gradients) + [med], "linear", W=ws["W"], theta=1, update=update, want_out=want_out, out=out)


_uniform = {}


def average(gradients, update=None, want_out=True, out=None):
    n = len(gradients)
    device = gradients[0].device
    key = (device, n)
    if key not in _uniform:
        _uniform[key] = torch.full((1, n), 1.0 / n, dtype=torch.float32, device=device)
    return coordwise(gradients, "linear", W=_uniform[key], theta=1, update=update, want_out=want_out, out=out)


def aggregate(name, gradients, f, m=None, update=None, want_out=True, out=None, **kw):
    """Run the registry rule `name` on device rows with an optional fused SGD epilogue.

    Same semantics as `aggregators.gars[name](gradients=..., f=...)` on CUDA tensors; the fused
    form additionally applies the optimizer step and publishes the parameters in the same pass.
    `out`: write the aggregate into this d-length tensor (it must not alias an input row).
    """
    n = len(gradients)
    common = dict(update=update, want_out=want_out, out=out)
    if name == "average":
        return average(gradients, **common)
    if name == "median":
        return coordwise(gradients, "median_lower", **common)
    if name == "native-median":
        return coordwise(gradients, "median_finite_upper", **common)
    if name == "trimmed_mean":
        return coordwise(gradients, "trimmed_mean", f=f, **common)
    if name == "averaged_median":
        return coordwise(gradients, "averaged_median", beta=kw.get("beta") or n - f, **common)
    if name == "condense":
        return coordwise(gradients, "condense", p=kw.get("p", 0.9), seed=kw.get("seed", 0), **common)
    if name == "krum":
        return krum(gradients, f, n - f - 2 if m is None else m, n - f - 1, False, **common)
    if name == "native-krum":
        return krum(gradients, f, n - f - 2 if m is None else m, n - f - 2, True, **common)
    if name == "bulyan":
        return bulyan(gradients, f, m, native=False, **common)
    if name == "native-bulyan":
        return bulyan(gradients, f, m, native=True, **common)
    if name in ("brute", "mda"):
        return mda(gradients, f, squared=False, **common)
    if name == "native-brute":
        return mda(gradients, f, squared=True, **common)
    if name == "aksel":
        mode = kw.get("mode", "mid")
        return aksel(gradients, (n + 1) // 2 if mode == "mid" else n - f, **common)
    raise KeyError("unknown aggregation rule %r" % (name,))


GOSSIP_RULES = {"median": ("median_lower", lambda n, f: 0), "trimmed_mean": ("trimmed_mean", lambda n, f: f),
                "averaged_median": ("averaged_median", lambda n, f: n - f)}


def gossip_rounds(name, rows0, rows1, out0, out1, f, rounds, first_parity, seq, counter, my_flag=None, peer_flags=(), timeout_ns=20_000_000_000,
                  late=None):
    """`rounds` agreement rounds of the per-coordinate rule `name` inside one persistent kernel (csrc/cuda/gar_gossip.cu).
    rows0 / rows1: all n nodes' vectors of parity 0 / 1 (local or peer-mapped); out0 / out1: the local nodes' buffers that
    receive the result of a round reading parity 0 / 1. `seq`, `counter`: int32 device words (counter is zeroed here)."""
    rule, rank_of = GOSSIP_RULES[name]
    n, d, device = len(rows0), rows0[0].numel(), rows0[0].device
    lib = _lib.require()
    counter.zero_()
    table = lambda ts: (ctypes.c_void_p * max(1, len(ts)))(*[t.data_ptr() if isinstance(t, torch.Tensor) else int(t) for t in ts])
    with torch.cuda.device(device):
        _lib.check(lib.gb_gossip_rounds(table(rows0), table(rows1), n, d, RULES[rule], int(rank_of(n, f)), 0, table(out0), table(out1), len(out0),
                                        int(rounds), int(first_parity), seq.data_ptr(), ctypes.c_void_p(my_flag) if my_flag else None,
                                        table(list(peer_flags)), len(peer_flags), counter.data_ptr(), int(timeout_ns),
                                        late.data_ptr() if late is not None else None, _lib.stream_ptr(device)))
    _lib.count()


def pull_median_operand(rows, rows_out, k, flat, operand):
    """Head of the MSMW model pull: lower median of the replicas' first-layer weights ([rows_out, k] each) into `flat` and,
    zero-padded, into `operand` [rows_out, kpad] (csrc/cuda/gar_pull.cu)."""
    lib = _lib.require()
    with torch.cuda.device(flat.device):
        _lib.check(lib.gb_pull_median_operand(_lib.row_table(rows), len(rows), rows_out, k, operand.shape[1], flat.data_ptr(), operand.data_ptr(),
                                              _lib.stream_ptr(flat.device)))
    _lib.count()


def select_only(ws, n, device, mode, f=0, m=0, neighbors=0, squared=True):
    """Run the selection logic on `ws["dist"]` (already summed over shards) → `ws["W"]`."""
    lib = _lib.require()
    with torch.cuda.device(device):
        _lib.check(lib.gb_select_only(ws["dist"].data_ptr(), n, MODES[mode], int(f), int(m), int(neighbors), int(bool(squared)),
                                      ws["W"].data_ptr(), ws["info"].data_ptr(), _lib.stream_ptr(device)))
    _lib.count()
    return ws


_DISTANCE_RULES = {
    # name: (selection mode, squared/native flag, neighbours(n, f), phase-2 builder)
    "krum": ("krum", False), "native-krum": ("krum", True), "bulyan": ("bulyan_python", False),
    "native-bulyan": ("bulyan_native", True), "brute": ("mda", False), "mda": ("mda", False), "native-brute": ("mda", True),
}


def distance_rules():
    """Names of the rules whose selection depends on pairwise distances (and can therefore be coordinate-sharded
    or have part of their distances computed ahead of time)."""
    return tuple(_DISTANCE_RULES)


def aggregate_sharded(name, slice_rows, f, allreduce, m=None, update=None, want_out=True, out=None, apply_rows=None, **kw):
    """Coordinate-sharded form of `aggregate`: `slice_rows` are this rank's coordinate slice of the
    n gradients; `allreduce(tensor)` sums a small device tensor over the ranks in place.

    Per-coordinate rules need no communication. Distance-based rules compute the partial squared
    distances of the slice, all-reduce the n×n matrix, run the (replicated) selection and apply the
    resulting weights to the slice — the robust analogue of a reduce-scatter.

    `apply_rows` (distance rules only): rows the selected weights are applied to instead of `slice_rows` — used when the
    distances of the remaining coordinates were computed elsewhere (e.g. during the backward pass) and `allreduce`
    adds them in, while the aggregate is still wanted over the whole vector."""
    n = len(slice_rows)
    common = dict(update=update, want_out=want_out, out=out)
    if name == "aksel":
        med = coordwise(slice_rows, "median_lower")
        rows = list(slice_rows) + [med]
        ws = select(rows, "none", engine="simt")
        allreduce(ws["dist"])
        mode = kw.get("mode", "mid")
        select_only(ws, n + 1, rows[0].device, "closest", m=(n + 1) // 2 if mode == "mid" else n - f, squared=True)
        return coordwise(rows, "linear", W=ws["W"], theta=1, **common)
    if name not in _DISTANCE_RULES:
        return aggregate(name, slice_rows, f, m=m, **common, **kw)
    mode, native = _DISTANCE_RULES[name]
    ws = select(slice_rows, "none", engine=kw.get("engine", "auto"))
    allreduce(ws["dist"])
    device = slice_rows[0].device
    target = slice_rows if apply_rows is None else apply_rows
    m_eff = n - f - 2 if m is None else m
    if mode == "krum":
        select_only(ws, n, device, "krum", f=f, m=m_eff, neighbors=n - f - 2 if native else n - f - 1, squared=native)
        return coordwise(target, "linear", W=ws["W"], theta=1, **common)
    if mode == "mda":
        select_only(ws, n, device, "mda", f=f, squared=native)
        return coordwise(target, "linear", W=ws["W"], theta=1, **common)
    theta = n - 2 * f - 2
    select_only(ws, n, device, mode, f=f, m=m_eff, squared=native)
    return coordwise(target, "averaged_median", beta=theta - 2 * f, upper=native, W=ws["W"], theta=theta, **common)
