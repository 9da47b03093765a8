"""`native` — C++ host implementations of the aggregation rules, bound through ctypes.

API parity with the reference's pybind11 modules (/root/reference/pytorch_impl/libs/native,
SURVEY.md Appendix A):

    native.krum.aggregate(list, f, m)     native.bulyan.aggregate(list, f, m)
    native.median.aggregate(list)         native.brute.aggregate(list, f)

Inputs: list of n contiguous 1-D float32/float64 tensors of equal length. CPU tensors run the
C++ host library (libgarfield_b200_cpu.so); CUDA tensors are routed to the sm_100a kernels
(garfield_b200.ops.gar) — the counterpart of the reference's `.cu` paths.
"""

from __future__ import annotations

import ctypes
import types

import torch

from .. import _build

__all__ = ["available", "supports", "krum", "bulyan", "median", "brute", "pairwise_sqdist", "coordwise",
           "combinations", "num_threads", "lib"]

_lib = None
_err = None


def lib():
    """The ctypes handle of libgarfield_b200_cpu.so (built on first use)."""
    global _lib, _err
    if _lib is None and _err is None:
        try:
            handle = _build.load_cpu()
            _declare(handle)
            _lib = handle
        except Exception as err:  # keep the import of `native` itself cheap and safe
            _err = err
    if _lib is None:
        raise RuntimeError("host native library unavailable: %s" % (_err,))
    return _lib


def available():
    try:
        lib()
        return True
    except RuntimeError:
        return False


def _declare(h):
    pp, i32, i64 = ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int64
    vp = ctypes.c_void_p
    h.gb_cpu_last_error.restype = ctypes.c_char_p
    h.gb_cpu_num_threads.restype = i32
    for suffix in ("f32", "f64"):
        getattr(h, "gb_cpu_krum_" + suffix).argtypes = [pp, i32, i64, i32, i32, vp]
        getattr(h, "gb_cpu_bulyan_" + suffix).argtypes = [pp, i32, i64, i32, i32, vp]
        getattr(h, "gb_cpu_median_" + suffix).argtypes = [pp, i32, i64, vp]
        getattr(h, "gb_cpu_brute_" + suffix).argtypes = [pp, i32, i64, i32, vp]
        getattr(h, "gb_cpu_pairwise_sqdist_" + suffix).argtypes = [pp, i32, i64, vp]
        getattr(h, "gb_cpu_selection_average_" + suffix).argtypes = [pp, i64, vp, i32, vp]
    h.gb_combinations_count.restype = ctypes.c_uint64
    h.gb_combinations_count.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.POINTER(ctypes.c_int)]
    h.gb_combinations_seek.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.c_uint64, ctypes.POINTER(ctypes.c_int)]
    h.gb_combinations_next.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.POINTER(ctypes.c_int)]
    h.gb_cpu_parallel_for_probe.argtypes = [i64, i64, vp]
    for suffix in ("f32", "f64"):
        getattr(h, "gb_cpu_coordwise_" + suffix).argtypes = [pp, i32, i64, i32, i32, vp]
    h.gb_cpu_axis_nth_f32.argtypes = [vp, i32, ctypes.POINTER(ctypes.c_int64), i32, i64, vp]
    h.gb_mreg_create.restype = vp
    h.gb_mreg_create.argtypes = [i32, i64, i32]
    h.gb_mreg_close.argtypes = [vp]
    h.gb_mreg_destroy.argtypes = [vp]
    h.gb_mreg_write.argtypes = [vp, vp, i64]
    h.gb_mreg_read.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_uint64), i64]


def supports(gradients):
    """True when the list can be handed to the host library as is (CPU, fp32/fp64)."""
    g0 = gradients[0]
    return isinstance(g0, torch.Tensor) and g0.device.type == "cpu" and g0.dtype in (torch.float32, torch.float64)


def _validate(inputs):
    if not isinstance(inputs, (list, tuple)) or len(inputs) < 1:
        raise ValueError("expected a non-empty list of tensors")
    first = inputs[0]
    for t in inputs:
        if not isinstance(t, torch.Tensor) or t.dim() != 1:
            raise ValueError("expected 1-D tensors")
        if t.shape != first.shape or t.dtype != first.dtype or t.device != first.device:
            raise ValueError("all tensors must share size, dtype and device")
    if first.numel() == 0:
        raise ValueError("expected non-empty tensors")
    if first.dtype not in (torch.float32, torch.float64):
        raise TypeError("unsupported dtype %s (float32 / float64 only)" % first.dtype)
    return [t if t.is_contiguous() else t.contiguous() for t in inputs]


def _rows(inputs):
    table = (ctypes.c_void_p * len(inputs))(*[t.data_ptr() for t in inputs])
    return table


def _suffix(t):
    return "f32" if t.dtype == torch.float32 else "f64"


def _call(name, inputs, *args):
    inputs = _validate(inputs)
    out = torch.empty_like(inputs[0])
    fn = getattr(lib(), "%s_%s" % (name, _suffix(inputs[0])))
    rc = fn(_rows(inputs), len(inputs), inputs[0].numel(), *args, out.data_ptr())
    if rc != 0:
        raise ValueError(lib().gb_cpu_last_error().decode())
    return out


def _device(inputs):
    from ..ops import gar
    if not gar.available():
        raise RuntimeError("CUDA tensors given but the sm_100a library is unavailable")
    return gar


def _krum(inputs, f, m=None):
    """Multi-Krum (native semantics: squared distances, n−f−2 neighbours)."""
    n = len(inputs)
    m = n - f - 2 if m is None else m
    if inputs[0].is_cuda:
        return _device(inputs).krum(_validate(inputs), f=f, m=m, neighbors=n - f - 2, squared=True)
    return _call("gb_cpu_krum", inputs, int(f), int(m))


def _bulyan(inputs, f, m=None):
    """Bulyan over Multi-Krum (native semantics)."""
    n = len(inputs)
    m = n - f - 2 if m is None else m
    if inputs[0].is_cuda:
        return _device(inputs).bulyan(_validate(inputs), f=f, m=m, native=True)
    return _call("gb_cpu_bulyan", inputs, int(f), int(m))


def _median(inputs):
    """Non-finite-resilient coordinate-wise median."""
    if inputs[0].is_cuda:
        return _device(inputs).coordwise(_validate(inputs), "median_finite_upper")
    return _call("gb_cpu_median", inputs)


def _brute(inputs, f):
    """Brute / MDA (native semantics)."""
    if inputs[0].is_cuda:
        return _device(inputs).mda(_validate(inputs), f=f, squared=True)
    return _call("gb_cpu_brute", inputs, int(f))


_COORDWISE_RULES = {"median_lower": 0, "trimmed_mean": 1, "averaged_median": 2}


def coordwise(inputs, rule, arg=0):
    """Python-semantics coordinate-wise rules on CPU tensors through the vectorised host kernel (n <= 32):
    `median_lower` (torch.median semantics), `trimmed_mean` (arg = f), `averaged_median` (arg = beta).
    Returns None when the fast path declines (too many rows, or a column with NaN / non-finite values whose
    reference semantics it does not reproduce) — the caller then uses the reference formulation."""
    inputs = _validate(inputs)
    out = torch.empty_like(inputs[0])
    fn = getattr(lib(), "gb_cpu_coordwise_" + _suffix(inputs[0]))
    rc = fn(_rows(inputs), len(inputs), inputs[0].numel(), _COORDWISE_RULES[rule], int(arg), out.data_ptr())
    if rc == 1:
        raise ValueError(lib().gb_cpu_last_error().decode())
    return out if rc == 0 else None


krum = types.SimpleNamespace(aggregate=_krum)
bulyan = types.SimpleNamespace(aggregate=_bulyan)
median = types.SimpleNamespace(aggregate=_median)
brute = types.SimpleNamespace(aggregate=_brute)


def pairwise_sqdist(inputs):
    """n×n float64 matrix of squared L2 distances (+inf diagonal), host library."""
    inputs = _validate(inputs)
    n = len(inputs)
    out = torch.empty(n, n, dtype=torch.float64)
    fn = getattr(lib(), "gb_cpu_pairwise_sqdist_" + _suffix(inputs[0]))
    if fn(_rows(inputs), n, inputs[0].numel(), out.data_ptr()) != 0:
        raise ValueError(lib().gb_cpu_last_error().decode())
    return out


def num_threads():
    return int(lib().gb_cpu_num_threads())


class _Combinations:
    """Lexicographic k-combinations of range(n): `count`, `seek(pos)`, iteration from a rank."""

    @staticmethod
    def count(n, k):
        over = ctypes.c_int(0)
        value = lib().gb_combinations_count(n, k, ctypes.byref(over))
        if over.value:
            raise OverflowError("C(%d, %d) does not fit 64 bits" % (n, k))
        return int(value)

    @staticmethod
    def seek(n, k, pos):
        idx = (ctypes.c_int * k)()
        lib().gb_combinations_seek(n, k, pos, idx)
        return tuple(idx)

    @staticmethod
    def iterate(n, k, start=0, stop=None):
        total = _Combinations.count(n, k)
        stop = total if stop is None else min(stop, total)
        if start >= stop:
            return
        idx = (ctypes.c_int * k)()
        lib().gb_combinations_seek(n, k, start, idx)
        for _ in range(start, stop):
            yield tuple(idx)
            lib().gb_combinations_next(n, k, idx)


combinations = _Combinations
