"""Small general-purpose helpers (registries, key:value parsing, timing, 1-D search, pair iteration).

Capability parity with /root/reference/pytorch_impl/libs/tools/misc.py:35-569 — same public
names and call signatures so reference user code keeps working — re-implemented from the
behaviour described in SURVEY.md §2.1 P14.
"""

from __future__ import annotations

import code
import itertools
import os
import pathlib
import site
import sys
import threading
import time

from .console import Context, UserException, fatal, trace

__all__ = [
    "make_unavailable_exception_text", "fatal_unavailable", "UnavailableException",
    "MethodCallReplicator", "ClassRegister", "parse_keyval_auto_convert", "parse_keyval",
    "fullqual", "onetime", "TimedContext", "interactive", "get_loaded_dependencies",
    "line_maximize", "pairwise", "localtime", "deltatime_point", "deltatime_format",
]


def make_unavailable_exception_text(data, name, what="entry"):
    """Explain that `name` is not one of the `what`s listed in `data`."""
    data = list(data)
    if not data:
        tail = "no %s available" % what
    else:
        bullet = "%s· " % os.linesep
        tail = "expected one of:" + "".join(bullet + str(item) for item in data)
    return "Unknown %s %r, %s" % (what, name, tail)


def fatal_unavailable(*args, **kwargs):
    fatal(make_unavailable_exception_text(*args, **kwargs))


class UnavailableException(UserException):
    """Raised when a user-selected name is missing from a set of choices."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        self._text = make_unavailable_exception_text(*args, **kwargs)

    def __str__(self):
        return self._text


class MethodCallReplicator:
    """`MethodCallReplicator(a, b).foo(x)` == `[a.foo(x), b.foo(x)]`."""

    def __init__(self, *instances):
        if not instances:
            raise AssertionError("Expected at least one instance on which to forward method calls")
        self._targets = instances

    def __getattr__(self, name):
        bound = [getattr(target, name) for target in self._targets]
        return lambda *a, **kw: [method(*a, **kw) for method in bound]


class ClassRegister:
    """Name -> class registry with a helpful error for unknown names."""

    def __init__(self, singular, optplural=None):
        self._one = singular
        self._many = optplural if optplural is not None else singular + "(s)"
        self._table = {}

    def itemize(self):
        return self._table.keys()

    def register(self, name, cls):
        if name in self._table:
            raise AssertionError("Name %r already in use while registering %r" % (
                name, getattr(cls, "__name__", "<unknown %s class name>" % self._one)))
        self._table[name] = cls

    def instantiate(self, name, *args, **kwargs):
        try:
            ctor = self._table[name]
        except KeyError:
            if self._table:
                hint = "available %s: '%s'" % (self._many, "', '".join(self._table))
            else:
                hint = "no registered %s" % self._one
            raise UserException("Unknown name %r, %s" % (name, hint)) from None
        return ctor(*args, **kwargs)


def parse_keyval_auto_convert(val):
    """Guess bool / int / float, else keep the string."""
    lowered = val.lower()
    if lowered in ("true", "false"):
        return lowered == "true"
    for caster in (int, float):
        try:
            return caster(val)
        except ValueError:
            pass
    return val


def parse_keyval(list_keyval, defaults={}):
    """Turn `["lr:0.1", "nesterov:true"]` into a dict; `defaults` fixes types and fills gaps."""
    out = {}
    for item in list_keyval:
        key, sep, raw = item.partition(":")
        if not sep:
            raise UserException("Expected list of %r, got %r as one entry" % ("<key>:<value>", item))
        if key in out:
            raise UserException("Key %r had already been specified with value %r" % (key, out[key]))
        if key in defaults:
            kind = type(defaults[key])
            try:
                out[key] = (raw.lower() not in ("", "0", "n", "false")) if kind is bool else kind(raw)
            except Exception:
                raise UserException("Required key %r expected a value of type %r" % (
                    key, getattr(kind, "__name__", "<unknown>"))) from None
        else:
            out[key] = parse_keyval_auto_convert(raw)
    for key, value in defaults.items():
        out.setdefault(key, value)
    return out


def fullqual(obj):
    """`'pkg.mod.Class'` for classes, `'instance of pkg.mod.Class'` otherwise."""
    prefix = ""
    if not isinstance(obj, type):
        prefix, obj = "instance of ", type(obj)
    return "%s%s.%s" % (prefix, getattr(obj, "__module__", "<unknown module>"),
                        getattr(obj, "__qualname__", "<unknown name>"))


_onetime_flags = {}
_onetime_guard = threading.Lock()


def onetime(name=None):
    """A thread-safe set-once flag: returns `(is_set, set)` closures (shared when named)."""
    with _onetime_guard:
        if name is not None and name in _onetime_flags:
            return _onetime_flags[name]
        event = threading.Event()
        pair = (lambda *a, **k: event.is_set(), lambda *a, **k: event.set())
        if name is not None:
            _onetime_flags[name] = pair
        return pair


class TimedContext(Context):
    """`Context` that prints the wall time spent inside it on exit."""

    def __enter__(self):
        self._t0 = time.perf_counter()
        return super().__enter__()

    def __exit__(self, *exc):
        span = (time.perf_counter() - self._t0) * 1e9
        unit = "ns"
        for unit in ("ns", "µs", "ms", "s"):
            if span < 1000.0 or unit == "s":
                break
            span /= 1000.0
        trace("Execution time: %.3f %s" % (span, unit))
        return super().__exit__(*exc)


def interactive(glbs=None, lcls=None, prompt=">>> ", cprmpt="... "):
    """Drop into a REPL over the caller's globals/locals (EOF to continue)."""
    frame = sys._getframe(1)
    namespace = dict(frame.f_globals if glbs is None else glbs)
    namespace.update(frame.f_locals if lcls is None else lcls)
    console = code.InteractiveConsole(namespace)
    sys.ps1, sys.ps2 = prompt, cprmpt
    try:
        console.interact(banner="", exitmsg="")
    except SystemExit:
        pass


def get_loaded_dependencies():
    """List `(name, version or None, flavor)` for every loaded top-level module.

    flavor: `IS_PYTHON` (stdlib), `IS_SITE` (site-packages) or `IS_LOCAL` (anything else).
    """
    std_root = pathlib.Path(os.__file__).resolve().parent
    site_roots = []
    for getter in (getattr(site, "getsitepackages", lambda: []), lambda: [site.getusersitepackages()]):
        try:
            site_roots += [pathlib.Path(p).resolve() for p in getter()]
        except Exception:
            pass

    def flavor_of(path):
        path = pathlib.Path(path).resolve()
        if any(root in path.parents for root in site_roots):
            return get_loaded_dependencies.IS_SITE
        if std_root in path.parents:
            return get_loaded_dependencies.IS_PYTHON
        return get_loaded_dependencies.IS_LOCAL

    found = []
    for name, module in list(sys.modules.items()):
        if "." in name or name.startswith("_"):
            continue
        origin = getattr(module, "__file__", None)
        flavor = get_loaded_dependencies.IS_PYTHON if origin is None else flavor_of(origin)
        version = getattr(module, "__version__", None)
        found.append((name, None if version is None else str(version), flavor))
    return sorted(found)


get_loaded_dependencies.IS_PYTHON = 0
get_loaded_dependencies.IS_SITE = 1
get_loaded_dependencies.IS_LOCAL = 2


def line_maximize(scape, evals=16, start=0., delta=1., ratio=0.8):
    """Best-effort argmax of `scape` over x ≥ 0 within an evaluation budget.

    Doubling steps while the value improves, then shrinking steps (factor `ratio`)
    bracketing the best point found. Same exploration schedule as the reference helper
    (tools/misc.py:467-513) so `influence`-style searches give comparable answers.
    """
    best_x, best_y = start, scape(start)
    budget = evals - 1
    probe = best_x
    while budget > 0:  # expansion
        probe = best_x + delta
        value = scape(probe)
        budget -= 1
        if value > best_y:
            best_x, best_y = probe, value
            delta *= 2
        else:
            delta *= ratio
            break
    while budget > 0:  # contraction around the best point
        if probe < best_x:
            probe += delta
        else:
            cand = probe - delta
            while cand < 0:
                cand = (cand + probe) / 2
            probe = cand
        value = scape(probe)
        budget -= 1
        if value > best_y:
            best_x, best_y = probe, value
        delta *= ratio
    return best_x


def pairwise(data):
    """All `(data[i], data[j])` with i < j, in lexicographic index order."""
    return itertools.combinations(tuple(data[i] for i in range(len(data))), 2)


def localtime():
    """Current local time as `YYYY/MM/DD hh:mm:ss`."""
    return time.strftime("%Y/%m/%d %H:%M:%S", time.localtime())


def deltatime_point():
    """Opaque time point for `deltatime_format`."""
    return time.monotonic()


def deltatime_format(a, b):
    """Human-readable span between two `deltatime_point()`s, e.g. `1 h 2 min 3 s 40 ms`."""
    millis = int(round(abs(b - a) * 1000))
    parts = []
    for label, size in (("day", 86400000), ("h", 3600000), ("min", 60000), ("s", 1000), ("ms", 1)):
        count, millis = divmod(millis, size)
        if count:
            parts.append("%d %s" % (count, label))
    return " ".join(parts) if parts else "0 ms"
