"""Console context, coloured logging helpers, user-facing exceptions and the plugin loader.

Capability parity with the reference's `tools` package front page
(/root/reference/pytorch_impl/libs/tools/__init__.py:26-305): `UserException`, a per-thread
nested `Context` that prefixes and colours everything printed inside it, the
`trace/info/success/warning/error/fatal` printers, an excepthook that prints `UserException`s
without a traceback, and `import_directory` (load every module of a directory and pull its
exported symbols into a scope).

Design differences: the context stack is a plain thread-local list consulted by a thin stream
proxy; nothing is installed on `sys.stdout` until `install()` is called (importing this module
has no global side effect, so it is safe under pytest and torchrun).
"""

from __future__ import annotations

import importlib.util
import os
import pathlib
import sys
import threading
import traceback

__all__ = [
    "UserException", "Context", "ContextIOWrapper", "trace", "info", "success", "warning",
    "error", "fatal", "install", "uninstall", "import_exported_symbols", "import_directory",
]


class UserException(Exception):
    """Error caused by the user's input; printed without a traceback when uncaught."""


_PALETTE = {
    "header": "\033[1;30m", "gray": "\033[1;30m", "trace": "\033[1;30m",
    "red": "\033[1;31m", "error": "\033[1;31m",
    "green": "\033[1;32m", "success": "\033[1;32m",
    "yellow": "\033[1;33m", "warning": "\033[1;33m",
    "blue": "\033[1;34m", "info": "\033[1;34m",
}
_RESET = "\033[0m"
_tls = threading.local()


def _frames():
    stack = getattr(_tls, "frames", None)
    if stack is None:
        stack = _tls.frames = []
    return stack


class Context:
    """Nested, per-thread output context: `with Context("server", "info"): print(...)`.

    The innermost non-None colour wins; names are concatenated outermost first as
    `[a] [b] `. Threads other than the main one get their name prepended.
    """

    def __init__(self, cntxtname, colorname):
        if colorname is not None and colorname not in _PALETTE:
            raise AssertionError("Unknown color name %r" % (colorname,))
        self._frame = (cntxtname, None if colorname is None else _PALETTE[colorname])

    def __enter__(self):
        _frames().append(self._frame)
        return self

    def __exit__(self, *exc):
        frames = _frames()
        if frames:
            frames.pop()
        return False

    @staticmethod
    def _get():
        """Returns (header text, header colour, body colour, reset code) for this thread."""
        names = [n for n, _ in _frames() if n is not None]
        colour = _RESET
        for _, c in reversed(_frames()):
            if c is not None:
                colour = c
                break
        head = "".join("[%s] " % n for n in names)
        th = threading.current_thread()
        if th is not threading.main_thread():
            head = "[%s] " % th.name + head
        return head, _PALETTE["header"], colour, _RESET


class ContextIOWrapper:
    """Text stream proxy that prefixes each new line with the thread's context header."""

    def __init__(self, output, nocolor=None):
        self._out = output
        self._bol = True
        if nocolor is None:
            isatty = getattr(output, "isatty", None)
            nocolor = not (callable(isatty) and isatty())
        self._nocolor = bool(nocolor)

    def __getattr__(self, name):
        return getattr(self._out, name)

    def write(self, text):
        if not text:
            return 0
        head, hclr, bclr, rst = Context._get()
        pieces = []
        for line in text.splitlines(True):
            if self._bol and (head or not self._nocolor):
                pieces.append(head if self._nocolor else hclr + head + rst + bclr)
            if self._nocolor:
                pieces.append(line)
            elif line.endswith("\n"):
                pieces.append(line[:-1] + rst + "\n")
            else:
                pieces.append(line)
            self._bol = line.endswith("\n")
        return self._out.write("".join(pieces))


def _printer(kind, stream_name="stdout"):
    def emit(*args, context=None, **kwargs):
        kwargs.setdefault("file", getattr(sys, stream_name))
        with Context(context, kind):
            print(*args, **kwargs)
    emit.__name__ = kind
    emit.__doc__ = "Print in the %r colour (optionally inside a named sub-context)." % kind
    return emit


trace = _printer("trace")
info = _printer("info")
success = _printer("success")
warning = _printer("warning", "stderr")
error = _printer("error", "stderr")


def fatal(*args, with_traceback=False, **kwargs):
    """Print an error (optionally with the current traceback) and exit(1)."""
    error(*args, **kwargs)
    if with_traceback:
        with Context("traceback", "trace"):
            traceback.print_exc()
    raise SystemExit(1)


_saved = {}


def _excepthook(etype, evalue, tb):
    if issubclass(etype, UserException):
        error("".join(str(evalue)).strip() or etype.__name__, context="fatal")
        return
    _saved.get("hook", sys.__excepthook__)(etype, evalue, tb)


def install():
    """Wrap sys.stdout/sys.stderr with `ContextIOWrapper` and hook uncaught `UserException`s."""
    if _saved:
        return
    _saved["stdout"], _saved["stderr"], _saved["hook"] = sys.stdout, sys.stderr, sys.excepthook
    sys.stdout = ContextIOWrapper(sys.stdout)
    sys.stderr = ContextIOWrapper(sys.stderr)
    sys.excepthook = _excepthook


def uninstall():
    if not _saved:
        return
    sys.stdout, sys.stderr, sys.excepthook = _saved.pop("stdout"), _saved.pop("stderr"), _saved.pop("hook")


def import_exported_symbols(name, module, scope):
    """Copy `module.__all__` (or, without one, every public name) into `scope`."""
    names = getattr(module, "__all__", None)
    if names is None:
        names = [n for n in vars(module) if not n.startswith("_")]
    for sym in names:
        if not hasattr(module, sym):
            warning("Symbol %r exported by %r does not exist" % (sym, name))
            continue
        if sym in scope and scope[sym] is not getattr(module, sym):
            warning("Symbol %r exported by %r overrides an existing symbol" % (sym, name))
        scope[sym] = getattr(module, sym)


def import_directory(dirpath, scope, post=import_exported_symbols, ignore=("__init__",)):
    """Import every `*.py` directly inside `dirpath` (sorted, minus `ignore`), then `post` it.

    Each module is executed with the caller's package as parent when `scope` has one, so
    relative imports inside plugins work; failures are reported and skipped, not raised.
    """
    dirpath = pathlib.Path(dirpath)
    package = scope.get("__package__") or scope.get("__name__")
    for path in sorted(dirpath.iterdir()):
        if path.suffix != ".py" or path.stem in ignore or not path.is_file():
            continue
        modname = path.stem
        qualname = ("%s.%s" % (package, modname)) if package else modname
        with Context(modname, None):
            try:
                if qualname in sys.modules:
                    module = sys.modules[qualname]
                else:
                    spec = importlib.util.spec_from_file_location(qualname, os.fspath(path))
                    module = importlib.util.module_from_spec(spec)
                    sys.modules[qualname] = module
                    try:
                        spec.loader.exec_module(module)
                    except BaseException:
                        sys.modules.pop(qualname, None)
                        raise
                if post is not None:
                    post(modname, module, scope)
            except Exception as err:  # a broken plugin must not take the registry down
                warning("Loading failed for module %r: %s" % (path.name, err))
                with Context("traceback", "trace"):
                    traceback.print_exc()
