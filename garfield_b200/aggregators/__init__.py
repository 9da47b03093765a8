"""Gradient aggregation rule (GAR) registry.

Public surface identical to the reference registry
(/root/reference/pytorch_impl/libs/aggregators/__init__.py:42-97):

    aggregators.gars                     name -> rule
    rule(gradients=[...], f=..., **kw)   keyword-only call, returns a NEW tensor
    rule.check / rule.checked / rule.unchecked / rule.upper_bound / rule.influence
    aggregators.register(name, unchecked, check, upper_bound=None, influence=None)
    aggregators.make_gar(...)

Registered names: average, median, krum, bulyan, brute (alias mda), aksel, condense,
native-median, native-krum, native-bulyan, native-brute, plus the additions trimmed_mean and
averaged_median (SURVEY.md §2.7 "Trimmed mean" note). Every rule accepts CPU or CUDA tensors;
on CUDA tensors the hand-written sm_100a kernels of `garfield_b200.ops` run (no host
round-trip), on CPU tensors the pure-torch / C++ host implementations run.
"""

from __future__ import annotations

import pathlib

from ..utils import console as _console

__all__ = ["gars", "register", "make_gar"]

gars = dict()


def make_gar(unchecked, check, upper_bound=None, influence=None, name=None):
    """Wrap a raw rule: the returned callable validates its keyword arguments first (unless
    Python runs with -O, as in the reference) and carries the auxiliary callables as attributes."""
    label = name if name is not None else getattr(unchecked, "__name__", "<gar>")

    def checked(**kwargs):
        problem = check(**kwargs)
        if problem is not None:
            raise _console.UserException(
                "Aggregation rule %r cannot be used with the given parameters: %s" % (label, problem))
        return unchecked(**kwargs)

    if __debug__:
        def rule(**kwargs):
            return checked(**kwargs)
    else:
        def rule(**kwargs):
            return unchecked(**kwargs)
    rule.__name__ = str(label)
    rule.__doc__ = unchecked.__doc__
    for attr, value in (("check", check), ("checked", checked), ("unchecked", unchecked),
                        ("upper_bound", upper_bound), ("influence", influence)):
        setattr(rule, attr, value)
    return rule


def register(name, unchecked, check, upper_bound=None, influence=None):
    """Add a rule to `gars` (first registration of a name wins, later ones warn)."""
    if name in gars:
        _console.warning("Unable to register %r GAR: name already in use" % name)
        return
    gars[name] = make_gar(unchecked, check, upper_bound=upper_bound, influence=influence, name=name)


def _export(name, module, scope):
    return None  # rule modules register themselves; nothing to copy into the package scope


with _console.Context("aggregators", None):
    _console.import_directory(pathlib.Path(__file__).parent, globals(), post=_export,
                              ignore=("__init__", "_host", "_device", "_common"))

for _name, _rule in gars.items():
    globals()[_name.replace("-", "_")] = _rule
    globals()[_name] = _rule
