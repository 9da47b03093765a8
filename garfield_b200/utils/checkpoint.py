"""torch `state_dict` checkpoints (SURVEY.md §5.4 — new feature; the PT reference has none, the TF
one has `tools.tf.Checkpoints`, tensorflow_impl/rsrcs/tools/tf.py:78-173).

A checkpoint is one `torch.save` file:
    {"model": model.state_dict(), "optimizer": optimizer.state_dict(), "step": int,
     "rng": {"cpu": ByteTensor, "cuda": [ByteTensor...]}, "config": dict, "format": 1}
Parameters are views of a flat arena, so loading copies INTO the views (aliasing is preserved)
and a resumed run continues bit-exactly (same arena contents, optimizer state and RNG streams).
"""

from __future__ import annotations

import os
import pathlib
import re

import torch

__all__ = ["save_checkpoint", "load_checkpoint", "Checkpoints"]


def _rng_state():
    state = {"cpu": torch.get_rng_state()}
    if torch.cuda.is_available():
        state["cuda"] = torch.cuda.get_rng_state_all()
    return state


def save_checkpoint(path, model, optimizer=None, step=0, config=None, extra=None):
    """Atomically write a checkpoint of `model` (module or FlatModel) and `optimizer`."""
    path = pathlib.Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    payload = {"format": 1, "step": int(step), "config": dict(config or {}), "rng": _rng_state(),
               "model": {k: v.detach().cpu() for k, v in model.state_dict().items()}}
    if optimizer is not None:
        payload["optimizer"] = optimizer.state_dict()
    if extra:
        payload["extra"] = extra
    tmp = path.with_suffix(path.suffix + ".tmp%d" % os.getpid())
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path, model, optimizer=None, restore_rng=True, map_location="cpu"):
    """Load a checkpoint into `model`/`optimizer` in place; returns the payload (step, config…)."""
    payload = torch.load(path, map_location=map_location, weights_only=False)
    if hasattr(model, "module") and hasattr(model, "flat"):  # FlatModel: copy into the arena views
        model.load_state_dict(payload["model"])
    else:
        own = model.state_dict()
        with torch.no_grad():
            for key, tensor in own.items():
                tensor.copy_(payload["model"][key])
    if optimizer is not None and "optimizer" in payload:
        optimizer.load_state_dict(payload["optimizer"])
    if restore_rng and "rng" in payload:
        torch.set_rng_state(payload["rng"]["cpu"].cpu())
        if torch.cuda.is_available() and "cuda" in payload["rng"]:
            states = [s.cpu() for s in payload["rng"]["cuda"]]
            if len(states) == torch.cuda.device_count():
                torch.cuda.set_rng_state_all(states)
    return payload


class Checkpoints:
    """Directory manager: files `<base>-<step>.pt`; `save`, `list`, `latest`, `restore`, pruning."""

    def __init__(self, directory, base="model", keep=3):
        self.dir, self.base, self.keep = pathlib.Path(directory), base, keep
        self._pattern = re.compile(re.escape(base) + r"-(\d+)\.pt$")

    def list(self):
        if not self.dir.exists():
            return []
        found = []
        for entry in self.dir.iterdir():
            match = self._pattern.match(entry.name)
            if match:
                found.append((int(match.group(1)), entry))
        return sorted(found)

    def latest(self):
        items = self.list()
        return items[-1] if items else None

    def save(self, model, optimizer, step, config=None):
        path = save_checkpoint(self.dir / ("%s-%d.pt" % (self.base, step)), model, optimizer, step, config)
        if self.keep:
            for _, old in self.list()[:-self.keep]:
                old.unlink()
        return path

    def restore(self, model, optimizer=None, step=None):
        items = self.list()
        if not items:
            return None
        chosen = items[-1] if step is None else next((it for it in items if it[0] == step), None)
        if chosen is None:
            raise FileNotFoundError("no checkpoint for step %r in %s" % (step, self.dir))
        return load_checkpoint(chosen[1], model, optimizer)
