"""Step-phase tracing and metrics.

`Tracer` times named phases (`get_gradients`, `gar`, `update_model`, `get_models`, `write_model`…)
with CUDA events on the current stream (or wall clock on CPU), emits NVTX ranges when CUDA is
present, aggregates per-phase statistics and can dump a chrome://tracing JSON. It replaces the
reference's `torch.autograd.profiler` tables and ad-hoc per-phase timer lists
(applications/Aggregathor/trainer.py:234-239, Garfield_CC/trainer.py:63-70,198-208).
`Metrics` is a tiny append-only JSONL logger for loss / accuracy / images-per-second records.
"""

from __future__ import annotations

import contextlib
import json
import time

import torch

__all__ = ["Tracer", "Metrics"]


class Tracer:
    def __init__(self, enabled=True, device=None):
        self.enabled = enabled
        self.cuda = torch.cuda.is_available() and (device is None or str(device).startswith("cuda"))
        self._pending, self.records, self._origin = [], [], time.perf_counter()

    @contextlib.contextmanager
    def phase(self, name, step=None):
        if not self.enabled:
            yield
            return
        t0 = time.perf_counter()
        if self.cuda:
            torch.cuda.nvtx.range_push(name)
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
        try:
            yield
        finally:
            if self.cuda:
                stop.record()
                torch.cuda.nvtx.range_pop()
                self._pending.append((name, step, t0 - self._origin, start, stop))
            else:
                self.records.append(dict(name=name, step=step, ts=t0 - self._origin, ms=(time.perf_counter() - t0) * 1e3))

    def flush(self):
        """Resolve pending CUDA-event timings (synchronises once)."""
        if self._pending:
            torch.cuda.synchronize()
            for name, step, ts, start, stop in self._pending:
                self.records.append(dict(name=name, step=step, ts=ts, ms=start.elapsed_time(stop)))
            self._pending = []
        return self.records

    def summary(self):
        stats = {}
        for rec in self.flush():
            s = stats.setdefault(rec["name"], dict(count=0, total_ms=0.0, max_ms=0.0))
            s["count"] += 1
            s["total_ms"] += rec["ms"]
            s["max_ms"] = max(s["max_ms"], rec["ms"])
        for s in stats.values():
            s["mean_ms"] = s["total_ms"] / s["count"]
        return stats

    def dump_chrome_trace(self, path, pid=0):
        events = [dict(name=r["name"], ph="X", ts=r["ts"] * 1e6, dur=r["ms"] * 1e3, pid=pid, tid=0, args={"step": r["step"]})
                  for r in self.flush()]
        with open(path, "w") as fh:
            json.dump({"traceEvents": events}, fh)
        return path


class Metrics:
    def __init__(self, path=None):
        self.path, self.rows = path, []

    def log(self, **fields):
        fields.setdefault("time", time.time())
        self.rows.append(fields)
        if self.path:
            with open(self.path, "a") as fh:
                fh.write(json.dumps(fields) + "\n")
        return fields
