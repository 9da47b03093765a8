#!/usr/bin/env python
"""GAR micro-benchmark: latency of every registered rule vs n, d and f.

Same sweeps as the reference script (/root/reference/pytorch_impl/applications/benchmarks/
gar_bench.py:40-89: n = 2^k at d=1e5, f=1; d = 10^k at n=23, f=5; f = 2^k at n=23, d=1e5) plus
the model-sized points of BASELINE.md (d = 62 006 / 23 528 522 / 134 301 514). Unlike the
reference, GPU timings use CUDA events with synchronisation on both sides, warm-up, and an L2
flush (a write larger than the 126 MB L2) between timed calls.

    python applications/benchmarks/gar_bench.py --device cuda --sweep models --json out.json
"""

import argparse
import json
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

import torch  # noqa: E402

import aggregators  # noqa: E402


def time_rule(rule, grads, f, device, reps, warmup=3, flush=None):
    """Median milliseconds per call."""
    times = []
    for it in range(warmup + reps):
        if device.type == "cuda":
            if flush is not None:
                flush.add_(1.0)
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            start.record()
            rule(gradients=grads, f=f)
            stop.record()
            torch.cuda.synchronize()
            ms = start.elapsed_time(stop)
        else:
            t0 = time.perf_counter()
            rule(gradients=grads, f=f)
            ms = (time.perf_counter() - t0) * 1e3
        if it >= warmup:
            times.append(ms)
    times.sort()
    return times[len(times) // 2]


def sweep_points(kind):
    if kind == "n":
        return [(2 ** k, 1, 100000) for k in range(0, 7)]
    if kind == "d":
        return [(23, 5, 10 ** k) for k in range(0, 6)]
    if kind == "f":
        return [(23, 2 ** k, 100000) for k in range(0, 4)]
    if kind == "models":
        return [(8, 1, 62006), (8, 1, 23528522), (8, 1, 134301514), (23, 5, 62006), (23, 5, 23528522)]
    raise SystemExit("unknown sweep %r" % kind)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--sweep", default="n,d,f", help="comma list of: n, d, f, models")
    ap.add_argument("--gars", default="", help="comma list (default: every registered rule)")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--max-bytes", type=float, default=12e9, help="skip points whose inputs exceed this many bytes")
    ap.add_argument("--json", default="")
    args = ap.parse_args(argv)
    device = torch.device(args.device)
    names = [g for g in args.gars.split(",") if g] or sorted(aggregators.gars)
    flush = torch.zeros(64 * 1024 * 1024, device=device) if device.type == "cuda" else None  # 256 MB > L2
    hbm = None
    if device.type == "cuda":
        peaks = pathlib.Path(__file__).resolve().parents[2] / "MEASURED_PEAKS.json"
        hbm = json.loads(peaks.read_text())["hbm_gbs"] if peaks.exists() else 6650.0
    rows = []
    for kind in args.sweep.split(","):
        for n, f, d in sweep_points(kind):
            if 4.0 * n * d > args.max_bytes:
                continue
            torch.manual_seed(1234)
            grads = [torch.randn(d, device=device) for _ in range(n)]
            for name in names:
                rule = aggregators.gars[name]
                if rule.check(gradients=grads, f=f) is not None:
                    continue
                if name in ("brute", "mda", "native-brute") and device.type == "cpu" and n > 16:
                    continue
                try:
                    ms = time_rule(rule, grads, f, device, args.reps, flush=flush)
                except Exception as err:  # keep sweeping
                    print("%-16s n=%-3d f=%-2d d=%-10d FAILED: %s" % (name, n, f, d, err), flush=True)
                    continue
                row = dict(sweep=kind, gar=name, n=n, f=f, d=d, ms=ms, device=str(device))
                if hbm is not None:
                    floor_ms = (4.0 * (n + 1) * d) / (hbm * 1e9) * 1e3  # one pass over inputs + output write
                    row["one_pass_hbm_floor_ms"] = floor_ms
                    row["frac_of_one_pass_floor"] = floor_ms / ms if ms > 0 else None
                rows.append(row)
                extra = "" if hbm is None else "  floor(1 pass) %.4f ms  frac %.2f" % (row["one_pass_hbm_floor_ms"], row["frac_of_one_pass_floor"])
                print("%-16s n=%-3d f=%-2d d=%-10d %10.4f ms%s" % (name, n, f, d, ms, extra), flush=True)
            del grads
    if args.json:
        pathlib.Path(args.json).write_text(json.dumps(rows, indent=1))
    return rows


if __name__ == "__main__":
    main()
