#!/usr/bin/env python
"""Exact pairwise-distance kernel: time and accuracy per row count (CUDA events, 20 timed launches after 3 warm-ups, inputs
larger than L2). Run once per switch setting, e.g. `GARFIELD_PAIRWISE_MULTI=0 python scripts/pairwise_sweep.py`.
"""
import argparse
import json
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch

from garfield_b200.ops import gar

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, nargs="+", default=[13, 16, 20, 23, 24, 28, 32])
ap.add_argument("--d", type=int, default=23528522)
ap.add_argument("--out", default=None)
args = ap.parse_args()
dev = torch.device("cuda")
torch.manual_seed(0)
results = []
for n in args.rows:
    d = args.d
    g = torch.randn(n, (d + 63) // 64 * 64, device=dev)
    rows = [g[i][:d] for i in range(n)]
    got = gar.pairwise_sqdist(rows).double()
    # fp64 oracle on a prefix and a suffix slice plus the norm identity on the whole (cheap enough for 23.5 M coordinates)
    want = torch.zeros(n, n, dtype=torch.float64, device=dev)
    step = 1 << 21
    for lo in range(0, d, step):
        x = g[:, lo:min(d, lo + step)].double()
        sq = (x * x).sum(1)
        want += (sq[:, None] + sq[None, :] - 2.0 * x @ x.T).clamp_min(0)
    off = ~torch.eye(n, dtype=torch.bool, device=dev)
    rel = ((got - want).abs()[off] / want[off]).max().item()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        gar.pairwise_sqdist(rows)
    torch.cuda.synchronize()
    start.record()
    for _ in range(20):
        gar.pairwise_sqdist(rows)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / 20
    floor = n * d * 4 / 6.5e12 * 1e3
    rec = {"n": n, "d": d, "ms": round(ms, 4), "hbm_floor_ms": round(floor, 4), "max_rel_err_vs_fp64": rel,
           "multi": os.environ.get("GARFIELD_PAIRWISE_MULTI", "1"), "packed": os.environ.get("GARFIELD_PW_PACKED", "1")}
    print(json.dumps(rec), flush=True)
    results.append(rec)
    del g, rows
if args.out:
    pathlib.Path(args.out).write_text(json.dumps(results, indent=1))
