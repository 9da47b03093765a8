#!/usr/bin/env python
"""Tiny driver for ncu / timing: the GAR kernels at the ResNet-50 size (d = 23 528 522).
usage: profile_gar.py [n] [d] [comma list of rules] [f]"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import aggregators  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
d = int(sys.argv[2]) if len(sys.argv) > 2 else 23528522
names = sys.argv[3].split(",") if len(sys.argv) > 3 else ["average", "median", "krum", "bulyan"]
f = int(sys.argv[4]) if len(sys.argv) > 4 else 1
torch.manual_seed(0)
g = [torch.randn(d, device="cuda") for _ in range(n)]
for _ in range(3):
    for name in names:
        aggregators.gars[name](gradients=g, f=f)
torch.cuda.synchronize()
for name in names:
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(10):
        aggregators.gars[name](gradients=g, f=f)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / 10
    print("%-14s n=%-2d f=%d  %.4f ms/call  (inputs %.0f MB > L2; one-pass HBM floor %.4f ms at 6542 GB/s)" % (
        name, n, f, ms, 4e-6 * n * d, 4.0 * (n + 1) * d / 6542.1e9 * 1e3))
