#!/usr/bin/env python
"""Tiny driver for ncu: a few calls of the GAR kernels at the ResNet-50 size (n=8, d=23 528 522)."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import aggregators  # noqa: E402

n, d = int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 23528522
names = sys.argv[3].split(",") if len(sys.argv) > 3 else ["average", "median", "krum", "bulyan"]
torch.manual_seed(0)
g = [torch.randn(d, device="cuda") for _ in range(n)]
for _ in range(3):
    for name in names:
        aggregators.gars[name](gradients=g, f=1)
torch.cuda.synchronize()
for name in names:
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(10):
        aggregators.gars[name](gradients=g, f=1)
    stop.record()
    torch.cuda.synchronize()
    print("%-10s %.4f ms/call (10 back-to-back calls, inputs %.0f MB > L2)" % (name, start.elapsed_time(stop) / 10, 4e-6 * n * d))
