#!/usr/bin/env python
"""A few launches of one registry rule at (n, f, d) — the target of an `ncu -k regex:coordwise -c 1` capture.
usage: ncu_coordwise.py RULE N F [D]"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch

from garfield_b200.ops import gar

rule = sys.argv[1] if len(sys.argv) > 1 else "median"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 23
f = int(sys.argv[3]) if len(sys.argv) > 3 else 5
d = int(sys.argv[4]) if len(sys.argv) > 4 else 23528522
g = torch.randn(n, (d + 63) // 64 * 64, device="cuda")
rows = [g[i][:d] for i in range(n)]
for _ in range(3):
    out = gar.aggregate(rule, rows, f)
torch.cuda.synchronize()
print("ok", float(out[0]))
