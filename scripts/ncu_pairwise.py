#!/usr/bin/env python
"""One launch of the exact pairwise-distance kernel at (n, d) — the target of an `ncu -k regex:pairwise -c 1` capture."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch

from garfield_b200.ops import gar

n = int(sys.argv[1]) if len(sys.argv) > 1 else 23
d = int(sys.argv[2]) if len(sys.argv) > 2 else 23528522
g = torch.randn(n, (d + 63) // 64 * 64, device="cuda")
rows = [g[i][:d] for i in range(n)]
for _ in range(2):
    out = gar.pairwise_sqdist(rows)
torch.cuda.synchronize()
print("ok", float(out[0, 1]))
