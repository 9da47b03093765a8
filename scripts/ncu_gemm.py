#!/usr/bin/env python
"""A few launches of `gemm_nt` at one shape — the target of an `ncu -k regex:gemm_tc -c 1` capture.
usage: ncu_gemm.py M N K [bf16|tf32]"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch

from garfield_b200.ops import nn as gnn

m, n, k = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 4096, 4096)
dtype = torch.bfloat16 if (len(sys.argv) > 4 and sys.argv[4] == "bf16") else torch.float32
a = torch.randn(m, k, device="cuda").to(dtype)
b = torch.randn(n, k, device="cuda").to(dtype)
for _ in range(3):
    c = gnn.gemm_nt(a, b)
torch.cuda.synchronize()
print("ok", float(c[0, 0]))
