"""ncu target: a few launches of the tcgen05 GEMM at 4096^3 (bf16 and tf32) — nothing else on the GPU."""
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.ops import nn as gnn

for dtype in (torch.bfloat16, torch.float32):
    a = torch.randn(4096, 4096, device="cuda").to(dtype)
    b = torch.randn(4096, 4096, device="cuda").to(dtype)
    for _ in range(3):
        c = gnn.gemm_nt(a, b)
    torch.cuda.synchronize()
print("done", float(c[0, 0]))
