#!/usr/bin/env python
"""Small driver for `ncu --set full -k regex:conv_tc_kernel`: one forward, data-gradient and weight-gradient launch of two
ResNet-50 layer shapes (3×3 64→64 on 8×8 planes, 1×1 1024→256 on 2×2 planes; batch 200 = 8 workers × 25), after a warm-up."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch

from garfield_b200.ops import conv as C

for (n, h, w, cin, cout, k, p) in ((200, 8, 8, 64, 64, 3, 1), (200, 2, 2, 1024, 256, 1, 0)):
    x = torch.randn(n, h, w, cin, device="cuda")
    wt = torch.randn(cout, k, k, cin, device="cuda") / (k * k * cin) ** 0.5
    dy = torch.randn(n, h, w, cout, device="cuda")
    for _ in range(3):
        C.conv_fprop(x, wt, None, 1, p)
        C.conv_dgrad(dy, wt, (h, w), 1, p)
        C.conv_wgrad(x, dy, 8, k, 1, p)
    torch.cuda.synchronize()
print("ok")
