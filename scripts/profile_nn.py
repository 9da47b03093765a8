#!/usr/bin/env python
"""Tiny driver for ncu: the tcgen05 GEMM at a large size and the fused NHWC BatchNorm kernels at ResNet-50 shapes."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

from garfield_b200.ops import nn as gnn  # noqa: E402

torch.manual_seed(0)
results = []
for dtype, name in ((torch.bfloat16, "bf16"), (torch.float32, "tf32")):
    m = n = k = 4096
    a = torch.randn(m, k, device="cuda").to(dtype)
    b = torch.randn(n, k, device="cuda").to(dtype)
    for _ in range(3):
        gnn.gemm_nt(a, b)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        gnn.gemm_nt(a, b)
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 10
    if dtype == torch.float32:
        torch.backends.cuda.matmul.allow_tf32 = True
    for _ in range(3):
        a @ b.t()
    torch.cuda.synchronize()
    s.record()
    for _ in range(10):
        a @ b.t()
    e.record(); torch.cuda.synchronize()
    ms_lib = s.elapsed_time(e) / 10
    print("gemm_nt %s 4096^3: ours %.3f ms (%.0f TFLOP/s)   cuBLAS %.3f ms (%.0f TFLOP/s)" % (
        name, ms, 2 * m * n * k / ms / 1e9, ms_lib, 2 * m * n * k / ms_lib / 1e9), flush=True)
for (nb, c, h) in ((25, 256, 8), (25, 64, 16), (25, 2048, 1)):
    x = torch.randn(nb, c, h, h, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
    bn = torch.nn.BatchNorm2d(c).cuda().train()
    for _ in range(3):
        y = gnn.fused_batch_norm(x, bn, relu=True)
        y.backward(torch.ones_like(y))
    torch.cuda.synchronize()
print("done")
