#!/usr/bin/env python
"""tcgen05 GEMM (`gemm_nt`) against cuBLAS over a few shapes and both operand types: CUDA events, 20 timed launches after 3 warm-ups.
`GARFIELD_GEMM_PAIR=0` times the single-CTA kernel on the same shapes."""
import json
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch

from garfield_b200.ops import nn as gnn

torch.backends.cuda.matmul.allow_tf32 = True
torch.manual_seed(0)
shapes = [(4096, 4096, 4096), (8192, 8192, 4096), (8192, 4096, 1024), (2048, 8192, 8192), (16384, 4096, 512)]
out = []


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / 20


for dtype, name in ((torch.bfloat16, "bf16"), (torch.float32, "tf32")):
    for m, n, k in shapes:
        a = torch.randn(m, k, device="cuda").to(dtype)
        b = torch.randn(n, k, device="cuda").to(dtype)
        got = gnn.gemm_nt(a, b)
        ref = (a @ b.t()).float()
        err = ((got - ref).abs().max() / ref.abs().max()).item()
        ms = timed(lambda: gnn.gemm_nt(a, b))
        lib = timed(lambda: a @ b.t())
        rec = {"dtype": name, "m": m, "n": n, "k": k, "ours_ms": round(ms, 4), "cublas_ms": round(lib, 4), "ours_tflops": round(2 * m * n * k / ms / 1e9),
               "cublas_tflops": round(2 * m * n * k / lib / 1e9), "ratio": round(lib / ms, 3), "max_rel_diff_vs_cublas": err,
               "pair": os.environ.get("GARFIELD_GEMM_PAIR", "1")}
        print(json.dumps(rec), flush=True)
        out.append(rec)
if len(sys.argv) > 1:
    pathlib.Path(sys.argv[1]).write_text(json.dumps(out, indent=1))
