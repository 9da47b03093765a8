"""Drive the grouped fused BatchNorm kernels on the flagship shapes (8 workers x batch 25) — timing, or an ncu target."""
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.ops import nn as gnn

dev = torch.device("cuda")
shapes = [(64, 16), (64, 8), (256, 8), (128, 4), (512, 4), (256, 2), (1024, 2), (512, 1), (2048, 1)]
G, B = 8, 25


def run(c, hw, relu=True, res=True, iters=1):
    x = torch.randn(G * B, c, hw, hw, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn_like(x).requires_grad_(True) if res else None
    gamma, beta = torch.ones(c, device=dev, requires_grad=True), torch.zeros(c, device=dev, requires_grad=True)
    rm, rv = torch.zeros(G, c, device=dev), torch.ones(G, c, device=dev)
    gy = torch.randn_like(x)
    got = {}
    sink = lambda dg, db: got.update(dg=dg, db=db)
    times = []
    for phase in ("fwd", "bwd"):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ys = []
        if phase == "fwd":
            for _ in range(3):
                gnn.grouped_batch_norm(x, gamma, beta, rm, rv, G, 1e-5, 0.1, r, relu, sink)
            a.record()
            for _ in range(iters):
                y = gnn.grouped_batch_norm(x, gamma, beta, rm, rv, G, 1e-5, 0.1, r, relu, sink)
            b.record()
        else:
            y.backward(gy, retain_graph=True)
            a.record()
            for _ in range(iters):
                y.backward(gy, retain_graph=True)
            b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b) / iters * 1e3)
    mb = x.numel() * 4 / 1e6
    print(f"C={c:5d} HW={hw:2d}  tensor {mb:6.2f} MB  fwd {times[0]:6.1f} us  bwd {times[1]:6.1f} us (eager incl. launch overhead)", flush=True)


if __name__ == "__main__":
    for c, hw in shapes:
        run(c, hw, iters=1 if len(sys.argv) > 1 else 20)
