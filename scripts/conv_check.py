#!/usr/bin/env python
"""GPU parity sweep of the tcgen05 convolution kernels against fp32 PyTorch (TF32 off) — every distinct layer shape of
ResNet-50 on 32×32 inputs plus awkward geometries. Prints one line per case; exit code 1 on any mismatch."""
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F

from garfield_b200.ops import conv as C

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda")
bad = 0


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-20)).item()


def case(name, n, h, w, cin, cout, k, s, p, groups=8, bias=False, tol=4e-3, time_it=False):
    global bad
    torch.manual_seed(hash(name) % 1000)
    x = torch.randn(n, h, w, cin, device=dev)
    wt = torch.randn(cout, k, k, cin, device=dev) / (k * k * cin) ** 0.5
    b = torch.randn(cout, device=dev) if bias else None
    xr = x.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wr = wt.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = F.conv2d(xr, wr, b, s, p)
    dy = torch.randn_like(yr)
    out = []
    y = C.conv_fprop(x, wt, b, s, p)
    out.append(("fprop", rel(y, yr.permute(0, 2, 3, 1))))
    dyn = dy.permute(0, 2, 3, 1).contiguous()
    dx = C.conv_dgrad(dyn, wt, (h, w), s, p)
    gx, = torch.autograd.grad(yr, xr, dy, retain_graph=True)
    out.append(("dgrad", rel(dx, gx.permute(0, 2, 3, 1))))
    g = groups if n % groups == 0 else 1
    dw = C.conv_wgrad(x, dyn, g, k, s, p)
    ref = []
    for i in range(g):
        xs = xr[i * (n // g):(i + 1) * (n // g)]
        gw, = torch.autograd.grad(F.conv2d(xs, wr, None, s, p), wr, dy[i * (n // g):(i + 1) * (n // g)])
        ref.append(gw.permute(0, 2, 3, 1).reshape(cout, -1))
    out.append(("wgrad", rel(dw, torch.stack(ref))))
    torch.cuda.synchronize()
    msg = "%-28s n=%d %dx%d cin=%d cout=%d k=%d s=%d p=%d : " % (name, n, h, w, cin, cout, k, s, p)
    ok = True
    for kind, e in out:
        flag = "" if e < tol else " <-- FAIL"
        ok &= e < tol
        msg += "%s %.2e%s  " % (kind, e, flag)
    if time_it:
        for kind, fn in (("fprop", lambda: C.conv_fprop(x, wt, b, s, p)), ("dgrad", lambda: C.conv_dgrad(dyn, wt, (h, w), s, p)),
                         ("wgrad", lambda: C.conv_wgrad(x, dyn, g, k, s, p)),
                         ("cudnn_f", lambda: F.conv2d(xcl, wcl, b, s, p)),
                         ("cudnn_d", lambda: torch.ops.aten.convolution_backward(dycl, xcl, wcl, None, [s, s], [p, p], [1, 1], False, [0, 0], 1, [True, False, False])),
                         ("cudnn_w", lambda: torch.ops.aten.convolution_backward(dycl, xcl, wcl, None, [s, s], [p, p], [1, 1], False, [0, 0], 1, [False, True, False]))):
            if kind == "cudnn_f":
                torch.backends.cudnn.allow_tf32 = True
                xcl = xr.detach().contiguous(memory_format=torch.channels_last)
                wcl = wr.detach().contiguous(memory_format=torch.channels_last)
                dycl = dy.contiguous(memory_format=torch.channels_last)
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            # device time of back-to-back launches (a CUDA graph of 20 calls: no host launch overhead in the number)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for _ in range(20):
                    fn()
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            msg += "%s %.1fus " % (kind, e0.elapsed_time(e1) * 10)
        torch.backends.cudnn.allow_tf32 = False
    print(msg, flush=True)
    bad += 0 if ok else 1


T = "--time" in sys.argv
quick = "--quick" in sys.argv
# ResNet-50 (torchvision) on 32×32 inputs, batch 200 = 8 workers × 25
case("l1.conv1 1x1", 200, 8, 8, 64, 64, 1, 1, 0, time_it=T)
case("l1.conv2 3x3", 200, 8, 8, 64, 64, 3, 1, 1, time_it=T)
case("l1.conv3 1x1", 200, 8, 8, 64, 256, 1, 1, 0, time_it=T)
case("l1.b1.conv1 1x1", 200, 8, 8, 256, 64, 1, 1, 0, time_it=T)
case("l2.conv2 3x3 s2", 200, 8, 8, 128, 128, 3, 2, 1, time_it=T)
case("l2.down 1x1 s2", 200, 8, 8, 256, 512, 1, 2, 0, time_it=T)
if not quick:
    case("l2.conv1 1x1", 200, 8, 8, 256, 128, 1, 1, 0, time_it=T)
    case("l2.conv3 1x1", 200, 4, 4, 128, 512, 1, 1, 0, time_it=T)
    case("l2.b1.conv1", 200, 4, 4, 512, 128, 1, 1, 0, time_it=T)
    case("l2.b1.conv2 3x3", 200, 4, 4, 128, 128, 3, 1, 1, time_it=T)
    case("l3.conv1 1x1", 200, 4, 4, 512, 256, 1, 1, 0, time_it=T)
    case("l3.conv2 3x3 s2", 200, 4, 4, 256, 256, 3, 2, 1, time_it=T)
    case("l3.conv3 1x1", 200, 2, 2, 256, 1024, 1, 1, 0, time_it=T)
    case("l3.down 1x1 s2", 200, 4, 4, 512, 1024, 1, 2, 0, time_it=T)
    case("l3.b1.conv1", 200, 2, 2, 1024, 256, 1, 1, 0, time_it=T)
    case("l3.b1.conv2 3x3", 200, 2, 2, 256, 256, 3, 1, 1, time_it=T)
    case("l4.conv1 1x1", 200, 2, 2, 1024, 512, 1, 1, 0, time_it=T)
    case("l4.conv2 3x3 s2", 200, 2, 2, 512, 512, 3, 2, 1, time_it=T)
    case("l4.conv3 1x1", 200, 1, 1, 512, 2048, 1, 1, 0, time_it=T)
    case("l4.down 1x1 s2", 200, 2, 2, 1024, 2048, 1, 2, 0, time_it=T)
    case("l4.b1.conv1", 200, 1, 1, 2048, 512, 1, 1, 0, time_it=T)
    case("l4.b1.conv2 3x3", 200, 1, 1, 512, 512, 3, 1, 1, time_it=T)
    case("fc", 200, 1, 1, 2048, 10, 1, 1, 0, bias=True, time_it=T)
    case("stem 7x7 s2 rgb", 200, 32, 32, 3, 64, 7, 2, 3, time_it=T)
    # awkward geometries: explicit patch path, odd planes, partial tiles, small channel counts
    case("cifarnet conv1 5x5", 56, 32, 32, 3, 6, 5, 1, 0)
    case("cifarnet conv2 5x5", 56, 14, 14, 6, 16, 5, 1, 0)
    case("mnist conv 5x5", 40, 28, 28, 1, 10, 5, 1, 0)
    case("vgg 3x3 32x32", 16, 32, 32, 64, 64, 3, 1, 1)
    case("odd plane 7x7", 24, 7, 7, 32, 48, 3, 1, 1)
    case("wide 3x3 cin96", 16, 16, 16, 96, 160, 3, 1, 1)
    case("1x1 cin 36", 16, 8, 8, 36, 20, 1, 1, 0)
    case("single group", 6, 8, 8, 64, 64, 3, 1, 1, groups=1)
print("FAILED %d" % bad if bad else "ALL OK")
sys.exit(1 if bad else 0)
