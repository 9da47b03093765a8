"""Where does the ResNet-50 step go? Time every distinct convolution of the flagship model (forward, data gradient,
weight gradient) through the library path at the per-worker batch (25) and at the all-workers batch (200), plus the
fused BN kernels, and print a table weighted by how often each shape occurs."""
import collections
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.garfieldpp import tools

dev = torch.device("cuda")
aten = torch.ops.aten
torch.backends.cudnn.benchmark = True


def timeit(fn, steps=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps * 1e3  # µs


def main():
    m = tools.select_model("resnet50", dev, "cifar10").to(memory_format=torch.channels_last)
    shapes = collections.Counter()

    def hook(mod, inp, out):
        x = inp[0]
        shapes[(mod.in_channels, mod.out_channels, mod.kernel_size[0], mod.stride[0], mod.padding[0], x.shape[2])] += 1

    hs = [c.register_forward_hook(hook) for c in m.modules() if isinstance(c, torch.nn.Conv2d)]
    with torch.no_grad():
        m(torch.randn(2, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last))
    for h in hs:
        h.remove()
    tot = collections.defaultdict(float)
    print(f"{'cin':>5} {'cout':>5} k s {'H':>3} cnt | batch 25: fwd dgrad wgrad (us) | batch 200: fwd dgrad wgrad (us) | GFLOP(200)")
    for (cin, cout, k, s, p, H), cnt in sorted(shapes.items(), key=lambda kv: -kv[1]):
        row = []
        for batch in (25, 200):
            x = torch.randn(batch, cin, H, H, device=dev).contiguous(memory_format=torch.channels_last)
            w = torch.randn(cout, cin, k, k, device=dev).contiguous(memory_format=torch.channels_last)
            y = torch.nn.functional.conv2d(x, w, None, s, p)
            gy = torch.randn_like(y)
            f = timeit(lambda: torch.nn.functional.conv2d(x, w, None, s, p))
            d = timeit(lambda: aten.convolution_backward(gy, x, w, None, [s, s], [p, p], [1, 1], False, [0, 0], 1, [True, False, False]))
            g = timeit(lambda: aten.convolution_backward(gy, x, w, None, [s, s], [p, p], [1, 1], False, [0, 0], 1, [False, True, False]))
            row += [f, d, g]
            tot[(batch, "fwd")] += f * cnt
            tot[(batch, "dgrad")] += d * cnt
            tot[(batch, "wgrad")] += g * cnt
        Ho = y.shape[2]
        gflop = 2.0 * 200 * Ho * Ho * cout * cin * k * k / 1e9
        print(f"{cin:5d} {cout:5d} {k} {s} {H:3d} {cnt:3d} | {row[0]:7.1f} {row[1]:7.1f} {row[2]:7.1f} | {row[3]:7.1f} {row[4]:7.1f} {row[5]:7.1f} | {gflop:6.3f}", flush=True)
    for batch in (25, 200):
        print(f"batch {batch}: sum over layers fwd {tot[(batch, 'fwd')] / 1e3:.3f} ms, dgrad {tot[(batch, 'dgrad')] / 1e3:.3f} ms, "
              f"wgrad {tot[(batch, 'wgrad')] / 1e3:.3f} ms")


if __name__ == "__main__":
    main()
