"""Feasibility timing: ONE forward/backward over the batches of all co-located workers (shared weights, batched
forward + data gradients, per-worker weight gradients) versus one chain per worker. Timing only (BN statistics over
the whole batch here); the real grouped path keeps per-worker statistics."""
import sys
import time

import torch

sys.path.insert(0, ".")
from garfield_b200.garfieldpp import tools
from garfield_b200.models.fused_resnet import fuse_resnet_blocks
from garfield_b200.parallel.engine import _skip_bn_counters

G, B = 8, 25
dev = torch.device("cuda")
aten = torch.ops.aten


class GroupConv(torch.autograd.Function):
    sides = None
    sink = []

    @staticmethod
    def forward(ctx, x, w, stride, padding):
        ctx.save_for_backward(x, w)
        ctx.cfg = (stride, padding)
        return torch.nn.functional.conv2d(x, w, None, stride, padding)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, padding = ctx.cfg
        sides = GroupConv.sides
        cur = torch.cuda.current_stream()
        if sides:
            for s in sides:
                s.wait_stream(cur)
        for g in range(G):
            xs, gs = x[g * B:(g + 1) * B], gy[g * B:(g + 1) * B]
            if sides:
                with torch.cuda.stream(sides[g % len(sides)]):
                    gw = aten.convolution_backward(gs, xs, w, None, stride, padding, [1, 1], False, [0, 0], 1, [False, True, False])[1]
            else:
                gw = aten.convolution_backward(gs, xs, w, None, stride, padding, [1, 1], False, [0, 0], 1, [False, True, False])[1]
            GroupConv.sink.append(gw)
        gx = None
        if ctx.needs_input_grad[0]:
            gx = aten.convolution_backward(gy, x, w, None, stride, padding, [1, 1], False, [0, 0], 1, [True, False, False])[0]
        return gx, None, None, None


def patch(module):
    import types

    def forward(self, x):
        return GroupConv.apply(x, self.weight, self.stride, self.padding)

    for m in module.modules():
        if isinstance(m, torch.nn.Conv2d):
            m.forward = types.MethodType(forward, m)


def build(grouped):
    torch.manual_seed(0)
    m = tools.select_model("resnet50", dev, "cifar10")
    _skip_bn_counters(m)
    fuse_resnet_blocks(m)
    m = m.to(memory_format=torch.channels_last)
    if grouped:
        patch(m)
    return m


def timeit(fn, steps=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def run(label, grouped, sides, batch):
    m = build(grouped)
    GroupConv.sides = [torch.cuda.Stream() for _ in range(sides)] if sides else None
    x = torch.randn(batch, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (batch,), device=dev)
    lossf = torch.nn.CrossEntropyLoss()

    def step():
        GroupConv.sink.clear()
        for p in m.parameters():
            p.grad = None
        loss = lossf(m(x), y)
        loss.backward()
        if GroupConv.sides:
            for s in GroupConv.sides:
                torch.cuda.current_stream().wait_stream(s)

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            step()
    torch.cuda.synchronize()
    eager = timeit(step, 10)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        step()
    t = timeit(g.replay)
    print(f"{label:44s} eager {eager:7.3f} ms   graph {t:7.3f} ms   -> {G * B / t * 1e3:9.0f} img/s", flush=True)


if __name__ == "__main__":
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    run("one chain, batch 25 (1 worker)", False, 0, 25)
    run("batch 200, plain autograd (wgrad over 200)", False, 0, 200)
    run("batch 200, per-group wgrad, same stream", True, 0, 200)
    run("batch 200, per-group wgrad, 4 side streams", True, 4, 200)
    run("batch 200, per-group wgrad, 8 side streams", True, 8, 200)
