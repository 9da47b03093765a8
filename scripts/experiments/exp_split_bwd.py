#!/usr/bin/env python
"""Experiment: take conv weight-gradient kernels off the backward critical path (side stream) and
see how much one worker's CUDA-graphed ResNet-50 step shrinks."""
import sys, types
import torch, torchvision

sys.path.insert(0, ".")
torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")
aten = torch.ops.aten


class SplitConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, stride, padding, dilation, groups, side):
        ctx.save_for_backward(x, w)
        ctx.cfg = (stride, padding, dilation, groups, side, None if b is None else list(b.shape))
        return torch.nn.functional.conv2d(x, w, b, stride, padding, dilation, groups)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, padding, dilation, groups, side, bias_sizes = ctx.cfg
        main = torch.cuda.current_stream()
        gx = gw = gb = None
        if ctx.needs_input_grad[1]:
            side.wait_stream(main)
            gy.record_stream(side); x.record_stream(side)
            with torch.cuda.stream(side):
                _, gw, gb = aten.convolution_backward(gy, x, w, bias_sizes, stride, padding, dilation, False, [0, 0], groups,
                                                      [False, True, bias_sizes is not None])
        if ctx.needs_input_grad[0]:
            gx = aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0], groups, [True, False, False])[0]
        return gx, gw, gb, None, None, None, None, None


def patch(module, side):
    def forward(self, x):
        return SplitConv.apply(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups, side)
    for m in module.modules():
        if isinstance(m, torch.nn.Conv2d) and m.padding_mode == "zeros":
            m.forward = types.MethodType(forward, m)


def bench(split):
    m = torchvision.models.resnet50(num_classes=10).to(dev).to(memory_format=torch.channels_last)
    side = torch.cuda.Stream()
    if split:
        patch(m, side)
    x = torch.randn(25, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (25,), device=dev)
    lossf = torch.nn.CrossEntropyLoss()
    def run():
        for p in m.parameters(): p.grad = None
        lossf(m(x), y).backward()
        if split:
            torch.cuda.current_stream().wait_stream(side)
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3): run()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    ref = torch.cat([p.grad.reshape(-1) for p in m.parameters()]).clone()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): run()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(30): g.replay()
    b.record(); torch.cuda.synchronize()
    print("split=%s: %.3f ms/step" % (split, a.elapsed_time(b) / 30), flush=True)
    return ref

torch.manual_seed(0); r0 = bench(False)
torch.manual_seed(0); r1 = bench(True)
print("max |diff| %.3e of max |g| %.3e" % ((r0 - r1).abs().max().item(), r0.abs().max().item()))
