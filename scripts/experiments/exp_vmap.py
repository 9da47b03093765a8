#!/usr/bin/env python
"""Experiment: per-worker gradients of ResNet-50 for 8 co-located workers in ONE batched pass
(torch.func vmap(grad) → convs fold the worker dim into the batch / groups) vs 8 streams."""
import copy
import sys
import torch, torchvision
from torch.func import functional_call, grad, vmap

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")
W, B = 8, 25
CL = "--nchw" not in sys.argv
model = torchvision.models.resnet50(num_classes=10).to(dev)
if CL:
    model = model.to(memory_format=torch.channels_last)
for m in model.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.track_running_stats = False
        m.running_mean = None; m.running_var = None; m.num_batches_tracked = None
model.train()
params = {k: v.detach() for k, v in model.named_parameters()}
buffers = {k: v for k, v in model.named_buffers()}
X = torch.randn(W * B, 3, 32, 32, device=dev)
if CL:
    X = X.contiguous(memory_format=torch.channels_last)
X = X.view(W, B, 3, 32, 32)
Y = torch.randint(0, 10, (W, B), device=dev)
lossf = torch.nn.functional.cross_entropy

def one(params, x, y):
    out = functional_call(model, (params, buffers), (x,))
    return lossf(out, y)

per_worker_grads = vmap(grad(one), in_dims=(None, 0, 0))
d = sum(p.numel() for p in params.values())
rows = torch.zeros(W, d, device=dev)

def run():
    g = per_worker_grads(params, X, Y)
    off = 0
    for k, p in params.items():
        n = p.numel()
        rows[:, off:off + n].copy_(g[k].reshape(W, n))
        off += n

try:
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): run()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): run()
    b.record(); torch.cuda.synchronize()
    print("vmap(grad) eager: %.3f ms/step" % (a.elapsed_time(b) / 10), flush=True)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): run()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    a.record()
    for _ in range(20): g.replay()
    b.record(); torch.cuda.synchronize()
    print("vmap(grad) CUDA graph: %.3f ms/step (%.0f img/s)" % (a.elapsed_time(b) / 20, W * B / (a.elapsed_time(b) / 20 / 1e3)), flush=True)
    # check against a plain per-worker gradient
    ref = copy.deepcopy(model)
    ref.zero_grad()
    lossf(ref(X[3]), Y[3]).backward()
    flat = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
    print("max |vmap - plain| worker 3: %.3e (|g|max %.3e)" % ((rows[3] - flat).abs().max().item(), flat.abs().max().item()))
except Exception as err:
    import traceback; traceback.print_exc()
    print("FAILED:", err)
