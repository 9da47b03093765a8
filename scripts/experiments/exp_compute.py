#!/usr/bin/env python
"""Experiment: CUDA-graphed fwd+bwd time of torchvision ResNet-50 (10 classes, 32x32, batch 25)
under different layouts / precisions, 1 and 8 concurrent replicas."""
import sys, time
import torch, torchvision

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")

def bench(tag, channels_last=False, autocast=None, replicas=1, batch=25, tf32_matmul=False):
    torch.backends.cuda.matmul.allow_tf32 = tf32_matmul
    models = [torchvision.models.resnet50(num_classes=10).to(dev) for _ in range(replicas)]
    if channels_last:
        models = [m.to(memory_format=torch.channels_last) for m in models]
    xs = [torch.randn(batch, 3, 32, 32, device=dev) for _ in range(replicas)]
    if channels_last:
        xs = [x.contiguous(memory_format=torch.channels_last) for x in xs]
    ys = [torch.randint(0, 10, (batch,), device=dev) for _ in range(replicas)]
    streams = [torch.cuda.Stream() for _ in range(replicas)]
    lossf = torch.nn.CrossEntropyLoss()
    for m in models:
        for p in m.parameters():
            p.grad = torch.zeros_like(p)
    def run():
        main = torch.cuda.current_stream()
        for m, x, y, s in zip(models, xs, ys, streams):
            s.wait_stream(main)
            with torch.cuda.stream(s):
                if autocast:
                    with torch.autocast("cuda", dtype=autocast):
                        loss = lossf(m(x), y)
                else:
                    loss = lossf(m(x), y)
                loss.backward()
        for s in streams:
            main.wait_stream(s)
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): run()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): run()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): g.replay()
    b.record(); torch.cuda.synchronize()
    print("%-40s replicas=%d  %.3f ms/step  (%.0f img/s)" % (tag, replicas, a.elapsed_time(b) / 20, replicas * batch / (a.elapsed_time(b) / 20 / 1e3)), flush=True)

for reps in (1, 8):
    bench("fp32 NCHW (reference numerics)", replicas=reps)
    bench("fp32 channels_last", channels_last=True, replicas=reps)
    bench("fp32 NCHW + tf32 matmul", tf32_matmul=True, replicas=reps)
    bench("bf16 autocast channels_last (info only)", channels_last=True, autocast=torch.bfloat16, replicas=reps)
