"""Gradient pass of ONE hosted worker (the 8-GPU situation): per-worker chain vs the fused path with G = 1."""
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.parallel.engine import SSMWEngine

torch.backends.cudnn.benchmark = True
for g in (1, 2, 4):
    for fuse in (False, True):
        eng = SSMWEngine("resnet50", "synthetic-cifar10", "cross-entropy", g, 0, "average", 25, "sgd", dict(lr=0.2, momentum=0.9, weight_decay=5e-4),
                         use_graphs=True, fuse_workers=fuse)
        eng.load_batches(0)
        eng._capture()
        for _ in range(5):
            eng._graph.replay()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50):
            eng._graph.replay()
        b.record()
        torch.cuda.synchronize()
        print("workers on this GPU %d  fused %-5s  gradient pass %.3f ms" % (g, fuse, a.elapsed_time(b) / 50), flush=True)
        eng.close()
        del eng
        torch.cuda.empty_cache()
