#!/usr/bin/env python
"""Device time of ONE rank's graphed gradient pass (ResNet-50, G co-located workers × B images) — `--groups 1 --batch 25` is what
each GPU of the 8-GPU run executes, `--groups 8` the single-GPU run."""
import argparse
import copy
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--groups", type=int, default=1)
ap.add_argument("--batch", type=int, default=25)
ap.add_argument("--model", default="resnet50")
args = ap.parse_args()
from garfield_b200.garfieldpp import tools
from garfield_b200.models.fused_resnet import fuse_resnet_blocks
from garfield_b200.parallel.arena import FlatModel
from garfield_b200.parallel.engine import _skip_bn_counters
from garfield_b200.parallel.grouped import GroupedWorkers

dev = torch.device("cuda")
torch.manual_seed(0)
module = tools.select_model(args.model, dev, "cifar10")
_skip_bn_counters(module)
fuse_resnet_blocks(module)
module = module.to(memory_format=torch.channels_last)
flat = FlatModel(module, dev, channels_last=True)
G, B = args.groups, args.batch
rows = torch.zeros(G, (flat.numel + 63) // 64 * 64, device=dev)
group = GroupedWorkers(flat, [rows[g][:flat.numel] for g in range(G)], torch.nn.CrossEntropyLoss())
x = torch.randn(G * B, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (G * B,), device=dev)
losses = torch.zeros(G, device=dev)
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(3):
        group.forward_backward(x, y, losses)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    group.forward_backward(x, y, losses)
for _ in range(5):
    graph.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    graph.replay()
e1.record()
torch.cuda.synchronize()
print("pass G=%d B=%d: %.4f ms" % (G, B, e0.elapsed_time(e1) / 50))
