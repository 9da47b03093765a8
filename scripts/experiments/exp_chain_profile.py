#!/usr/bin/env python
"""Where does one worker's ResNet-50 forward+backward chain (batch 25, 32x32, NHWC) spend its time?"""
import torch, torchvision
from torch.profiler import profile, ProfilerActivity

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")
import sys
sys.path.insert(0, ".")
m = torchvision.models.resnet50(num_classes=10).to(dev).to(memory_format=torch.channels_last)
if "--fused" in sys.argv:
    from garfield_b200.models.fused_resnet import fuse_resnet_blocks
    print("patched blocks:", fuse_resnet_blocks(m))
x = torch.randn(25, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (25,), device=dev)
lossf = torch.nn.CrossEntropyLoss()
def run():
    for p in m.parameters():
        p.grad = None
    lossf(m(x), y).backward()
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): run()
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g): run()
for _ in range(3): g.replay()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20): g.replay()
b.record(); torch.cuda.synchronize()
wall = a.elapsed_time(b) / 20
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    g.replay(); torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages()]
rows.sort(key=lambda r: -r[2])
total = sum(r[2] for r in rows)
print("graph replay wall %.3f ms; sum of kernel times %.3f ms; kernels %d" % (wall, total / 1e3, sum(r[1] for r in rows)))
for k, c, t in rows[:25]:
    print("%8.1f us  x%-4d  avg %6.2f us  %s" % (t, c, t / c, k[:110]))
