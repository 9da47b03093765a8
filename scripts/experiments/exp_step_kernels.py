"""One fused-BN ResNet-50 forward+backward at a given batch inside a cudaProfilerStart/Stop window (for an
`ncu --profile-from-start off --metrics gpu__time_duration.sum` kernel list)."""
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.garfieldpp import tools
from garfield_b200.models.fused_resnet import fuse_resnet_blocks
from garfield_b200.parallel.engine import _skip_bn_counters

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True
m = tools.select_model("resnet50", dev, "cifar10")
_skip_bn_counters(m)
fuse_resnet_blocks(m)
m = m.to(memory_format=torch.channels_last)
x = torch.randn(batch, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (batch,), device=dev)
lossf = torch.nn.CrossEntropyLoss()


def step():
    for p in m.parameters():
        p.grad = None
    lossf(m(x), y).backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
