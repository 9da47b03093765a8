"""One fused-workers (G = 8, B = 25) ResNet-50 gradient pass inside a cudaProfilerStart/Stop window, for an
`ncu --profile-from-start off --metrics gpu__time_duration.sum` kernel list; without ncu prints the CUDA-graph time."""
import sys

import torch

sys.path.insert(0, ".")
from garfield_b200.parallel.engine import SSMWEngine

torch.backends.cudnn.benchmark = True
eng = SSMWEngine("resnet50", "synthetic-cifar10", "cross-entropy", 8, 1, "krum", 25, "sgd", dict(lr=0.2, momentum=0.9, weight_decay=5e-4),
                 use_graphs=False, fuse_workers=True)
eng.load_batches(0)
for _ in range(3):
    eng._run_workers()
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng._run_workers()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
if len(sys.argv) > 1 and sys.argv[1] == "time":
    eng.use_graphs = True
    eng._capture()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        eng._graph.replay()
    a.record()
    for _ in range(50):
        eng._graph.replay()
    b.record()
    torch.cuda.synchronize()
    print("fused gradient pass (graph): %.3f ms" % (a.elapsed_time(b) / 50))
print("done")
