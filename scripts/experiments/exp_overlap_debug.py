import os, sys, traceback
os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
sys.path.insert(0, ".")
import torch
from garfield_b200.parallel.engine import SSMWEngine
for model, batch in (("cnn", 8), ("resnet50", 25)):
    eng = SSMWEngine(model, "synthetic-cifar10", "cross-entropy", 8, 1, "krum", batch, "sgd", dict(lr=0.02, momentum=0.9), seed=3, overlap_distances=True)
    print(model, "tail_at", eng._tail_at, "d", eng.d, flush=True)
    orig = eng._step_body
    try:
        for i in range(2):
            eng.step(i)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            eng._step_body()
        print(model, "capture ok", flush=True)
    except Exception as e:
        print(model, "capture FAILED:", type(e).__name__, str(e)[:600], flush=True)
        traceback.print_exc(limit=6)
    break_ = False
