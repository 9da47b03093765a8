#!/usr/bin/env python
"""One small call of every hand-written kernel family, for `compute-sanitizer --tool memcheck|racecheck`."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import aggregators  # noqa: E402
from garfield_b200.ops import attacks, gar, nn as gnn  # noqa: E402
from garfield_b200.ops.optim import FlatOptimizer  # noqa: E402

torch.manual_seed(0)
for n, f, d in ((8, 1, 5000), (11, 2, 3001), (23, 5, 4100)):
    g = [torch.randn(d, device="cuda") for _ in range(n)]
    for name in ("average", "median", "native-median", "trimmed_mean", "averaged_median", "krum", "native-krum", "bulyan",
                 "native-bulyan", "brute", "aksel", "condense"):
        rule = aggregators.gars[name]
        if rule.check(gradients=g, f=f) is None and not (name == "brute" and n > 11):
            rule(gradients=g, f=f)
    gar.pairwise_sqdist(g, engine="tc")
    gar.pairwise_sqdist([t[1:] for t in g], engine="simt")       # unaligned rows → direct-load path
flat = torch.randn(10001, device="cuda")
for name in ("sgd", "adam", "adamw", "rmsprop", "adagrad"):
    FlatOptimizer(flat.clone(), name, lr=0.01).step(torch.randn(10001, device="cuda"))
opt = FlatOptimizer(flat.clone(), "sgd", lr=0.1, momentum=0.9)
rows = [torch.randn(10001, device="cuda") for _ in range(8)]
gar.krum(rows, 1, 5, 5, True, update=opt.sgd_fusion(mirrors=[torch.zeros(10001, device="cuda")]), want_out=False)
for fn in (attacks.random_like, attacks.reverse, attacks.drop):
    fn(flat)
attacks.lie(rows[:3]); attacks.empire(rows[:3])
a, b = torch.randn(130, 100, device="cuda"), torch.randn(70, 100, device="cuda")
gnn.gemm_nt(a, b, bias=torch.randn(70, device="cuda"), relu=True)
gnn.gemm_nt(a.bfloat16()[:, :96].contiguous(), b.bfloat16()[:, :96].contiguous())
x = torch.randn(4, 3, 12, 12, device="cuda", requires_grad=True)
w = torch.randn(5, 3, 3, 3, device="cuda", requires_grad=True)
y = gnn.max_pool2d(gnn.conv2d(x, w, torch.zeros(5, device="cuda"), 1, 1, relu=True), 2)
gnn.cross_entropy(gnn.global_avg_pool(y), torch.randint(0, 5, (4,), device="cuda")).backward()
xb = torch.randn(6, 8, 5, 5, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
bn = torch.nn.BatchNorm2d(8).cuda().train()
gnn.fused_batch_norm(xb, bn, residual=torch.randn_like(xb), relu=True).sum().backward()
gnn.batch_norm_train(torch.randn(6, 8, 5, 5, device="cuda", requires_grad=True), bn.weight, bn.bias).sum().backward()
gnn.dropout(torch.randn(1000, device="cuda", requires_grad=True), 0.3, True, seed=1).sum().backward()
gnn.binary_cross_entropy(torch.rand(16, 1, device="cuda").clamp(0.01, 0.99).requires_grad_(True), torch.ones(16, 1, device="cuda")).backward()
torch.cuda.synchronize()
print("sanitize driver finished")
