"""Worker fusion (parallel/grouped.py): one pass over the concatenated batches must give every worker exactly the
gradient its own replica would compute (per-worker BatchNorm statistics, per-worker weight gradients, per-worker loss)."""
import copy

import pytest
import torch

from garfield_b200.parallel.arena import FlatModel
from garfield_b200.parallel.grouped import GroupedWorkers


class _Block(torch.nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = torch.nn.Conv2d(cin, cout, 1, bias=False)
        self.bn1 = torch.nn.BatchNorm2d(cout)
        self.conv2 = torch.nn.Conv2d(cout, cout, 3, stride=stride, padding=1, bias=True)
        self.bn2 = torch.nn.BatchNorm2d(cout)
        self.down = torch.nn.Conv2d(cin, cout, 1, stride=stride, bias=False)

    def forward(self, x):
        out = torch.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return torch.relu(out + self.down(x))


class _Net(torch.nn.Module):
    def __init__(self, classes=5):
        super().__init__()
        self.stem = torch.nn.Conv2d(3, 8, 3, padding=1)
        self.b1 = _Block(8, 16, 1)
        self.b2 = _Block(16, 32, 2)
        self.pool = torch.nn.AdaptiveAvgPool2d(1)
        self.fc = torch.nn.Linear(32, classes)

    def forward(self, x):
        x = self.b2(self.b1(torch.relu(self.stem(x))))
        return self.fc(torch.flatten(self.pool(x), 1))


def _reference(module, xs, ys, loss_fn):
    grads, losses, stats = [], [], []
    for x, y in zip(xs, ys):
        m = copy.deepcopy(module)
        m.train()
        loss = loss_fn(m(x), y)
        loss.backward()
        grads.append(torch.cat([p.grad.reshape(-1) for p in m.parameters()]))
        losses.append(loss.detach())
        stats.append({k: v.clone() for k, v in m.named_buffers() if "running" in k})
    return grads, losses, stats


def _run_grouped(module, xs, ys, loss_fn, device, channels_last):
    G = len(xs)
    flat_model = FlatModel(copy.deepcopy(module), device, channels_last=channels_last)
    d = flat_model.numel
    matrix = torch.zeros(G, d, device=device)
    group = GroupedWorkers(flat_model, [matrix[g] for g in range(G)], loss_fn, side_streams=2)
    data = torch.cat(xs).to(device)
    if channels_last:
        data = data.contiguous(memory_format=torch.channels_last)
    losses = torch.zeros(G, device=device)
    for _ in range(2):  # twice: the sink must not keep state between steps
        for name, (rm, rv) in group.running_tables.items():
            ref = dict(module.named_modules())[name]
            rm.copy_(ref.running_mean.expand_as(rm)); rv.copy_(ref.running_var.expand_as(rv))
        group.forward_backward(data, torch.cat(ys).to(device), losses)
    return flat_model, matrix, losses, group


def _logical(flat_model, row):
    """Row of the gradient matrix → the reference's flat order (undo the OHWI storage of 4-D parameters)."""
    return torch.cat([flat_model._view(row, off, n, shape).reshape(-1) for off, n, shape in flat_model.slices])


def test_grouped_workers_match_separate_replicas_cpu():
    torch.manual_seed(0)
    module = _Net()
    G, B = 3, 4
    xs = [torch.randn(B, 3, 8, 8) for _ in range(G)]
    ys = [torch.randint(0, 5, (B,)) for _ in range(G)]
    loss_fn = torch.nn.CrossEntropyLoss()
    grads, losses, stats = _reference(module, xs, ys, loss_fn)
    flat_model, matrix, glosses, group = _run_grouped(module, xs, ys, loss_fn, torch.device("cpu"), channels_last=False)
    for g in range(G):
        assert torch.allclose(_logical(flat_model, matrix[g]), grads[g], rtol=1e-4, atol=1e-5), g
        assert torch.allclose(glosses[g], losses[g], rtol=1e-5)
        for name, (rm, rv) in group.running_tables.items():
            assert torch.allclose(rm[g], stats[g][name + ".running_mean"], rtol=1e-4, atol=1e-6)
            assert torch.allclose(rv[g], stats[g][name + ".running_var"], rtol=1e-4, atol=1e-6)


def test_grouped_workers_channels_last_storage_cpu():
    torch.manual_seed(1)
    module = _Net()
    xs = [torch.randn(2, 3, 8, 8) for _ in range(2)]
    ys = [torch.randint(0, 5, (2,)) for _ in range(2)]
    loss_fn = torch.nn.CrossEntropyLoss()
    grads, _, _ = _reference(module, xs, ys, loss_fn)
    flat_model, matrix, _, _ = _run_grouped(module, xs, ys, loss_fn, torch.device("cpu"), channels_last=True)
    for g in range(2):
        assert torch.allclose(_logical(flat_model, matrix[g]), grads[g], rtol=1e-4, atol=1e-5)


def test_unsupported_models_are_reported():
    loss_fn = torch.nn.CrossEntropyLoss()
    assert GroupedWorkers.unsupported_reason(_Net(), loss_fn) is None
    odd = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 1), torch.nn.GroupNorm(2, 4))
    assert "GroupNorm" in GroupedWorkers.unsupported_reason(odd, loss_fn)
    assert "loss" in GroupedWorkers.unsupported_reason(_Net(), torch.nn.MSELoss())


@pytest.mark.gpu
@pytest.mark.parametrize("own_kernels,pointwise_gemm", [(True, False), (False, False), (False, True)])
def test_grouped_workers_match_separate_replicas_gpu(own_kernels, pointwise_gemm):
    """own_kernels: convolutions / Linear / pooling / loss on the tcgen05 implicit-GEMM kernels (TF32 operands → looser
    tolerance); otherwise the library kernels in exact fp32, optionally with 1×1 convolutions on the tcgen05 GEMM."""
    from garfield_b200.parallel import grouped
    previous = grouped.POINTWISE_GEMM[0], grouped.OWN_KERNELS[0]
    grouped.POINTWISE_GEMM[0], grouped.OWN_KERNELS[0] = pointwise_gemm, own_kernels
    try:
        tf32 = own_kernels or pointwise_gemm
        _check_grouped_gpu(8e-2 if tf32 else 2e-4, tf32)   # TF32 everywhere vs exact fp32
    finally:
        grouped.POINTWISE_GEMM[0], grouped.OWN_KERNELS[0] = previous


def _check_grouped_gpu(tolerance, allow_tf32):
    torch.manual_seed(0)
    dev = torch.device("cuda")
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = allow_tf32     # exact comparison unless the TF32 GEMM path is under test
    try:
        module = _Net().to(dev)
        G, B = 4, 6
        xs = [torch.randn(B, 3, 16, 16, device=dev) for _ in range(G)]
        ys = [torch.randint(0, 5, (B,), device=dev) for _ in range(G)]
        loss_fn = torch.nn.CrossEntropyLoss()
        grads, losses, stats = _reference(module, xs, ys, loss_fn)
        flat_model, matrix, glosses, group = _run_grouped(module, xs, ys, loss_fn, dev, channels_last=True)
        torch.cuda.synchronize()
        for g in range(G):
            got, want = _logical(flat_model, matrix[g]), grads[g]
            assert (got - want).norm() <= tolerance * want.norm(), (g, float((got - want).norm() / want.norm()))
            loose = allow_tf32
            assert torch.allclose(glosses[g], losses[g], rtol=2e-2 if loose else 1e-4)
            for name, (rm, rv) in group.running_tables.items():
                assert torch.allclose(rm[g], stats[g][name + ".running_mean"], rtol=2e-2 if loose else 1e-3, atol=1e-3 if loose else 1e-5)
                assert torch.allclose(rv[g], stats[g][name + ".running_var"], rtol=2e-2 if loose else 1e-3, atol=1e-3 if loose else 1e-5)
    finally:
        torch.backends.cudnn.allow_tf32 = prev


@pytest.mark.gpu
@pytest.mark.parametrize("groups,batch,channels,hw", [(1, 25, 64, 16), (3, 4, 64, 20), (8, 25, 256, 8), (2, 40, 8, 32), (1, 3, 2048, 1),
                                                      (4, 2, 12, 3), (1, 200, 64, 16)])
def test_grouped_batchnorm_kernel_matches_torch(groups, batch, channels, hw):
    """Cluster-split / register-resident / re-read variants of the fused NHWC BatchNorm against the per-group torch formula."""
    from garfield_b200.ops import nn as gnn
    torch.manual_seed(groups * 1000 + channels)
    dev = torch.device("cuda")
    n = groups * batch
    x = (torch.randn(n, channels, hw, hw, device=dev) * 2 + 3).contiguous(memory_format=torch.channels_last)
    res = torch.randn_like(x).contiguous(memory_format=torch.channels_last)
    gamma, beta = torch.rand(channels, device=dev) + 0.5, torch.randn(channels, device=dev)
    gy = torch.randn_like(x)
    for relu, use_res in ((True, True), (False, False), (True, False)):
        outs = []
        for device in (dev, torch.device("cpu")):
            xi = x.to(device).detach().requires_grad_(True)
            ri = res.to(device).detach().requires_grad_(True) if use_res else None
            g, b = gamma.to(device).clone().requires_grad_(True), beta.to(device).clone().requires_grad_(True)
            rm, rv = torch.zeros(groups, channels, device=device), torch.ones(groups, channels, device=device)
            got = {}
            y = gnn.grouped_batch_norm(xi, g, b, rm, rv, groups, 1e-5, 0.1, ri, relu, sink=lambda dg, db: got.update(dg=dg, db=db))
            y.backward(gy.to(device))
            outs.append((y.detach().cpu(), xi.grad.cpu(), None if ri is None else ri.grad.cpu(), got["dg"].reshape(groups, channels).cpu(),
                         got["db"].reshape(groups, channels).cpu(), rm.cpu(), rv.cpu()))
        for a, b in zip(*outs):
            if a is not None:   # a ReLU output within rounding of zero may flip its mask: tolerate isolated elements
                bad = (a - b).abs() > 2e-3 + 2e-3 * b.abs()
                assert int(bad.sum()) <= max(3, 2e-6 * bad.numel()), (int(bad.sum()), float((a - b).abs().max()))   # (and the sums they feed)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,k,stride,pad", [((6, 8, 9, 7), 3, 1, 1), ((4, 64, 8, 8), 3, 2, 1), ((3, 12, 5, 5), 5, 1, 2), ((2, 4, 4, 4), 1, 2, 0)])
def test_im2col_kernel_matches_unfold(shape, k, stride, pad):
    from garfield_b200.ops import nn as gnn
    x = torch.randn(*shape, device="cuda").contiguous(memory_format=torch.channels_last)
    got = gnn.im2col_nhwc(x, (k, k), (stride, stride), (pad, pad))
    want = gnn.im2col_nhwc(x.cpu().contiguous(), (k, k), (stride, stride), (pad, pad))
    assert torch.equal(got.cpu(), want)


def test_scatter_sink_slices_every_gradient_to_its_owner_cpu():
    """ScatterSink: worker g's gradient lands slice by slice in the arena of the rank owning each coordinate range
    (in-place GEMM outputs for parameters inside one range, copies for the ones straddling a boundary)."""
    from garfield_b200.parallel.grouped import ScatterSink
    torch.manual_seed(3)
    module = _Net()
    G, B, world = 3, 4, 3
    xs = [torch.randn(B, 3, 8, 8) for _ in range(G)]
    ys = [torch.randint(0, 5, (B,)) for _ in range(G)]
    loss_fn = torch.nn.CrossEntropyLoss()
    grads, _, _ = _reference(module, xs, ys, loss_fn)
    flat_model = FlatModel(copy.deepcopy(module), torch.device("cpu"))
    d = flat_model.numel
    per = ((d + 63) // 64 + world - 1) // world * 64
    bounds = [(min(d, r * per), min(d, r * per + per)) for r in range(world)]
    first_row, rows_total = 2, 2 + G + 1                      # this "rank" owns arena rows 2 .. 2+G-1
    arenas = [torch.full((rows_total * per,), float("nan")) for _ in range(world)]
    sink = lambda params, slices, view_of: ScatterSink(params, slices, view_of, G, arenas, first_row, per, bounds, rank=1)
    group = GroupedWorkers(flat_model, [None] * G, loss_fn, side_streams=0, sink=sink)
    straddling = sum(1 for pieces in group.sink._pieces if len(pieces) > 1)
    assert straddling >= 1 and any(len(pieces) == 1 for pieces in group.sink._pieces)
    losses = torch.zeros(G)
    group.forward_backward(torch.cat(xs), torch.cat(ys), losses)
    for g in range(G):
        got = torch.cat([arenas[r][(first_row + g) * per:(first_row + g) * per + (hi - lo)] for r, (lo, hi) in enumerate(bounds)])
        assert torch.allclose(got, grads[g], rtol=1e-4, atol=1e-5), g
    untouched = torch.cat([a[:first_row * per] for a in arenas] + [a[(first_row + G) * per:] for a in arenas])
    assert torch.isnan(untouched).all()                      # other ranks' rows are never written


def test_milestone_hook_sees_final_tail_gradients_cpu():
    """add_milestone: when the hook fires, every coordinate from the milestone module's first parameter on is already
    in the rows (large gradients written in place, small ones by the partial flush) and is not touched again."""
    torch.manual_seed(5)
    module = _Net()
    G, B = 2, 3
    xs = [torch.randn(B, 3, 8, 8) for _ in range(G)]
    ys = [torch.randint(0, 5, (B,)) for _ in range(G)]
    loss_fn = torch.nn.CrossEntropyLoss()
    flat_model = FlatModel(copy.deepcopy(module), torch.device("cpu"))
    matrix = torch.full((G, flat_model.numel), float("nan"))
    group = GroupedWorkers(flat_model, [matrix[g] for g in range(G)], loss_fn, side_streams=0)
    seen = {}
    start = group.add_milestone(group.module.b2, lambda: seen.update(tail=matrix[:, start:].clone(), head_nan=bool(torch.isnan(matrix[:, :start]).any())))
    assert 0 < start < flat_model.numel
    losses = torch.zeros(G)
    group.forward_backward(torch.cat(xs), torch.cat(ys), losses)
    assert "tail" in seen and seen["head_nan"]                      # fired during the pass, before the head was complete
    assert not torch.isnan(seen["tail"]).any()
    assert torch.equal(seen["tail"], matrix[:, start:])             # final already at hook time
    grads, _, _ = _reference(module, xs, ys, loss_fn)
    for g in range(G):
        assert torch.allclose(matrix[g], grads[g], rtol=1e-4, atol=1e-5)
    matrix.fill_(float("nan"))                                      # a second pass behaves the same (pending state is reset)
    group.forward_backward(torch.cat(xs), torch.cat(ys), losses)
    for g in range(G):
        assert torch.allclose(matrix[g], grads[g], rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
def test_fused_resnet_blocks_on_own_kernels_match_separate_replicas_gpu():
    """A torchvision ResNet through the whole own-kernel path — tcgen05 convolutions, fused grouped BatchNorm, NHWC max-pool, the
    skip gradient summed in the data-gradient epilogue (fork node), grouped cross-entropy — against separate fp32 replicas."""
    torchvision = pytest.importorskip("torchvision")
    from garfield_b200.models.fused_resnet import fuse_resnet_blocks
    torch.manual_seed(4)
    dev = torch.device("cuda")
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        module = torchvision.models.resnet18(num_classes=10).to(dev)
        G, B = 4, 6
        xs = [torch.randn(B, 3, 32, 32, device=dev) for _ in range(G)]
        ys = [torch.randint(0, 10, (B,), device=dev) for _ in range(G)]
        loss_fn = torch.nn.CrossEntropyLoss()
        grads, losses, _ = _reference(module, xs, ys, loss_fn)
        from garfield_b200.models import fused_resnet
        fused = copy.deepcopy(module)
        assert fuse_resnet_blocks(fused) == 8
        fused_resnet.FORK_SKIP[0] = True
        try:
            flat_model, matrix, glosses, group = _run_grouped(fused, xs, ys, loss_fn, dev, channels_last=True)
        finally:
            fused_resnet.FORK_SKIP[0] = False
        torch.cuda.synchronize()
        assert any(hasattr(m, "_gb_fork") for m in group.module.modules())
        errs = [float((_logical(flat_model, matrix[g]) - grads[g]).norm() / grads[g].norm()) for g in range(G)]
        # the same pass with the skip gradients summed by a separate add kernel (no fork node) must agree with the fork path to
        # rounding: the fork only moves the addition into the data-gradient epilogue
        plain = copy.deepcopy(module)
        fuse_resnet_blocks(plain)
        flat2, matrix2, _, _ = _run_grouped(plain, xs, ys, loss_fn, dev, channels_last=True)
        torch.cuda.synchronize()
        for g in range(G):
            a, b = _logical(flat_model, matrix[g]), _logical(flat2, matrix2[g])
            assert (a - b).norm() <= 1e-4 * b.norm(), (g, float((a - b).norm() / b.norm()))
        # against exact fp32 replicas: TF32 operand rounding, amplified by BatchNorm over 6-sample batches on 1×1 planes
        assert max(errs) < 0.3, errs
        for g in range(G):
            assert torch.allclose(glosses[g], losses[g], rtol=2e-2)
    finally:
        torch.backends.cudnn.allow_tf32 = prev
