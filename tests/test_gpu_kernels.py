"""sm_100a kernels vs plain PyTorch fp32/fp64 references of the same ops (run with -m gpu)."""

import itertools
import math

import pytest
import torch

import aggregators
from garfield_b200.aggregators import _host

pytestmark = pytest.mark.gpu


def cuda_lib():
    from garfield_b200.ops import gar
    assert gar.available(), "sm_100a library must be loaded on a GPU box"
    return gar


def make(n, d, seed=0, spread=1.0, device="cuda"):
    gen = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=gen)
    return [(base + spread * torch.randn(d, generator=gen)).to(device) for _ in range(n)]


SHAPES = [(3, 1), (5, 7), (8, 1000), (8, 3 * 1024 + 5), (11, 50001), (16, 4096), (23, 9999), (33, 2049), (64, 700)]


def test_native_library_is_loaded():
    import garfield_b200._build as b
    cuda_lib()
    with open("/proc/self/maps") as maps:
        assert "libgarfield_b200_cuda.so" in maps.read()


@pytest.mark.parametrize("n,d", SHAPES)
def test_coordwise_rules(n, d):
    gar = cuda_lib()
    g = make(n, d, seed=n * 7 + d)
    mat = torch.stack(g).cpu()
    torch.testing.assert_close(gar.coordwise(g, "median_lower").cpu(), _host.median_lower(mat), rtol=0, atol=0)
    torch.testing.assert_close(gar.coordwise(g, "median_finite_upper").cpu(), _host.median_finite_upper(mat), rtol=0, atol=0)
    f = (n - 1) // 4
    torch.testing.assert_close(gar.coordwise(g, "trimmed_mean", f=f).cpu(), _host.trimmed_mean(mat, f), rtol=1e-5, atol=1e-6)
    beta = max(1, n - f)
    torch.testing.assert_close(gar.coordwise(g, "averaged_median", beta=beta).cpu(), _host.averaged_median(mat, beta), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gar.coordwise(g, "averaged_median", beta=beta, upper=True).cpu(),
                               _host.averaged_median(mat, beta, upper=True), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gar.average(g).cpu(), mat.mean(0), rtol=1e-5, atol=1e-6)


def test_coordwise_unaligned_rows_and_nonfinite():
    gar = cuda_lib()
    n, d = 9, 5000
    backing = [torch.randn(d + 3, device="cuda") for _ in range(n)]
    g = [b[1:d + 1] for b in backing]  # 4-byte aligned only → direct-load path
    mat = torch.stack(g).cpu()
    torch.testing.assert_close(gar.coordwise(g, "median_lower").cpu(), _host.median_lower(mat), rtol=0, atol=0)
    bad = [x.clone() for x in make(n, d, seed=5)]
    bad[0][10] = float("nan"); bad[1][10] = float("inf"); bad[2][11] = -float("inf")
    for b in bad:
        b[12] = float("nan")
    cpu = torch.stack(bad).cpu()
    got = gar.coordwise(bad, "median_finite_upper").cpu()
    torch.testing.assert_close(got, _host.median_finite_upper(cpu), rtol=0, atol=0)
    assert got[12].item() == 0.0
    low = gar.coordwise(bad, "median_lower").cpu()
    assert math.isnan(low[10].item()) and math.isnan(low[12].item())
    ok = torch.ones(d, dtype=torch.bool); ok[10] = ok[12] = False
    torch.testing.assert_close(low[ok], _host.median_lower(cpu)[ok], rtol=0, atol=0)


@pytest.mark.parametrize("n,d", SHAPES[2:])
def test_pairwise_distances(n, d):
    gar = cuda_lib()
    g = make(n, d, seed=d)
    got = gar.pairwise_sqdist(g).cpu().double()
    mat = torch.stack(g).cpu().double()
    want = torch.cdist(mat, mat).pow(2)
    want.fill_diagonal_(float("inf"))
    torch.testing.assert_close(got, want, rtol=2e-5, atol=1e-5)
    again = gar.pairwise_sqdist(g).cpu().double()
    assert torch.equal(got, again), "distances must be run-to-run deterministic"


@pytest.mark.parametrize("n", [9, 12, 13, 16, 17, 20, 21, 24, 25, 28, 29, 32])
def test_pairwise_distances_many_tiles_per_cta(n):
    """The 13-32-row kernel (register-resident sums, packed FP32, padded stages) and the packed general kernel over enough
    coordinates for several tiles per CTA plus a remainder; phantom rows of the padded stages must not leak into the sums."""
    gar = cuda_lib()
    d = 148 * 1024 * 3 + 77
    g = make(n, d, seed=n, spread=0.05)
    got = gar.pairwise_sqdist(g).cpu().double()
    want = torch.zeros(n, n, dtype=torch.float64)
    mat = torch.stack(g).double()
    for lo in range(0, d, 1 << 18):
        x = mat[:, lo:lo + (1 << 18)]
        want += ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1).cpu()
    want.fill_diagonal_(float("inf"))
    torch.testing.assert_close(got, want, rtol=5e-6, atol=0)
    assert torch.equal(got, gar.pairwise_sqdist(g).cpu().double()), "distances must be run-to-run deterministic"


@pytest.mark.parametrize("n", [13, 23, 32])
def test_pairwise_multi_kernel_unaligned_and_nonfinite_rows(n):
    gar = cuda_lib()
    d = 70001
    backing = [torch.randn(d + 3, device="cuda") for _ in range(n)]
    g = [b[1:d + 1] for b in backing]                    # 4-byte aligned only: every coordinate takes the direct-load path
    mat = torch.stack(g).cpu().double()
    want = torch.cdist(mat, mat).pow(2)
    want.fill_diagonal_(float("inf"))
    torch.testing.assert_close(gar.pairwise_sqdist(g).cpu().double(), want, rtol=2e-5, atol=1e-5)
    bad = make(n, d, seed=3)
    bad[n - 1][d // 2] = float("nan")
    bad[1][5] = float("inf")
    got = gar.pairwise_sqdist(bad).cpu()
    assert torch.isinf(got[n - 1]).all() and torch.isinf(got[:, n - 1]).all() and torch.isinf(got[1]).all()
    clean = [i for i in range(n) if i not in (1, n - 1)]
    sub = torch.stack([bad[i] for i in clean]).cpu().double()
    ref = torch.cdist(sub, sub).pow(2)
    ref.fill_diagonal_(float("inf"))
    torch.testing.assert_close(got[clean][:, clean].double(), ref, rtol=2e-5, atol=1e-5)


@pytest.mark.parametrize("n,beta", [(8, 4), (13, 9), (23, 18), (24, 3), (32, 20)])
def test_averaged_median_with_repeated_values(n, beta):
    """Columns full of repeated values (ties at the selection radius): the result must be the mean of β values that are all at
    least as close to the median as every value left out."""
    gar = cuda_lib()
    d = 5000
    gen = torch.Generator().manual_seed(n)
    mat = torch.randint(-3, 4, (n, d), generator=gen).float()
    g = [mat[i].cuda() for i in range(n)]
    got = gar.coordwise(g, "averaged_median", beta=beta).cpu().double()
    med = mat.sort(0).values[(n - 1) // 2]
    dist = (mat - med).abs()
    order = dist.sort(0).values
    radius = order[beta - 1]
    inside = (mat * (dist < radius)).double().sum(0)
    missing = beta - (dist < radius).sum(0)
    edge_vals = torch.where(dist == radius, mat, torch.full_like(mat, float("nan")))
    lo = inside + missing * torch.nan_to_num(edge_vals, nan=float("inf")).min(0).values.double()
    hi = inside + missing * torch.nan_to_num(edge_vals, nan=-float("inf")).max(0).values.double()
    assert ((got * beta >= lo - 1e-4) & (got * beta <= hi + 1e-4)).all()


RULE_CASES = [(7, 1, 3000), (8, 1, 40000), (11, 2, 12345), (15, 3, 5000), (23, 5, 2048), (33, 7, 1500)]


@pytest.mark.parametrize("n,f,d", RULE_CASES)
@pytest.mark.parametrize("name", ["krum", "native-krum", "bulyan", "native-bulyan", "brute", "native-brute", "aksel",
                                  "average", "median", "native-median", "trimmed_mean", "averaged_median"])
def test_registry_rules_on_cuda_match_cpu(name, n, f, d):
    cuda_lib()
    if name in ("brute", "native-brute") and math.comb(n, f) > 40000:
        pytest.skip("host oracle too slow")
    g = make(n, d, seed=n + f + d)
    cpu = [x.cpu() for x in g]
    rule = aggregators.gars[name]
    if rule.check(gradients=cpu, f=f) is not None:
        pytest.skip("invalid (n, f) for this rule")
    got = rule(gradients=g, f=f)
    assert got.is_cuda and got.dtype == torch.float32
    want = rule(gradients=cpu, f=f)
    torch.testing.assert_close(got.cpu(), want, rtol=3e-5, atol=3e-6)


def test_rules_ignore_byzantine_nonfinite_rows():
    cuda_lib()
    honest = make(9, 4000, seed=3, spread=0.05)
    attack = [torch.full((4000,), float("nan"), device="cuda"), torch.full((4000,), float("inf"), device="cuda")]
    g = honest + attack
    cpu = [x.cpu() for x in g]
    for name in ("krum", "native-krum", "bulyan", "native-bulyan", "brute", "native-median"):
        got = aggregators.gars[name](gradients=g, f=2)
        assert torch.isfinite(got).all(), name
        torch.testing.assert_close(got.cpu(), aggregators.gars[name](gradients=cpu, f=2), rtol=3e-5, atol=3e-6)


def test_krum_m_and_condense():
    gar = cuda_lib()
    g = make(11, 3000, seed=1)
    cpu = [x.cpu() for x in g]
    for m in (1, 3, 7):
        torch.testing.assert_close(aggregators.krum(gradients=g, f=2, m=m).cpu(), aggregators.krum(gradients=cpu, f=2, m=m), rtol=3e-5, atol=3e-6)
    out = aggregators.condense(gradients=g, f=2, p=0.9, seed=11).cpu()
    med = torch.stack(cpu).median(0).values
    from_med, from_first = out == med, out == cpu[0]
    assert bool((from_med | from_first).all()) and 0.85 < from_med.float().mean().item() < 0.95
    assert torch.equal(out, aggregators.condense(gradients=g, f=2, p=0.9, seed=11).cpu())


@pytest.mark.parametrize("momentum,nesterov,wd", [(0.0, False, 0.0), (0.9, False, 5e-4), (0.9, True, 1e-3)])
def test_fused_sgd_epilogue_matches_torch(momentum, nesterov, wd):
    gar = cuda_lib()
    from garfield_b200.ops.optim import FlatOptimizer
    n, f, d = 8, 1, 20000 + 3
    p0 = torch.randn(d)
    ref_p = torch.nn.Parameter(p0.clone())
    ref = torch.optim.SGD([ref_p], lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    flat = p0.clone().cuda()
    mirror = torch.zeros(d, device="cuda")
    opt = FlatOptimizer(flat, "sgd", lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    for step in range(3):
        g = make(n, d, seed=100 + step)
        agg_cpu = aggregators.gars["native-krum"](gradients=[x.cpu() for x in g], f=f)
        ref_p.grad = agg_cpu.clone()
        ref.step()
        out = gar.krum(g, f=f, m=n - f - 2, neighbors=n - f - 2, squared=True, update=opt.sgd_fusion(mirrors=[mirror]))
        opt.note_fused_step()
        torch.testing.assert_close(out.cpu(), agg_cpu, rtol=3e-5, atol=3e-6)
        torch.testing.assert_close(flat.cpu(), ref_p.detach(), rtol=3e-5, atol=3e-6)
        assert torch.equal(mirror, flat)


def test_flat_optimizers_cuda_match_torch():
    from test_optim_attacks import CONFIGS, run_pair
    cuda_lib()
    for name, kwargs in CONFIGS:
        got, want = run_pair(name, kwargs, "cuda")
        torch.testing.assert_close(got, want, rtol=3e-5, atol=3e-6, msg=lambda m: "%s %s: %s" % (name, kwargs, m))


def test_attacks_cuda():
    from garfield_b200.ops import attacks
    cuda_lib()
    g = torch.randn(100003, device="cuda")
    r = attacks.random_like(g, seed=1)
    assert r.min() >= 0 and r.max() < 1 and abs(r.mean().item() - 0.5) < 0.01
    assert torch.equal(r, attacks.random_like(g, seed=1)) and not torch.equal(r, attacks.random_like(g, seed=2))
    torch.testing.assert_close(attacks.reverse(g), g * -100)
    dropped = attacks.drop(g, p=0.3, seed=3)
    zeroed = dropped == 0
    assert 0.29 < zeroed.float().mean().item() < 0.31 and torch.equal(dropped[~zeroed], g[~zeroed])
    est = [torch.randn(100003, device="cuda") for _ in range(4)]
    stack = torch.stack(est)
    torch.testing.assert_close(attacks.lie(est), stack.mean(0) + 1.035 * stack.std(0), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(attacks.empire(est), -10 * stack.mean(0), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(attacks.lie(est[:1]), est[0])


@pytest.mark.parametrize("n,d", [(2, 4096), (8, 100000), (13, 70001), (16, 300003), (17, 50000), (23, 123457), (32, 65536 + 5)])
def test_tcgen05_gram_distances(n, d):
    """Tensor-core (tcgen05 / TMEM, TF32 operands) Gram distances vs the fp64 reference."""
    gar = cuda_lib()
    g = make(n, d, seed=n * 3 + d, spread=1.0)
    got = gar.pairwise_sqdist(g, engine="tc").cpu().double()
    mat = torch.stack(g).cpu().double()
    want = torch.cdist(mat, mat).pow(2)
    norms = mat.pow(2).sum(1)
    scale = norms[:, None] + norms[None, :]
    err = (got - want).abs()
    err.fill_diagonal_(0.0)
    assert torch.isinf(got.diagonal()).all()
    assert (err / scale).max().item() < 2e-3, (err / scale).max().item()
    exact = gar.pairwise_sqdist(g, engine="simt").cpu().double()
    rel = ((got - exact).abs() / scale)
    rel.fill_diagonal_(0.0)
    assert rel.max().item() < 2e-3


def test_tcgen05_gram_selection_matches_exact_kernel():
    gar = cuda_lib()
    n, f, d = 23, 5, 200000
    g = make(n, d, seed=77, spread=1.0)
    for mode, kw in (("krum", dict(f=f, m=n - f - 2, neighbors=n - f - 2, squared=True)),
                     ("bulyan_native", dict(f=f, m=n - f - 2, squared=True))):
        theta = n - 2 * f - 2 if mode.startswith("bulyan") else 1
        w_exact = gar.select(g, mode, engine="simt", **kw)["W"][:theta].clone()
        w_tc = gar.select(g, mode, engine="tc", **kw)["W"][:theta].clone()
        assert torch.equal(w_exact != 0, w_tc != 0), mode


def test_device_topology_engines_single_gpu():
    """MSMW and decentralized engines with every replica co-located on one GPU: honest replicas
    stay in agreement and the loss is finite under Byzantine workers and servers."""
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.parallel.topologies import DecentralizedEngine, MSMWEngine
    cuda_lib()
    eng = MSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 5, 3, 1, 1, "krum", 16, mar="median",
                     opt_args=dict(lr=0.05, momentum=0.9), byzantine_workers={4: "reverse"}, byzantine_servers={2: "random"})
    losses = [eng.step(i).item() for i in range(4)]
    assert all(l == l for l in losses)
    a = eng._smodel(0, 0, eng.cur).clone()
    b = eng._smodel(0, 1, eng.cur).clone()
    torch.testing.assert_close(a, b)          # the two honest servers agree after the model GAR
    assert torch.isfinite(a).all()
    eng.close()
    dec = DecentralizedEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 6, 1, "median", 16, opt_args=dict(lr=0.05),
                              byzantine={5: "random"}, non_iid=True)
    losses = [dec.step(i).item() for i in range(3)]
    assert all(l == l for l in losses)
    m0, m1 = dec.model_of(0).clone(), dec.model_of(1).clone()
    torch.testing.assert_close(m0, m1)        # contract step: every node holds the same aggregated model
    dec.close()


@pytest.mark.parametrize("gar", ["krum", "bulyan", "brute"])
def test_engine_distances_during_backward_match_plain_aggregation(gar):
    """Pairwise distances of the coordinate tail computed from a backward-pass hook + the head afterwards == one call."""
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.parallel.engine import SSMWEngine
    cuda_lib()
    finals = {}
    for overlap in (False, True):
        for graphs in (False, True):
            eng = SSMWEngine("cnn", "synthetic-cifar10", "cross-entropy", 8, 1, gar, 8, "sgd", dict(lr=0.02, momentum=0.9),
                             seed=3, overlap_distances=overlap, use_graphs=graphs)
            assert (eng._tail_at is not None) == overlap
            for module in [w.replica.module for w in eng.workers] + [eng.fused.module]:
                for m in module.modules():
                    if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                        m.p = 0.0
            for i in range(4):
                eng.step(i)
            torch.cuda.synchronize()
            finals[(overlap, graphs)] = eng.master.to_reference_order(eng.flat).clone()
            eng.close()
    for key in ((True, False), (True, True), (False, True)):
        diff = (finals[key] - finals[(False, False)]).abs()
        assert diff.max().item() < 2e-3 and (diff > 2e-4).float().mean().item() < 1e-3, (key, diff.max().item())


def test_msmw_worker_fusion_matches_per_worker_path():
    """ByzSGD engine: one fused pass for a rank's workers (shared median model, channels-last arenas, CUDA graph)
    gives the same server models as the worker-by-worker path."""
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.parallel.topologies import MSMWEngine
    cuda_lib()
    finals = {}
    for fuse in (False, True):
        eng = MSMWEngine("cnn", "synthetic-cifar10", "cross-entropy", 5, 3, 1, 1, "krum", 8, mar="median", opt_args=dict(lr=0.02, momentum=0.9),
                         byzantine_workers={4: "reverse"}, byzantine_servers={2: "reverse"}, fuse_workers=fuse, seed=5)
        assert (eng.fused is not None) == fuse
        for module in [w["replica"].module for w in eng.workers] + ([eng.fused.module] if fuse else []):
            for m in module.modules():
                if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                    m.p = 0.0
        losses = [eng.step(i).item() for i in range(4)]
        assert all(l == l for l in losses)
        torch.cuda.synchronize()
        finals[fuse] = eng.layout.to_reference_order(eng.server_model(0)).clone()
        eng.close()
    diff = (finals[True] - finals[False]).abs()
    assert diff.max().item() < 2e-3 and (diff > 2e-4).float().mean().item() < 1e-4, diff.max().item()


def test_decentralized_worker_fusion_matches_per_node_path():
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.parallel.topologies import DecentralizedEngine
    cuda_lib()
    finals = {}
    for fuse in (False, True):
        eng = DecentralizedEngine("cnn", "synthetic-cifar10", "cross-entropy", 6, 1, "median", 8, opt_args=dict(lr=0.02, momentum=0.9),
                                  byzantine={5: "reverse"}, non_iid=True, fuse_workers=fuse, seed=7)
        assert (eng.fused is not None) == fuse
        for module in [node["replica"].module for node in eng.nodes] + ([eng.fused.module] if fuse else []):
            for m in module.modules():
                if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                    m.p = 0.0
        losses = [eng.step(i).item() for i in range(4)]
        assert all(l == l for l in losses)
        torch.cuda.synchronize()
        finals[fuse] = eng.layout.to_reference_order(eng.model_of(0)).clone()
        torch.testing.assert_close(eng.model_of(0), eng.model_of(1))
        eng.close()
    diff = (finals[True] - finals[False]).abs()
    assert diff.max().item() < 2e-3 and (diff > 2e-4).float().mean().item() < 1e-3, diff.max().item()


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (25, 10, 84), (200, 120, 400), (77, 300, 1000), (512, 1024, 4096), (1000, 4096, 25088 // 8),
                                   (4096, 4096, 512), (2048 + 7, 4096 + 20, 1000), (300, 40000, 136)])   # the last three: CTA-pair kernel
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_tcgen05_gemm_nt(m, n, k, dtype):
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    gen = torch.Generator().manual_seed(m + n + k)
    a = torch.randn(m, k, generator=gen).cuda().to(dtype)
    b = torch.randn(n, k, generator=gen).cuda().to(dtype)
    bias = torch.randn(n, generator=gen).cuda()
    want = a.double() @ b.double().t() + bias.double()
    got = gnn.gemm_nt(a, b, bias=bias)
    scale = (a.double().abs() @ b.double().abs().t()).clamp_min(1.0)   # error budget ∝ Σ|a||b|
    tol = 2e-3 if dtype == torch.float32 else 1e-5                      # TF32 operand rounding vs exact bf16 products
    assert ((got.double() - want).abs() / scale).max().item() < tol
    relu = gnn.gemm_nt(a, b, bias=bias, relu=True)
    torch.testing.assert_close(relu, got.clamp_min(0.0))


def test_tcgen05_linear_autograd():
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    torch.manual_seed(0)
    x = torch.randn(50, 400, device="cuda", requires_grad=True)
    layer = gnn.Linear(400, 120).cuda()
    ref = torch.nn.Linear(400, 120).cuda()
    ref.load_state_dict(layer.state_dict())
    x2 = x.detach().clone().requires_grad_(True)
    y = layer(x)
    y.square().sum().backward()
    torch.backends.cuda.matmul.allow_tf32 = False
    y2 = ref(x2)
    y2.square().sum().backward()
    torch.testing.assert_close(y, y2, rtol=5e-3, atol=5e-3)
    torch.testing.assert_close(x.grad, x2.grad, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(layer.weight.grad, ref.weight.grad, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(layer.bias.grad, ref.bias.grad, rtol=2e-2, atol=5e-2)   # inherits the TF32 error of y


def test_small_cnn_layer_kernels_match_torch():
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    torch.manual_seed(3)
    F = torch.nn.functional
    for (n, cin, h, cout, k, stride, pad) in [(25, 3, 32, 6, 5, 1, 0), (8, 6, 14, 16, 5, 1, 0), (4, 5, 11, 7, 3, 2, 1), (2, 1, 28, 10, 5, 1, 0)]:
        x = torch.randn(n, cin, h, h, device="cuda", requires_grad=True)
        w = torch.randn(cout, cin, k, k, device="cuda", requires_grad=True)
        b = torch.randn(cout, device="cuda", requires_grad=True)
        xr, wr, br = (t.detach().double().cpu().requires_grad_(True) for t in (x, w, b))
        y = gnn.conv2d(x, w, b, stride, pad, relu=True)
        yr = torch.relu(F.conv2d(xr, wr, br, stride, pad))
        torch.testing.assert_close(y.cpu().double(), yr, rtol=1e-4, atol=1e-4)
        g = torch.randn_like(y)
        y.backward(g)
        yr.backward(g.cpu().double())
        for got, want in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
            torch.testing.assert_close(got.cpu().double(), want, rtol=1e-3, atol=1e-3)
    x = torch.randn(6, 4, 13, 13, device="cuda", requires_grad=True)
    xr = x.detach().cpu().requires_grad_(True)
    for kernel, stride in ((2, 2), (3, 2)):
        y, yr = gnn.max_pool2d(x, kernel, stride), F.max_pool2d(xr, kernel, stride)
        torch.testing.assert_close(y.cpu(), yr)
        g = torch.randn_like(y)
        x.grad = None; xr.grad = None
        y.backward(g); yr.backward(g.cpu())
        torch.testing.assert_close(x.grad.cpu(), xr.grad)
    logits = torch.randn(50, 10, device="cuda", requires_grad=True)
    target = torch.randint(0, 10, (50,), device="cuda")
    lr = logits.detach().cpu().requires_grad_(True)
    loss, loss_r = gnn.cross_entropy(logits, target), F.cross_entropy(lr, target.cpu())
    loss.backward(); loss_r.backward()
    torch.testing.assert_close(loss.cpu(), loss_r, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(logits.grad.cpu(), lr.grad, rtol=1e-5, atol=1e-6)
    logp = torch.log_softmax(torch.randn(33, 10, device="cuda"), 1).requires_grad_(True)
    lp = logp.detach().cpu().requires_grad_(True)
    a, b2 = gnn.nll_loss(logp, target[:33]), F.nll_loss(lp, target[:33].cpu())
    a.backward(); b2.backward()
    torch.testing.assert_close(a.cpu(), b2, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(logp.grad.cpu(), lp.grad, rtol=1e-5, atol=1e-6)
    x = torch.randn(16, 12, 7, 7, device="cuda", requires_grad=True)
    gamma, beta = torch.rand(12, device="cuda", requires_grad=True), torch.randn(12, device="cuda", requires_grad=True)
    rm, rv = torch.zeros(12, device="cuda"), torch.ones(12, device="cuda")
    xr, gr, br = (t.detach().cpu().requires_grad_(True) for t in (x, gamma, beta))
    rmr, rvr = torch.zeros(12), torch.ones(12)
    y = gnn.batch_norm_train(x, gamma, beta, rm, rv)
    yr = F.batch_norm(xr, rmr, rvr, gr, br, True, 0.1, 1e-5)
    torch.testing.assert_close(y.cpu(), yr, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rm.cpu(), rmr, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rv.cpu(), rvr, rtol=1e-4, atol=1e-5)
    g = torch.randn_like(y)
    y.backward(g); yr.backward(g.cpu())
    for got, want in ((x.grad, xr.grad), (gamma.grad, gr.grad), (beta.grad, br.grad)):
        torch.testing.assert_close(got.cpu(), want, rtol=1e-3, atol=1e-4)


def test_native_cifarnet_matches_torch_model():
    """CifarNet on hand-written kernels only (conv/pool/tcgen05 GEMM) vs the torch-op model: same
    state_dict, same loss and gradients up to TF32 rounding of the linear layers."""
    from garfield_b200.garfieldpp import tools
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    torch.manual_seed(0)
    torch.backends.cuda.matmul.allow_tf32 = False
    ours = tools.select_model("cifarnet-native", torch.device("cuda"), "cifar10")
    ref = tools.select_model("cifarnet", torch.device("cuda"), "cifar10")
    ref.load_state_dict(ours.state_dict())
    x, t = torch.randn(25, 3, 32, 32, device="cuda"), torch.randint(0, 10, (25,), device="cuda")
    la = gnn.cross_entropy(ours(x), t)
    lb = torch.nn.functional.cross_entropy(ref(x), t)
    la.backward(); lb.backward()
    assert abs(la.item() - lb.item()) < 5e-3
    for (name, p), q in zip(ours.named_parameters(), ref.parameters()):
        denom = q.grad.abs().max().item() + 1e-6
        assert (p.grad - q.grad).abs().max().item() / denom < 3e-2, name


def test_engine_gradient_gather_matches_bound_grads():
    """One multi-tensor gather launch after backward == autograd accumulating into rows bound as .grad;
    channels-last arenas == NCHW arenas (in reference order); CUDA graph == eager."""
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.parallel.engine import SSMWEngine
    cuda_lib()
    finals = {}
    for tag, kw in (("bound", dict(gather_grads=False, fast_bn=False, use_graphs=False, split_backward=False)),
                    ("gather", dict(gather_grads=True, fast_bn=True, use_graphs=False, fuse_workers=False)),
                    ("graph", dict(gather_grads=True, fast_bn=True, use_graphs=True, fuse_workers=False)),
                    ("nchw", dict(gather_grads=True, fast_bn=True, use_graphs=False, channels_last=False, fuse_workers=False)),
                    ("fused", dict(use_graphs=False, fuse_workers=True)),
                    ("fused_graph", dict(use_graphs=True, fuse_workers=True)),
                    ("fused_nchw", dict(use_graphs=False, fuse_workers=True, channels_last=False))):
        eng = SSMWEngine("cnn", "synthetic-cifar10", "cross-entropy", 7, 1, "median", 8, "sgd", dict(lr=0.02, momentum=0.9),
                         byzantine={6: "reverse"}, seed=3, **kw)
        for m in eng.master.module.modules():   # dropout would make the runs incomparable
            if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                m.p = 0.0
        assert (eng.fused is not None) == tag.startswith("fused")
        for module in [w.replica.module for w in eng.workers] + ([eng.fused.module] if eng.fused is not None else []):
            for m in module.modules():
                if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                    m.p = 0.0
        for i in range(3):
            eng.step(i)
        torch.cuda.synchronize()
        finals[tag] = eng.master.to_reference_order(eng.flat).clone()
        bn = [b.clone() for b in eng.workers[0].replica.module.buffers() if b.dtype == torch.float32]
        finals[tag + "_bn"] = torch.cat([b.reshape(-1) for b in bn])
        eng.close()
    for tag in ("gather", "graph", "nchw", "fused", "fused_graph", "fused_nchw"):
        diff = (finals[tag] - finals["bound"]).abs()
        assert diff.max().item() < 2e-3 and (diff > 2e-4).float().mean().item() < 1e-4, "%s vs bound: max %g" % (tag, diff.max().item())
        if "graph" not in tag:  # graph capture replays its warm-up iterations, which also move the BN running statistics
            torch.testing.assert_close(finals[tag + "_bn"], finals["bound_bn"], rtol=2e-3, atol=5e-4)   # TF32 convolutions, different batch shapes


def test_avgpool_dropout_bce_kernels():
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    F = torch.nn.functional
    x = torch.randn(9, 7, 5, 5, device="cuda", requires_grad=True)
    xr = x.detach().cpu().requires_grad_(True)
    y, yr = gnn.global_avg_pool(x), xr.mean(dim=(2, 3))
    torch.testing.assert_close(y.cpu(), yr, rtol=1e-5, atol=1e-6)
    g = torch.randn_like(y)
    y.backward(g); yr.backward(g.cpu())
    torch.testing.assert_close(x.grad.cpu(), xr.grad, rtol=1e-5, atol=1e-6)
    v = torch.randn(100003, device="cuda", requires_grad=True)
    d = gnn.dropout(v, 0.3, True, seed=5)
    kept = d != 0
    assert 0.68 < kept.float().mean().item() < 0.72
    torch.testing.assert_close(d[kept], v.detach()[kept] / 0.7)
    d.sum().backward()
    torch.testing.assert_close(v.grad, kept.float() / 0.7)            # same mask regenerated in backward
    assert torch.equal(gnn.dropout(v, 0.3, False), v)
    prob = torch.rand(64, 1, device="cuda").clamp(1e-4, 1 - 1e-4).requires_grad_(True)
    tgt = (torch.rand(64, 1, device="cuda") > 0.5).float()
    pr = prob.detach().cpu().requires_grad_(True)
    a, b = gnn.binary_cross_entropy(prob, tgt), F.binary_cross_entropy(pr, tgt.cpu())
    a.backward(); b.backward()
    torch.testing.assert_close(a.cpu(), b, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(prob.grad.cpu(), pr.grad, rtol=1e-4, atol=1e-6)


def test_engine_checkpoint_resume_and_accuracy(tmp_path):
    import os
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")
    from garfield_b200.garfieldpp.datasets import DatasetManager
    from garfield_b200.parallel.engine import SSMWEngine
    cuda_lib()
    torch.backends.cudnn.deterministic = True
    try:
        def make_engine():
            return SSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 7, 1, "krum", 16, "sgd", dict(lr=0.05, momentum=0.9),
                              byzantine={6: "reverse"}, use_graphs=False, seed=11)
        a = make_engine()
        for i in range(3):
            a.step(i)
        a.save_checkpoint(str(tmp_path / "ck"))
        for i in range(3, 6):
            a.step(i)
        torch.cuda.synchronize()
        want = a.flat.clone()
        test_set = DatasetManager("synthetic-cifar10", 16, 7, 8, 1).get_test_set()
        acc = a.accuracy(test_set[:3])
        assert 0.0 <= acc <= 100.0
        a.close()
        b = make_engine()
        assert b.load_checkpoint(str(tmp_path / "ck")) == 3
        for i in range(3, 6):
            b.step(i)
        torch.cuda.synchronize()
        assert torch.equal(b.flat, want), (b.flat - want).abs().max().item()
        b.close()
    finally:
        torch.backends.cudnn.deterministic = False


def test_fused_nhwc_batchnorm_matches_torch():
    from garfield_b200.ops import nn as gnn
    cuda_lib()
    torch.manual_seed(1)
    for (n, c, h) in [(25, 64, 16), (25, 256, 8), (25, 2048, 1), (3, 12, 5)]:
        for relu, with_res in ((True, False), (True, True), (False, False)):
            x = torch.randn(n, c, h, h, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
            res = torch.randn(n, c, h, h, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True) if with_res else None
            bn = torch.nn.BatchNorm2d(c).cuda().train()
            with torch.no_grad():
                bn.weight.uniform_(0.5, 1.5); bn.bias.normal_()
            ref = torch.nn.BatchNorm2d(c).cuda().train()
            ref.load_state_dict(bn.state_dict())
            xr = x.detach().clone().requires_grad_(True)
            rr = res.detach().clone().requires_grad_(True) if with_res else None
            y = gnn.fused_batch_norm(x, bn, residual=res, relu=relu)
            yr = ref(xr)
            if with_res:
                yr = yr + rr
            if relu:
                yr = torch.relu(yr)
            torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
            torch.testing.assert_close(bn.running_mean, ref.running_mean, rtol=1e-4, atol=1e-5)
            torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-4, atol=1e-5)
            g = torch.randn_like(y)
            y.backward(g); yr.backward(g)
            torch.testing.assert_close(x.grad, xr.grad, rtol=1e-3, atol=1e-4)
            torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=1e-3, atol=1e-3)
            torch.testing.assert_close(bn.bias.grad, ref.bias.grad, rtol=1e-3, atol=1e-3)
            if with_res:
                torch.testing.assert_close(res.grad, rr.grad, rtol=1e-4, atol=1e-5)


def test_fused_resnet50_matches_torchvision():
    import copy
    import torchvision
    from garfield_b200.models.fused_resnet import fuse_resnet_blocks
    cuda_lib()
    torch.manual_seed(0)
    # full-precision, deterministic convolutions: the only difference left between the two models is the
    # BatchNorm/ReLU/add implementation (with TF32 convolutions a 50-layer net at batch 25 amplifies the
    # convolution rounding noise into O(1) relative gradient differences between ANY two runs)
    tf32, det = torch.backends.cudnn.allow_tf32, torch.backends.cudnn.deterministic
    torch.backends.cudnn.allow_tf32, torch.backends.cudnn.deterministic = False, True
    try:
        _compare_fused_resnet(torchvision, fuse_resnet_blocks, copy)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cudnn.deterministic = tf32, det


def _compare_fused_resnet(torchvision, fuse_resnet_blocks, copy):
    import types
    from garfield_b200.models.fused_resnet import _bottleneck_forward
    ref = torchvision.models.resnet50(num_classes=10).cuda().to(memory_format=torch.channels_last).train()
    ours = copy.deepcopy(ref)
    assert fuse_resnet_blocks(ours) == 16
    x = torch.randn(25, 3, 32, 32, device="cuda").contiguous(memory_format=torch.channels_last)
    t = torch.randint(0, 10, (25,), device="cuda")

    def flat_grads(model, inp):
        for p in model.parameters():
            p.grad = None
        loss = torch.nn.functional.cross_entropy(model(inp), t)
        loss.backward()
        return loss.item(), torch.cat([p.grad.reshape(-1) for p in model.parameters()])

    loss_ref, g_ref = flat_grads(ref, x)
    loss_ours, g_ours = flat_grads(ours, x)
    # a 50-layer BatchNorm net at batch 25 is ill-conditioned at initialisation: measure how much the REFERENCE's
    # own gradient moves under a 1e-6 relative input perturbation and require the fused model to be that close
    _, g_pert = flat_grads(copy.deepcopy(ref), x * (1 + 1e-6))
    noise = ((g_pert - g_ref).norm() / g_ref.norm()).item()
    err = ((g_ours - g_ref).norm() / g_ref.norm()).item()
    assert abs(loss_ours - loss_ref) < 1e-3
    assert err < 3 * noise + 1e-4, (err, noise)
    assert torch.nn.functional.cosine_similarity(g_ours, g_ref, dim=0).item() > 0.999
    # one block in isolation is well conditioned: tight agreement
    blk = copy.deepcopy(ref.layer2[1])
    fused = copy.deepcopy(blk)
    fused.forward = types.MethodType(_bottleneck_forward, fused)
    xb = torch.randn(25, 512, 4, 4, device="cuda").contiguous(memory_format=torch.channels_last)
    outs = []
    for b in (blk, fused):
        xi = xb.clone().requires_grad_(True)
        b(xi).square().mean().backward()
        outs.append(torch.cat([p.grad.reshape(-1) for p in b.parameters()] + [xi.grad.reshape(-1)]))
    torch.testing.assert_close(outs[1], outs[0], rtol=1e-4, atol=1e-7)
    ours.eval(); ref.eval()
    torch.testing.assert_close(ours(x), ref(x), rtol=1e-3, atol=1e-3)


def test_gossip_rounds_kernel_matches_round_by_round():
    """Path 4: several agreement rounds inside one persistent kernel == one coordwise launch per round. Two of the six nodes
    are hosted here (their buffers follow the rounds), the other four keep whatever their buffers hold (stale peers)."""
    cuda_lib()
    from garfield_b200.ops import gar
    torch.manual_seed(12)
    n, d, local = 6, 70_001, (1, 4)
    for name, f in (("median", 1), ("trimmed_mean", 1), ("averaged_median", 2)):
        for rounds in (1, 2, 3):
            bufs = [[torch.randn(d, device="cuda") for _ in range(n)] for _ in range(2)]      # bufs[parity][node]
            ref = [[t.clone() for t in row] for row in bufs]
            first = 1
            parity = first
            for _ in range(rounds):
                out = gar.aggregate(name, ref[parity], f)
                for k in local:
                    ref[1 - parity][k].copy_(out)
                parity = 1 - parity
            seq = torch.zeros(1, dtype=torch.int32, device="cuda")
            counter = torch.zeros(1, dtype=torch.int32, device="cuda")
            gar.gossip_rounds(name, bufs[0], bufs[1], [bufs[1][k] for k in local], [bufs[0][k] for k in local], f, rounds, first, seq, counter)
            torch.cuda.synchronize()
            for p in range(2):
                for k in range(n):
                    assert torch.equal(bufs[p][k], ref[p][k]), (name, rounds, p, k)


def test_pull_median_operand_matches_torch():
    """Path 3 head: the median of the replicas' first-layer weights lands in the flat model and, zero-padded, in the GEMM operand."""
    cuda_lib()
    from garfield_b200.ops import gar
    torch.manual_seed(13)
    cout, k, kpad = 64, 147, 148
    for replicas in (2, 3, 5):
        rows = [torch.randn(cout * k + 33, device="cuda") for _ in range(replicas)]
        flat = torch.full((cout * k,), float("nan"), device="cuda")
        operand = torch.full((cout, kpad), float("nan"), device="cuda")
        gar.pull_median_operand([r[:cout * k] for r in rows], cout, k, flat, operand)
        want = torch.stack([r[:cout * k] for r in rows]).median(0)[0]
        assert torch.equal(flat, want)
        assert torch.equal(operand[:, :k], want.view(cout, k)) and (operand[:, k:] == 0).all()


def test_registry_rules_stay_exact_on_tightly_clustered_gradients():
    """Honest gradients that differ by 1e-3 of their norm on top of a large common offset (late training, large batches) plus f
    outliers: the default engine is the exact direct-difference kernel for every n, so the selected set equals the fp64
    reference; TF32 Gram distances cannot resolve these rows, which is why that kernel is opt-in."""
    cuda_lib()
    from garfield_b200.ops import gar
    torch.manual_seed(77)
    n, f, d = 20, 3, 200_000
    common = torch.randn(d, dtype=torch.float64) * 10
    rows64 = [common + 1e-3 * torch.randn(d, dtype=torch.float64) * (1 + 0.05 * i) for i in range(n - f)]
    rows64 += [common + 0.5 * torch.randn(d, dtype=torch.float64) for _ in range(f)]
    rows = [r.float().cuda() for r in rows64]
    ws = gar.select(rows, "krum", f=f, m=n - f - 2, neighbors=n - f - 2, squared=True)          # engine="auto"
    torch.cuda.synchronize()
    picked = sorted(ws["info"][1:1 + n - f - 2].tolist())
    stack = torch.stack([r.double() for r in rows]).cpu()
    dist = torch.cdist(stack, stack).pow(2)
    dist.fill_diagonal_(float("inf"))
    scores = dist.sort(1)[0][:, :n - f - 2].sum(1)
    want = sorted(scores.argsort()[:n - f - 2].tolist())
    assert picked == want
    assert all(i < n - f for i in picked)                         # no outlier selected
    exact = gar.pairwise_sqdist(rows, engine="simt").cpu().double()
    off = ~torch.eye(n, dtype=torch.bool)
    assert ((exact - dist)[off].abs() / dist[off]).max().item() < 1e-4
    tc = gar.pairwise_sqdist(rows, engine="tc").cpu().double()
    honest = off.clone(); honest[n - f:] = False; honest[:, n - f:] = False
    assert ((tc - dist)[honest].abs() / dist[honest]).max().item() > 1e-2      # what the TF32 Gram form does to such rows


def test_select_only_rejects_out_of_range_parameters():
    cuda_lib()
    from garfield_b200.ops import gar
    ws = gar.select([torch.randn(1000, device="cuda") for _ in range(7)], "none")
    for kwargs in (dict(mode="krum", f=1, m=9, neighbors=5), dict(mode="krum", f=1, m=3, neighbors=9), dict(mode="bulyan_native", f=2, m=1),
                   dict(mode="closest", m=12), dict(mode="mda", f=4)):
        with pytest.raises(RuntimeError):
            gar.select_only(ws, 7, torch.device("cuda"), **kwargs)


@pytest.mark.parametrize("name,n,f", [("average", 5, 0), ("median", 7, 1), ("native-median", 6, 1), ("trimmed_mean", 9, 2), ("averaged_median", 8, 2),
                                      ("krum", 9, 2), ("native-krum", 9, 2), ("bulyan", 11, 2), ("native-bulyan", 11, 2), ("brute", 7, 2),
                                      ("native-brute", 7, 2), ("aksel", 7, 2)])
def test_float64_rows_run_on_the_fp64_device_kernels(name, n, f):
    """CUDA float64 inputs (the reference natives' gpu_double dispatch): same registry call, fp64 SIMT kernels, results equal
    the CPU registry on the same doubles to double precision."""
    cuda_lib()
    from garfield_b200.ops import _lib
    torch.manual_seed(n * 7 + f)
    d = 10_007
    rows = [torch.randn(d, dtype=torch.float64) for _ in range(n)]
    if name.startswith("native-median"):
        rows[2][5] = float("nan"); rows[3][9] = float("inf")
    want = aggregators.gars[name](gradients=[r.clone() for r in rows], f=f)
    before = _lib.launches[0]
    got = aggregators.gars[name](gradients=[r.cuda() for r in rows], f=f)
    assert _lib.launches[0] > before, "the fp64 device kernels must have run"
    assert got.dtype == torch.float64 and got.is_cuda
    # sums run in a different order than on the host: a few ulps of double
    assert torch.allclose(got.cpu(), want, rtol=1e-11, atol=1e-12, equal_nan=True), (got.cpu() - want).abs().max()
