"""tcgen05 implicit-GEMM convolutions (csrc/cuda/conv_tc.cu) and the channels-last helper kernels (nn_nhwc.cu) against a
plain fp32 PyTorch reference on the CPU: forward, data gradient and per-worker weight gradients over the geometry
classes the kernels distinguish (1×1 / k×k, strides, tiles covering several images / parts of one image, explicit patch
path for channel counts the TMA boxes cannot address), plus pooling, the grouped cross-entropy and column sums."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TF32_TOL = 4e-3     # operands are read as TF32 (10-bit mantissa), accumulation is fp32


def _rel(a, b):
    return ((a.cpu() - b).abs().max() / b.abs().max().clamp_min(1e-20)).item()


CASES = [
    # n, h, w, cin, cout, k, stride, pad, groups
    (16, 8, 8, 64, 64, 1, 1, 0, 4),        # 1×1, two images per tile
    (16, 8, 8, 64, 96, 3, 1, 1, 4),        # 3×3, N tile wider than 64
    (16, 8, 8, 128, 128, 3, 2, 1, 4),      # strided 3×3 (traversal stride in the tensor map; upsampled data gradient)
    (16, 8, 8, 64, 160, 1, 2, 0, 2),       # strided 1×1 (scattered data gradient)
    (24, 2, 2, 256, 64, 3, 1, 1, 3),       # tiny planes: 32 images per tile, chunks of 8 images in wgrad
    (50, 1, 1, 512, 40, 3, 1, 1, 2),       # 1×1 plane under a 3×3 filter: only the centre tap is in range
    (50, 1, 1, 2048, 10, 1, 1, 0, 2),      # Linear geometry with a ragged N
    (4, 32, 32, 16, 32, 3, 1, 1, 2),       # wide plane: 4 rows per tile, 1 row per wgrad chunk
    (6, 16, 16, 3, 16, 7, 2, 3, 2),        # RGB stem: explicit patch matrix (49 taps, Cin = 3)
    (8, 14, 14, 6, 16, 5, 1, 0, 2),        # CifarNet conv2: explicit path everywhere (Cin = 6, plane 10×10)
    (6, 7, 7, 32, 48, 3, 1, 1, 3),         # odd plane: partial tiles, explicit wgrad
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,s,p,groups", CASES)
def test_conv_matches_fp32_reference(n, h, w, cin, cout, k, s, p, groups):
    from garfield_b200.ops import conv as C
    torch.manual_seed(n * 131 + h * 17 + cin)
    x = torch.randn(n, h, w, cin)
    wt = torch.randn(cout, k, k, cin) / (k * k * cin) ** 0.5
    bias = torch.randn(cout)
    xr = x.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wr = wt.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = F.conv2d(xr, wr, bias, s, p)
    dy = torch.randn_like(yr)
    gx, = torch.autograd.grad(yr, xr, dy, retain_graph=True)
    xd, wd, dyd = x.cuda(), wt.cuda(), dy.permute(0, 2, 3, 1).contiguous().cuda()
    y = C.conv_fprop(xd, wd, bias.cuda(), s, p)
    assert _rel(y, yr.detach().permute(0, 2, 3, 1)) < TF32_TOL
    dx = C.conv_dgrad(dyd, wd, (h, w), s, p)
    assert _rel(dx, gx.permute(0, 2, 3, 1)) < TF32_TOL
    dw = C.conv_wgrad(xd, dyd, groups, k, s, p)
    b = n // groups
    ref = []
    for g in range(groups):
        gw, = torch.autograd.grad(F.conv2d(xr[g * b:(g + 1) * b], wr, None, s, p), wr, dy[g * b:(g + 1) * b])
        ref.append(gw.permute(0, 2, 3, 1).reshape(cout, -1))
    assert _rel(dw, torch.stack(ref)) < TF32_TOL


def test_wgrad_writes_strided_rows_in_place():
    """The weight-gradient kernel writes [G, Cout, K] matrices with an arbitrary group stride (gradient rows)."""
    from garfield_b200.ops import conv as C
    torch.manual_seed(3)
    G, b, cin, cout = 3, 4, 32, 24
    x, dy = torch.randn(G * b, 4, 4, cin).cuda(), torch.randn(G * b, 4, 4, cout).cuda()
    pitch, off = 9 * cin * cout + 777, 5                       # unaligned offset inside each row
    rows = torch.full((G * pitch,), float("nan"), device="cuda")
    out = torch.as_strided(rows, (G, cout, 9 * cin), (pitch, 9 * cin, 1), off)
    C.conv_wgrad(x, dy, G, 3, 1, 1, out=out)
    dense = C.conv_wgrad(x, dy, G, 3, 1, 1)
    assert torch.equal(out, dense)
    untouched = torch.ones(G * pitch, dtype=torch.bool)
    for g in range(G):
        untouched[g * pitch + off:g * pitch + off + cout * 9 * cin] = False
    assert torch.isnan(rows.cpu()[untouched]).all()


def test_relu_epilogue_and_output_pitch():
    from garfield_b200.ops import conv as C
    torch.manual_seed(4)
    x, w = torch.randn(8, 4, 4, 64), torch.randn(20, 1, 1, 64) / 8
    wide = torch.full((8, 4, 4, 24), float("nan"), device="cuda")
    C.conv_fprop(x.cuda(), w.cuda(), None, 1, 0, relu=True, out=wide[..., :20])
    ref = torch.relu(F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2))).permute(0, 2, 3, 1)
    assert _rel(wide[..., :20], ref) < TF32_TOL and torch.isnan(wide[..., 20:]).all()


@pytest.mark.parametrize("n,h,w,c,k,s,p", [(6, 16, 16, 64, 3, 2, 1), (3, 9, 9, 8, 2, 2, 0), (2, 8, 8, 4, 3, 1, 1)])
def test_maxpool_nhwc(n, h, w, c, k, s, p):
    from garfield_b200.ops import conv as C
    torch.manual_seed(n + h)
    x = torch.randn(n, c, h, w, requires_grad=True)
    yr = F.max_pool2d(x, k, s, p)
    dy = torch.randn_like(yr)
    gx, = torch.autograd.grad(yr, x, dy)
    y, arg = C.max_pool_nhwc(x.detach().permute(0, 2, 3, 1).contiguous().cuda(), k, s, p)
    assert torch.equal(y.cpu(), yr.detach().permute(0, 2, 3, 1))
    dx = C.max_pool_nhwc_backward(dy.permute(0, 2, 3, 1).contiguous().cuda(), arg, (n, h, w, c), k, s, p)
    assert torch.allclose(dx.cpu(), gx.permute(0, 2, 3, 1), atol=1e-6)


def test_grouped_cross_entropy_and_colsum():
    from garfield_b200.ops import conv as C
    torch.manual_seed(9)
    G, b, classes = 4, 7, 10
    logits = torch.randn(G * b, classes, requires_grad=True)
    target = torch.randint(0, classes, (G * b,))
    per = F.cross_entropy(logits, target, reduction="none")
    per.reshape(G, b).mean(1).sum().backward()
    row_loss, dlogits = C.softmax_ce_grouped(logits.detach().cuda(), target.cuda(), G)
    assert dlogits.shape == (G * b, 12) and (dlogits[:, classes:] == 0).all()
    assert torch.allclose(row_loss.cpu(), per.detach(), atol=1e-5)
    assert torch.allclose(dlogits[:, :classes].cpu(), logits.grad, atol=1e-6)
    losses = C.colsum_grouped(row_loss.view(-1, 1), G, scale=1.0 / b)
    assert torch.allclose(losses.view(-1).cpu(), per.detach().reshape(G, b).mean(1), atol=1e-5)
    sums = C.colsum_grouped(dlogits, G, cols=classes)
    assert torch.allclose(sums.cpu(), logits.grad.reshape(G, b, classes).sum(1), atol=1e-6)


def test_avgpool_nhwc():
    from garfield_b200.ops import conv as C
    x = torch.randn(5, 3, 3, 12)
    y = C.avg_pool_nhwc(x.cuda())
    assert torch.allclose(y.cpu(), x.mean((1, 2)), atol=1e-6)
    dx = C.avg_pool_nhwc_backward(y, (3, 3))
    assert torch.allclose(dx.cpu(), (y.cpu() / 9)[:, None, None, :].expand(5, 3, 3, 12), atol=1e-7)


def test_relu_and_scalar_maxpool_kernels():
    """ReLU forward / backward and max pooling on a channel count that is not a multiple of 4 (CifarNet: 6 channels)."""
    from garfield_b200.ops import conv as C
    torch.manual_seed(21)
    x = torch.randn(5, 7, 3, 6)
    y = C.relu(x.cuda())
    assert torch.equal(y.cpu(), torch.relu(x))
    dy = torch.randn_like(x)
    assert torch.equal(C.relu_backward(dy.cuda(), y).cpu(), dy * (x > 0))
    xp = torch.randn(4, 6, 14, 14, requires_grad=True)
    yr = F.max_pool2d(xp, 2, 2)
    g = torch.randn_like(yr)
    gx, = torch.autograd.grad(yr, xp, g)
    yk, arg = C.max_pool_nhwc(xp.detach().permute(0, 2, 3, 1).contiguous().cuda(), 2, 2, 0)
    assert torch.equal(yk.cpu(), yr.detach().permute(0, 2, 3, 1))
    dx = C.max_pool_nhwc_backward(g.permute(0, 2, 3, 1).contiguous().cuda(), arg, (4, 14, 14, 6), 2, 2, 0)
    assert torch.allclose(dx.cpu(), gx.permute(0, 2, 3, 1), atol=1e-6)


def test_cifarnet_pass_runs_on_own_kernels_only():
    """Config 2's network through the fused pass: every convolution, Linear layer, ReLU, pooling and the loss is one of this
    library's kernels (counted launches), and the gradients match separate fp32 replicas within TF32 tolerance."""
    import copy
    from garfield_b200.models.small import Cifarnet
    from garfield_b200.ops import _lib
    from garfield_b200.parallel.arena import FlatModel
    from garfield_b200.parallel.grouped import GroupedWorkers
    torch.manual_seed(8)
    dev = torch.device("cuda")
    module = Cifarnet().to(dev)
    G, B = 3, 8
    xs = [torch.randn(B, 3, 32, 32, device=dev) for _ in range(G)]
    ys = [torch.randint(0, 10, (B,), device=dev) for _ in range(G)]
    loss_fn = torch.nn.CrossEntropyLoss()
    want = []
    for x, y in zip(xs, ys):
        m = copy.deepcopy(module)
        loss_fn(m(x), y).backward()
        want.append(torch.cat([p.grad.reshape(-1) for p in m.parameters()]))
    flat = FlatModel(copy.deepcopy(module), dev, channels_last=True)
    rows = torch.zeros(G, flat.numel, device=dev)
    group = GroupedWorkers(flat, [rows[g] for g in range(G)], loss_fn, side_streams=2)
    losses = torch.zeros(G, device=dev)
    data = torch.cat(xs).contiguous(memory_format=torch.channels_last)
    before = _lib.launches[0]
    group.forward_backward(data, torch.cat(ys), losses)
    torch.cuda.synchronize()
    assert _lib.launches[0] - before >= 2 * 3 + 5 * 3 + 4 + 2 + 2      # convs, linears (fwd/dgrad/wgrad), ReLUs, pools, loss
    for g in range(G):
        got = torch.cat([flat._view(rows[g], off, n, shape).reshape(-1) for off, n, shape in flat.slices])
        assert (got - want[g]).norm() <= 3e-2 * want[g].norm(), float((got - want[g]).norm() / want[g].norm())
