"""Perf regression guards (last level): coarse lower bounds on the hot kernels, 2.5–3× away from the
measured values in profiles/README.md so that only a real regression (a kernel falling off its fast path) trips them.
Timed with CUDA events after warm-up, median of 5; inputs ≫ L2 for the streaming kernels."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _median_ms(fn, reps=5, inner=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(inner):
            fn()
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b) / inner)
    return sorted(times)[len(times) // 2]


def test_streaming_rules_stay_near_the_hbm_roofline():
    from garfield_b200.ops import gar
    if not gar.available():
        pytest.fail("sm_100a library missing on a GPU box")
    n, d = 8, 23_528_522                                        # the flagship's gradient matrix: 753 MB
    rows = [torch.randn(d, device="cuda") for _ in range(n)]
    bytes_read = n * d * 4
    median = _median_ms(lambda: gar.coordwise(rows, "median_lower"))
    krum = _median_ms(lambda: gar.aggregate("krum", rows, 1))
    bulyan = _median_ms(lambda: gar.aggregate("bulyan", rows, 1))
    # measured: 0.137 / 0.232 / 0.319 ms (6.4 TB/s effective on the median pass)
    assert median < 0.40, "coordinate-wise median took %.3f ms (%.2f TB/s)" % (median, bytes_read / median / 1e9)
    assert krum < 0.70, "Krum took %.3f ms" % krum
    assert bulyan < 0.95, "Bulyan took %.3f ms" % bulyan


def test_tcgen05_gemm_throughput():
    from garfield_b200.ops import nn as gnn
    m = n = k = 4096
    for dtype, floor_tflops in ((torch.bfloat16, 450.0), (torch.float32, 250.0)):      # measured: 1 197 and 638 TFLOP/s
        a = torch.randn(m, k, device="cuda").to(dtype)
        b = torch.randn(n, k, device="cuda").to(dtype)
        ms = _median_ms(lambda: gnn.gemm_nt(a, b))
        tflops = 2.0 * m * n * k / ms / 1e9
        assert tflops > floor_tflops, "%s GEMM at %.0f TFLOP/s" % (dtype, tflops)


def test_grouped_batchnorm_latency():
    from garfield_b200.ops import nn as gnn
    groups, batch = 8, 25
    x = torch.randn(groups * batch, 256, 8, 8, device="cuda").contiguous(memory_format=torch.channels_last)      # 13 MB, the largest flagship shape
    gamma, beta = torch.ones(256, device="cuda"), torch.zeros(256, device="cuda")
    rm, rv = torch.zeros(groups, 256, device="cuda"), torch.ones(groups, 256, device="cuda")
    ms = _median_ms(lambda: gnn.grouped_batch_norm(x, gamma, beta, rm, rv, groups, 1e-5, 0.1, None, True))
    assert ms < 0.15, "grouped BN forward took %.1f us (measured 13-17 us on the device; the bound leaves room for host launch overhead)" % (ms * 1e3)


def test_tcgen05_convolution_latency():
    """ResNet-50 layer shapes at the flagship batch (8 workers × 25): coarse bounds ≈3× above the measured device times
    (profiles/r2/conv_check_b200.log: 3×3 64→64 on 8×8: 7.6 / 7.3 / 20.1 µs; these eager timings include host launch overhead)."""
    from garfield_b200.ops import conv as C
    x = torch.randn(200, 8, 8, 64, device="cuda")
    w = torch.randn(64, 3, 3, 64, device="cuda") / 24
    dy = torch.randn(200, 8, 8, 64, device="cuda")
    assert _median_ms(lambda: C.conv_fprop(x, w, None, 1, 1)) < 0.08
    assert _median_ms(lambda: C.conv_dgrad(dy, w, (8, 8), 1, 1)) < 0.08
    assert _median_ms(lambda: C.conv_wgrad(x, dy, 8, 3, 1, 1)) < 0.10
