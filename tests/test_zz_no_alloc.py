"""Allocation-free pairwise/selection entry point, called through the C ABI.

`gb_pairwise_select_noalloc` exists for hosts that cannot allocate device memory when they need
an aggregation (all HBM already owned by their own pool, a fixed memory plan, or no tolerance for
cudaMalloc's device-wide synchronisation). It keeps its scratch in a fixed array inside the
library, so this test checks the two things such a caller relies on: it produces what the regular
entry point produces, and a second call is not disturbed by what the first one left in the shared
scratch. The file name makes pytest collect it last, so a `-x` run has already executed every
other test if this one fails.
"""

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MODES = {"none": 0, "krum": 1, "bulyan_native": 3, "mda": 4}


def _lib():
    from garfield_b200.ops import _lib as binding
    lib = binding.require()
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
    lib.gb_pairwise_select_noalloc.argtypes = [ctypes.POINTER(vp), i32, i64, i32, i32, i32, i32, i32, vp, vp, vp, vp]
    lib.gb_pairwise_select_noalloc.restype = i32
    return binding, lib


def _rows(n, d, seed):
    gen = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=gen)
    return [(base + torch.randn(d, generator=gen)).cuda() for _ in range(n)]


def _call_both(binding, lib, rows, mode, f, m, neighbors, squared):
    """Run the regular entry point (caller-provided scratch) and the no-alloc one on the same rows."""
    n, d = len(rows), rows[0].numel()
    table = binding.row_table(rows)
    stream = binding.stream_ptr(rows[0].device)
    outs = []
    for noalloc in (False, True):
        dist = torch.full((n, n), -1.0, device="cuda")
        W = torch.full((n, n), -1.0, device="cuda")
        info = torch.full((1 + n,), -1, dtype=torch.int32, device="cuda")
        if noalloc:
            rc = lib.gb_pairwise_select_noalloc(table, n, d, mode, f, m, neighbors, squared, dist.data_ptr(), W.data_ptr(),
                                                info.data_ptr(), stream)
        else:
            nbytes = int(lib.gb_pairwise_scratch_bytes(n))
            scratch = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
            rc = lib.gb_pairwise_select(table, n, d, mode, f, m, neighbors, squared, scratch.data_ptr(), nbytes, dist.data_ptr(),
                                        W.data_ptr(), info.data_ptr(), stream)
        assert rc == 0, lib.gb_cuda_last_error().decode()
        torch.cuda.synchronize()
        outs.append((dist.cpu(), W.cpu(), info.cpu()))
    return outs


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    binding, lib = _lib()
    # two different problems back to back: the second call starts from the first one's leftovers
    cases = [(_rows(8, 300001, seed=1), MODES["krum"], 1, 5, 5, 1, 1),
             (_rows(11, 70003, seed=2), MODES["bulyan_native"], 2, 7, 0, 1, 5),
             (_rows(16, 4096, seed=3), MODES["mda"], 2, 0, 0, 0, 1)]
    for rows, mode, f, m, neighbors, squared, theta in cases:
        (dist_a, w_a, info_a), (dist_b, w_b, info_b) = _call_both(binding, lib, rows, mode, f, m, neighbors, squared)
        n = len(rows)
        torch.testing.assert_close(dist_b, dist_a, rtol=2e-5, atol=1e-5)
        torch.testing.assert_close(w_b[:theta], w_a[:theta], rtol=2e-5, atol=1e-6)
        assert info_b[0].item() == info_a[0].item() == theta
        # and both agree with a plain PyTorch fp64 reference of the distances
        mat = torch.stack(rows).cpu().double()
        want = torch.cdist(mat, mat).pow(2)
        want.fill_diagonal_(float("inf"))
        torch.testing.assert_close(dist_b.double(), want, rtol=2e-5, atol=1e-5)
        assert w_b[0].sum().item() == pytest.approx(1.0, rel=1e-5) and n >= 1


def test_noalloc_rejects_more_rows_than_its_scratch_holds():
    binding, lib = _lib()
    rows = _rows(17, 64, seed=4)
    dist = torch.empty(17, 17, device="cuda")
    rc = lib.gb_pairwise_select_noalloc(binding.row_table(rows), 17, 64, 0, 0, 0, 0, 1, dist.data_ptr(), None, None,
                                        binding.stream_ptr(dist.device))
    assert rc != 0
    assert "16 rows" in lib.gb_cuda_last_error().decode()
