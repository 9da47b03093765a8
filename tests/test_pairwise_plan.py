"""Host-side launch plan of the exact pairwise-distance kernels (`gb_pairwise_plan`): which kernel a row count gets, the tile it
streams and — for the register-resident kernel — the assignment of block pairs to warps. No device work: runs without a GPU
(the sm_100a library itself loads on a CPU box; skipped only if it cannot be built or loaded there)."""
import ctypes

import pytest

from garfield_b200.ops import _lib


def plan(n):
    lib = _lib.load()
    if lib is None:
        pytest.skip("sm_100a library not loadable on this machine")
    kind, tile = ctypes.c_int(-1), ctypes.c_int(-1)
    slots = (ctypes.c_int * 2)()
    owner = ctypes.create_string_buffer(36)
    assert lib.gb_pairwise_plan(n, ctypes.byref(kind), ctypes.byref(tile), slots, owner) == 0
    return kind.value, tile.value, (slots[0], slots[1]), owner.raw


def test_kernel_choice_per_row_count():
    for n in range(1, 65):
        kind, tile, _, _ = plan(n)
        assert kind == (2 if 13 <= n <= 32 else 1 if n > 8 else 0), n
        assert tile % 128 == 0 and 128 <= tile <= 1024
        rows = (n + 3) // 4 * 4 if kind == 2 else n
        slots = ((8 * n * n + 31) // 32) * 32
        assert 4 * (slots + 3 * rows * tile) <= 227 * 1024, (n, tile)       # three stages + per-warp sums fit one CTA


@pytest.mark.parametrize("n", range(13, 33))
def test_every_block_pair_has_one_owner_within_the_slot_limits(n):
    kind, _, (qo, qd), owner = plan(n)
    assert kind == 2
    B = (n + 3) // 4
    pairs = [(i, j) for i in range(B) for j in range(i, B)]
    off, diag, load = [0] * 8, [0] * 8, [0.0] * 8
    for p, (i, j) in enumerate(pairs):
        w = owner[p]
        assert 0 <= w < 8
        if i == j:
            diag[w] += 1
            load[w] += 6
        else:
            off[w] += 1
            load[w] += 16
    assert max(off) <= qo and max(diag) <= qd
    assert qo + qd <= 5                                   # what the 168 registers of a 288-thread CTA hold (packed: 32 per slot)
    # balance: the busiest warp sets the pace of the CTA
    assert max(load) <= 1.36 * (sum(load) / 8), (n, load)
