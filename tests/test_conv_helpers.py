"""Host-side logic of the tensor-core convolution drivers (ops/conv.py) and of the scatter sink: geometry helpers and the rule
that library kernels only ever receive this GPU's memory. Runs without a GPU."""
import torch

from garfield_b200.ops import conv as C


def test_out_size_matches_torch():
    for h, w, k, s, p in ((32, 32, 7, 2, 3), (8, 8, 3, 1, 1), (8, 8, 3, 2, 1), (28, 28, 5, 1, 0), (1, 1, 3, 1, 1), (14, 9, (3, 5), (2, 1), (1, 2))):
        y = torch.nn.functional.conv2d(torch.zeros(1, 1, h, w), torch.zeros(1, 1, *C._pair(k)), None, s, p)
        assert C.out_size(h, w, k, s, p) == tuple(y.shape[2:])


def test_pixel_pitch_and_density_of_views():
    dense = torch.zeros(6, 4, 4, 16)
    assert C._pitch(dense) == 16 and C._dense_pixels(dense)
    wide = torch.zeros(6, 4, 4, 24)[..., :20]                  # channel slice of a wider pixel: still consecutive pixels
    assert C._pitch(wide) == 24 and C._dense_pixels(wide)
    logits = torch.zeros(10, 12)[:, :10].unsqueeze(1).unsqueeze(1)     # padded logit gradients viewed as a 1×1 plane
    assert C._pitch(logits) == 12 and C._dense_pixels(logits)
    nchw_view = torch.zeros(6, 16, 4, 4).permute(0, 2, 3, 1)           # NCHW storage seen as NHWC: channels are not contiguous
    assert not C._dense_pixels(nchw_view)
    cl = torch.zeros(6, 16, 4, 4).contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
    assert C._dense_pixels(cl) and C._pitch(cl) == 16
    strided_rows = torch.zeros(6, 8, 4, 16)[:, ::2]                    # every other row: not one pitch
    assert not C._dense_pixels(strided_rows)
    single = torch.zeros(1, 1, 1, 10)
    assert C._pitch(single) == 12                                       # a single pixel: pitch rounded up to 16 bytes


def test_implicit_path_conditions():
    assert C._implicit_ok(64, 3, 3, 8) and C._implicit_ok(4, 1, 1, 128)
    assert not C._implicit_ok(3, 7, 7, 16)            # RGB stem: 147 floats per patch row, 49 taps
    assert not C._implicit_ok(6, 5, 5, 10)            # CifarNet conv2: 150 floats per row
    assert not C._implicit_ok(64, 3, 3, 224)          # plane wider than one tile
    assert C._wgrad_implicit_ok(8, 8) and C._wgrad_implicit_ok(2, 2) and C._wgrad_implicit_ok(1, 1) and C._wgrad_implicit_ok(32, 32)
    assert not C._wgrad_implicit_ok(7, 7) and not C._wgrad_implicit_ok(14, 14) and not C._wgrad_implicit_ok(16, 64)


def test_scatter_sink_hands_library_kernels_local_memory_only():
    """direct(): the local arena when this rank aggregates the whole parameter, else a local staging tensor — never a peer's arena."""
    from garfield_b200.parallel.grouped import ScatterSink
    world, per, G = 3, 64, 2
    params = [torch.nn.Parameter(torch.zeros(8, 4)), torch.nn.Parameter(torch.zeros(16, 4)), torch.nn.Parameter(torch.zeros(20, 4))]
    slices, off = [], 0
    for p in params:
        slices.append((off, p.numel(), tuple(p.shape)))
        off += p.numel()
    bounds = [(min(off, r * per), min(off, r * per + per)) for r in range(world)]
    arenas = [torch.zeros(4 * per) for _ in range(world)]
    sink = ScatterSink(params, slices, lambda vec, o, n, shape: vec[o:o + n].view(shape), G, arenas, 1, per, bounds, rank=0)
    first = sink.direct(0, 8, 4)                      # coordinates 0..31: rank 0's own range -> a view of ITS arena
    assert first.data_ptr() >= arenas[0].data_ptr() and first.data_ptr() < arenas[0].data_ptr() + 4 * arenas[0].numel()
    second = sink.direct(1, 16, 4)                    # coordinates 32..95 straddle ranks 0 and 1 -> staging
    third = sink.direct(2, 20, 4)                     # coordinates 96..175: ranks 1 and 2 -> staging
    for staged in (second, third):
        for arena in arenas:
            assert not (arena.data_ptr() <= staged.data_ptr() < arena.data_ptr() + 4 * arena.numel())
    second.fill_(1.0); third.fill_(2.0)
    sink.done(1); sink.done(2)
    for g in range(G):
        row = torch.cat([arenas[r][(1 + g) * per:(1 + g) * per + (hi - lo)] for r, (lo, hi) in enumerate(bounds)])
        assert torch.equal(row[32:96], torch.ones(64)) and torch.equal(row[96:176], torch.full((80,), 2.0))
