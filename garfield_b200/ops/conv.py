"""Python drivers of the tcgen05 implicit-GEMM convolutions (csrc/cuda/conv_tc.cu) and of the channels-last helper
kernels around them (csrc/cuda/nn_nhwc.cu).

All tensors are channels-last: activations `[N, H, W, C]` contiguous, weights `[Cout, kh, kw, Cin]` contiguous (OHWI —
what a `channels_last` nn.Conv2d weight is in memory). Three GEMM-shaped entry points

    conv_fprop(x, w, bias, stride, padding)                  -> y  [N, Ho, Wo, Cout]
    conv_dgrad(dy, w, in_hw, stride, padding)                -> dx [N, H, W, Cin]
    conv_wgrad(x, dy, groups, ksize, stride, padding, out=)  -> dw [G, Cout, kh, kw, Cin]   (one matrix per co-located worker)

cover every convolution and Linear layer (1×1 geometry). Layers whose operands the tiled-TMA boxes cannot address (Cin·kh·kw
not a multiple of 4 — the RGB stem, CifarNet —, output planes that do not tile into 32-pixel chunks) go through an explicit
zero-padded patch matrix and the same kernels in GEMM geometry.

Reference counterpart: cuDNN / cuBLAS behind nn.Conv2d / nn.Linear and their autograd
(/root/reference/pytorch_impl/libs/garfieldpp/worker.py:85-91, models/nets.py:40-78).
"""

from __future__ import annotations

import torch

from . import _lib

__all__ = ["conv_fprop", "conv_dgrad", "conv_wgrad", "upsample_zero", "relu", "relu_backward", "max_pool_nhwc", "max_pool_nhwc_backward", "softmax_ce_grouped",
           "colsum_grouped", "avg_pool_nhwc", "avg_pool_nhwc_backward", "out_size"]


def _pair(v):
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


def out_size(h, w, ksize, stride, padding):
    (kh, kw), (sh, sw), (ph, pw) = _pair(ksize), _pair(stride), _pair(padding)
    return (h + 2 * ph - kh) // sh + 1, (w + 2 * pw - kw) // sw + 1


def _st(t):
    return _lib.stream_ptr(t.device)


def _pitch(t):
    """Floats between consecutive pixels of a channels-last tensor [..., C] (C may be a slice of a wider pixel)."""
    for size, stride in zip(reversed(t.shape[:-1]), reversed(t.stride()[:-1])):
        if size > 1:
            return stride
    return (t.shape[-1] + 3) // 4 * 4


def _dense_pixels(t):
    """True when `t` [..., C] is laid out as consecutive pixels of one pitch (C contiguous)."""
    if t.shape[-1] > 1 and t.stride(-1) != 1:
        return False
    expect = _pitch(t)
    if expect < t.shape[-1]:
        return False
    for size, stride in zip(reversed(t.shape[:-1]), reversed(t.stride()[:-1])):
        if size > 1 and stride != expect:
            return False
        expect *= size
    return True


def _tma_ok(t):
    """Can a tiled-TMA tensor map address `t` directly? (consecutive pixels, 16-byte pitch and base)"""
    return t.dtype == torch.float32 and _dense_pixels(t) and _pitch(t) % 4 == 0 and t.data_ptr() % 16 == 0


def _prep(t):
    """`t` itself when a tensor map can address it, else a copy with a 16-byte pixel pitch."""
    if _tma_ok(t):
        return t
    t = t.float().contiguous()
    if t.shape[-1] % 4 or t.data_ptr() % 16:
        wide = torch.zeros(t.shape[:-1] + ((t.shape[-1] + 3) // 4 * 4,), device=t.device)
        wide[..., :t.shape[-1]] = t
        t = wide[..., :t.shape[-1]]
    return t


def _pixels(t):
    """`t` with consecutive pixels of one pitch (any pitch): what the SIMT gather kernels need."""
    return t if (t.dtype == torch.float32 and _dense_pixels(t)) else t.float().contiguous()


# ------------------------------------------------------------------------------------------------ explicit patch path ---

def _im2col(x, ksize, stride, padding):
    """cols [N·Ho·Wo, kpad] ((kh, kw, c) order, zero-padded to a multiple of 4 columns)."""
    (kh, kw), (sh, sw), (ph, pw) = _pair(ksize), _pair(stride), _pair(padding)
    n, h, w, c = x.shape
    ho, wo = out_size(h, w, ksize, stride, padding)
    kpad = (kh * kw * c + 3) // 4 * 4
    cols = torch.empty((n * ho * wo, kpad), device=x.device)
    _lib.check(_lib.require().gb_im2col_pad_nhwc(x.data_ptr(), cols.data_ptr(), n, h, w, c, _pitch(x), kh, kw, sh, sw, ph, pw, kpad, _st(x)))
    _lib.count()
    return cols, ho, wo


def _pad_weight(w2d, kpad):
    """[Cout, K] → [Cout, kpad] zero-padded copy (row stride a multiple of 16 bytes)."""
    out = torch.empty((w2d.shape[0], kpad), device=w2d.device)
    _lib.check(_lib.require().gb_pad_rows(w2d.data_ptr(), out.data_ptr(), w2d.shape[0], w2d.shape[1], w2d.stride(0), kpad, _st(w2d)))
    _lib.count()
    return out


def _implicit_ok(cin, kh, kw, wo):
    return (kh * kw * cin) % 4 == 0 and kh * kw <= 32 and 1 <= wo <= 128


# --------------------------------------------------------------------------------------------------------- the three ops ---

def conv_fprop(x, w, bias=None, stride=1, padding=0, relu=False, out=None, cache=None):
    """y = conv(x, w) (+ bias) (ReLU). x [N, H, W, Cin], w [Cout, kh, kw, Cin] → y [N, Ho, Wo, Cout].
    `cache`: a dict that receives the explicit patch matrix (key "cols") when this layer needs one, for `conv_wgrad(cols=)`."""
    lib = _lib.require()
    w = w if w.is_contiguous() else w.contiguous()
    n, h, wd, cin = x.shape
    cout, kh, kw, _ = w.shape
    (sh, sw), (ph, pw) = _pair(stride), _pair(padding)
    ho, wo = out_size(h, wd, (kh, kw), stride, padding)
    y = out if out is not None else torch.empty((n, ho, wo, cout), device=x.device)
    b = bias.data_ptr() if bias is not None else None
    implicit = _implicit_ok(cin, kh, kw, wo) and max(sh, sw) <= 8 and w.data_ptr() % 16 == 0
    x = _prep(x) if implicit else _pixels(x)
    with torch.cuda.device(x.device):
        if implicit:
            _lib.check(lib.gb_conv_fprop_tc(x.data_ptr(), w.data_ptr(), b, y.data_ptr(), n, h, wd, cin, _pitch(x), cout, _pitch(y), kh, kw, sh, sw,
                                            ph, pw, int(relu), _st(x)))
            _lib.count()
        else:
            cols, _, _ = _im2col(x, (kh, kw), stride, padding)
            if cache is not None:
                cache["cols"] = cols
            w2 = cache.get("w2") if cache is not None else None      # operand emitted by the model-pull kernel (MSMW path 3)
            if w2 is None or tuple(w2.shape) != (cout, cols.shape[1]):
                w2 = _pad_weight(w.view(cout, -1), cols.shape[1])
            _lib.check(lib.gb_conv_fprop_tc(cols.data_ptr(), w2.data_ptr(), b, y.data_ptr(), cols.shape[0], 1, 1, cols.shape[1], cols.shape[1], cout,
                                            _pitch(y), 1, 1, 1, 1, 0, 0, int(relu), _st(x)))
            _lib.count()
    return y


def upsample_zero(t, stride, out_hw):
    """out[n, h·s, w·s, :] = t[n, h, w, :], zero elsewhere."""
    (sh, sw) = _pair(stride)
    n, ho, wo, c = t.shape
    t = t if t.is_contiguous() else t.contiguous()
    out = torch.empty((n, out_hw[0], out_hw[1], c), device=t.device)
    _lib.check(_lib.require().gb_upsample_zero_nhwc(t.data_ptr(), out.data_ptr(), n, ho, wo, c, sh, sw, out_hw[0], out_hw[1], _st(t)))
    _lib.count()
    return out


def conv_dgrad(dy, w, in_hw, stride=1, padding=0, addend=None):
    """dx [N, H, W, Cin] of y = conv(x, w); dy [N, Ho, Wo, Cout], w [Cout, kh, kw, Cin]. `addend` (dx's shape, contiguous): a
    second gradient of the same tensor (skip connection), summed in the kernel's epilogue where the path allows it."""
    lib = _lib.require()
    dy, w = _prep(dy), w if w.is_contiguous() else w.contiguous()
    n, ho, wo, cout = dy.shape
    _, kh, kw, cin = w.shape
    (sh, sw), (ph, pw) = _pair(stride), _pair(padding)
    h, wd = in_hw
    dx = torch.empty((n, h, wd, cin), device=dy.device)
    fused_add = addend is not None and (sh, sw) == (1, 1) and addend.is_contiguous() and tuple(addend.shape) == tuple(dx.shape)
    result = _conv_dgrad(lib, dy, w, dx, n, ho, wo, cout, kh, kw, cin, sh, sw, ph, pw, h, wd, addend if fused_add else None)
    if addend is not None and not fused_add:
        result = result + addend
    return result


def _conv_dgrad(lib, dy, w, dx, n, ho, wo, cout, kh, kw, cin, sh, sw, ph, pw, h, wd, addend):
    with torch.cuda.device(dy.device):
        if (kh * kw * cin) % 4 == 0 and kh * kw <= 32 and wd <= 128 and w.data_ptr() % 16 == 0 and (sh == sw == 1 or cout % 4 == 0):
            if (kh, kw) == (1, 1) and (ph, pw) == (0, 0) and (sh, sw) != (1, 1):
                # strided 1×1: the GEMM at output resolution, scattered onto the sampled pixels
                small = torch.empty((n, ho, wo, cin), device=dy.device)
                _lib.check(lib.gb_conv_dgrad_tc(dy.data_ptr(), w.data_ptr(), small.data_ptr(), n, ho, wo, cin, cin, cout, _pitch(dy), 1, 1, 0, 0, None, _st(dy)))
                _lib.count()
                if cin % 4 == 0:
                    return upsample_zero(small, (sh, sw), (h, wd))
                dx.zero_()
                dx[:, ::sh, ::sw] = small
                return dx
            if (sh, sw) != (1, 1):
                # strided k×k: stride-1 data gradient of the zero-upsampled output gradient
                full_h, full_w = h + 2 * ph - kh + 1, wd + 2 * pw - kw + 1
                dy = upsample_zero(dy if dy.is_contiguous() else dy.contiguous(), (sh, sw), (full_h, full_w))
            _lib.check(lib.gb_conv_dgrad_tc(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), n, h, wd, cin, cin, cout, _pitch(dy), kh, kw, ph, pw,
                                            addend.data_ptr() if addend is not None else None, _st(dy)))
            _lib.count()
            addend = None
        else:
            # explicit: dcols[pix, (kh, kw, c)] = dy · W2d (a 1×1 data-gradient GEMM on zero-padded weights), then col2im
            k = kh * kw * cin
            kpad = (k + 3) // 4 * 4
            w2 = _pad_weight(w.view(cout, k), kpad)
            m = n * ho * wo
            dcols = torch.empty((m, kpad), device=dy.device)
            _lib.check(lib.gb_conv_dgrad_tc(dy.data_ptr(), w2.data_ptr(), dcols.data_ptr(), m, 1, 1, kpad, kpad, cout, _pitch(dy), 1, 1, 0, 0, None, _st(dy)))
            _lib.check(lib.gb_col2im_nhwc(dcols.data_ptr(), dx.data_ptr(), n, h, wd, cin, cin, kh, kw, sh, sw, ph, pw, kpad, _st(dy)))
            _lib.count(2)
    return dx if addend is None else dx + addend


def _wgrad_implicit_ok(ho, wo):
    hw = ho * wo
    return wo <= 32 and 32 % wo == 0 and (hw % 32 == 0 or 32 % hw == 0)


def conv_wgrad(x, dy, groups, ksize, stride=1, padding=0, out=None, cols=None):
    """Per-worker weight gradients dw [G, Cout, kh, kw, Cin]: images g·B … g·B+B−1 of x / dy belong to worker g.
    `out`: an optional [G, Cout, kh·kw·Cin] destination whose matrices are row-major with any group stride (a strided view of
    the gradient rows). `cols`: the patch matrix the forward pass already built for this layer (explicit path)."""
    lib = _lib.require()
    dy = _prep(dy)
    n, h, wd, cin = x.shape
    _, ho, wo, cout = dy.shape
    (kh, kw), (sh, sw), (ph, pw) = _pair(ksize), _pair(stride), _pair(padding)
    implicit = kh * kw <= 32 and max(sh, sw) <= 8 and _wgrad_implicit_ok(ho, wo) and (_tma_ok(x) or cin % 4 == 0)
    x = _prep(x) if implicit else _pixels(x)
    k = kh * kw * cin
    if n % groups:
        raise ValueError("batch %d is not a multiple of %d groups" % (n, groups))
    b = n // groups
    if out is None:
        out = torch.empty((groups, cout, k), device=x.device)
    if out.stride(2) != 1 or out.stride(1) < k:
        raise ValueError("conv_wgrad: destination matrices must be row-major")
    gstride = out.stride(0) if groups > 1 else cout * out.stride(1)
    with torch.cuda.device(x.device):
        if implicit:
            _lib.check(lib.gb_conv_wgrad_tc(x.data_ptr(), dy.data_ptr(), out.data_ptr(), groups, b, h, wd, cin, _pitch(x), cout, _pitch(dy), kh, kw,
                                            sh, sw, ph, pw, cin, out.stride(1), gstride, _st(x)))
            _lib.count()
        else:
            if cols is None:
                cols, _, _ = _im2col(x, (kh, kw), stride, padding)
            _lib.check(lib.gb_conv_wgrad_tc(cols.data_ptr(), dy.data_ptr(), out.data_ptr(), groups, b * ho * wo, 1, 1, cols.shape[1], cols.shape[1], cout,
                                            _pitch(dy), 1, 1, 1, 1, 0, 0, k, out.stride(1), gstride, _st(x)))
            _lib.count()
    return out


# ----------------------------------------------------------------------------------------------------- non-GEMM layers ---

def relu(x):
    """max(x, 0) on a dense tensor of any shape (same strides in, same strides out)."""
    y = torch.empty_like(x)
    _lib.check(_lib.require().gb_relu_forward(x.data_ptr(), y.data_ptr(), x.numel(), _st(x)))
    _lib.count()
    return y


def add(a, b):
    """a + b for two dense tensors of the same shape AND element order (a new tensor with a's strides)."""
    out = torch.empty_like(a)
    _lib.check(_lib.require().gb_add(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _st(a)))
    _lib.count()
    return out


def relu_backward(dy, y):
    """dy where y > 0, else 0; `dy` must have y's element order."""
    dx = torch.empty_like(y)
    _lib.check(_lib.require().gb_relu_backward(dy.data_ptr(), y.data_ptr(), dx.data_ptr(), y.numel(), _st(y)))
    _lib.count()
    return dx


def max_pool_nhwc(x, kernel, stride, padding, out=None):
    n, h, w, c = x.shape
    ho, wo = out_size(h, w, kernel, stride, padding)
    x = x if x.is_contiguous() else x.contiguous()
    y = out if out is not None else torch.empty((n, ho, wo, c), device=x.device)
    arg = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device)
    _lib.check(_lib.require().gb_maxpool_nhwc_forward(x.data_ptr(), y.data_ptr(), arg.data_ptr(), n, h, w, c, kernel, stride, padding, _st(x)))
    _lib.count()
    return y, arg


def max_pool_nhwc_backward(dy, arg, in_shape, kernel, stride, padding):
    n, h, w, c = in_shape
    dy = dy if dy.is_contiguous() else dy.contiguous()
    dx = torch.empty(in_shape, device=dy.device)
    _lib.check(_lib.require().gb_maxpool_nhwc_backward(dy.data_ptr(), arg.data_ptr(), dx.data_ptr(), n, h, w, c, kernel, stride, padding, _st(dy)))
    _lib.count()
    return dx


def avg_pool_nhwc(x, out=None):
    n, h, w, c = x.shape
    x = x if x.is_contiguous() else x.contiguous()
    y = out if out is not None else torch.empty((n, c), device=x.device)
    _lib.check(_lib.require().gb_avgpool_nhwc_forward(x.data_ptr(), y.data_ptr(), n, h * w, c, _st(x)))
    _lib.count()
    return y


def avg_pool_nhwc_backward(dy, hw):
    n, c = dy.shape
    dy = dy if dy.is_contiguous() else dy.contiguous()
    dx = torch.empty((n, hw[0], hw[1], c), device=dy.device)
    _lib.check(_lib.require().gb_avgpool_nhwc_backward(dy.data_ptr(), dx.data_ptr(), n, hw[0] * hw[1], c, _st(dy)))
    _lib.count()
    return dx


def softmax_ce_grouped(logits, target, groups):
    """Cross-entropy of G workers' logits in one launch: returns (row_loss [rows], dlogits [rows, pad4(classes)]) where
    dlogits is the gradient of Σ_g mean_g(row losses) — every sample's logit gradient is its own worker's."""
    rows, classes = logits.shape
    ldd = (classes + 3) // 4 * 4
    row_loss = torch.empty(rows, device=logits.device)
    dlogits = torch.empty((rows, ldd), device=logits.device)
    target = target if target.dtype == torch.int64 else target.long()
    _lib.check(_lib.require().gb_softmax_ce_grouped(logits.data_ptr(), logits.stride(0), target.data_ptr(), row_loss.data_ptr(), dlogits.data_ptr(),
                                                   ldd, rows, classes, rows // groups, _st(logits)))
    _lib.count()
    return row_loss, dlogits


def colsum_grouped(t, groups, cols=None, scale=1.0, out=None):
    """out[g, c] = scale · Σ over the rows of group g of t[row, c] (t: [rows, ≥cols], row pitch stride(0))."""
    rows = t.shape[0]
    cols = t.shape[1] if cols is None else cols
    if out is None:
        out = torch.empty((groups, cols), device=t.device)
    _lib.check(_lib.require().gb_colsum_grouped(t.data_ptr(), t.stride(0), out.data_ptr(), groups, rows // groups, cols, float(scale), _st(t)))
    _lib.count()
    return out
