"""Worker fusion: ONE forward/backward pass computes the gradients of all G workers hosted by a GPU.

The reference runs every worker as its own process with its own pickled model copy (garfieldpp/worker.py:77-96);
the first device engine here ran one CUDA-graphed chain per worker on its own stream. But co-located workers of one
step all evaluate the SAME parameters, so everything except the weight gradients can be batched:

* forward and data gradients run once over the concatenated batch [G·B, …] (same weights → one convolution / GEMM
  instead of G small ones, weights read once);
* BatchNorm keeps per-worker statistics: `ops.nn.grouped_batch_norm` normalises each of the G slices with its own
  mean/variance/running buffers in one launch (thread-block clusters + DSMEM reduction, csrc/cuda/bn_nhwc.cu);
* weight gradients stay per worker (that is what the aggregation rule consumes): 1×1 convolutions and Linear
  layers become ONE batched GEMM [G, out, B·HW] × [G, B·HW, in]; k×k convolutions run G weight-gradient kernels on
  side streams, off the data-gradient chain;
* the loss is the sum of the G per-worker mean losses, so every sample's logit gradient is exactly its own
  worker's; each worker's gradient lands in its own row of the gradient matrix through ONE multi-tensor scatter.

The result is the same G gradients as G separate replicas (tests/test_grouped.py), not a large-batch gradient.
"""

from __future__ import annotations

import ctypes
import os
import types

import torch
import torch.nn.functional as F

from ..ops import _lib, conv as convops
from ..ops.nn import gemm_nt, grouped_batch_norm, im2col_nhwc

__all__ = ["GroupedWorkers", "ScatterSink"]

_aten = torch.ops.aten

# 1×1 convolutions forward on our tcgen05 GEMM instead of the library convolution. Opt-in: at the flagship's shapes
# (M = 200·HW ≤ 12 800, K ≤ 2048: one tile per CTA, so set-up and epilogue are exposed) it measured 2.53 ms per
# gradient pass against 2.34 ms with the library kernels (profiles/README.md §3); it needs a small-tile variant
# before it can be the default.
POINTWISE_GEMM = [os.environ.get("GARFIELD_POINTWISE_GEMM", "0") != "0"]

# Convolutions, Linear layers, pooling and the loss of the fused pass on this library's own sm_100a kernels (tcgen05
# implicit GEMMs, ops/conv.py) — the default on CUDA with channels-last replicas. GARFIELD_OWN_CONV=0 switches back to the
# library (cuDNN / cuBLAS) kernels for A/B measurements.
OWN_KERNELS = [os.environ.get("GARFIELD_OWN_CONV", "1") != "0"]


def _same_layout(a, b):
    """Same shape and same element order in memory (strides of size-1 dimensions are irrelevant)."""
    if a.shape != b.shape:
        return False
    return all(sa == sb for n, sa, sb in zip(a.shape, a.stride(), b.stride()) if n > 1)


class _GradSink:
    """Per-group gradients of every parameter, scattered into the G gradient rows after the backward pass."""

    def __init__(self, params, slices, rows, view_of):
        self.params, self.slices, self.rows, self.view_of = params, slices, rows, view_of
        self.groups = len(rows)
        self.pending = {}
        base = rows[0]
        self.row_offsets = [(r.data_ptr() - base.data_ptr()) // 4 for r in rows]
        self.contiguous_rows = base.is_cuda and all(o >= 0 for o in self.row_offsets)
        pitch = self.row_offsets[1] if self.groups > 1 else base.numel()
        self.pitch = pitch if self.contiguous_rows and all(o == g * pitch for g, o in enumerate(self.row_offsets)) else None

    DONE = object()   # marker: the gradient was written straight into the rows

    def direct(self, index, rows, cols):
        """[G, rows, cols] strided view of parameter `index` inside the G gradient rows (row-major matrix layout), so a
        batched GEMM can write the per-worker gradients in place; None when the rows are not uniformly pitched."""
        if self.pitch is None:
            return None
        off, n, _ = self.slices[index]
        base = self.rows[0]
        return torch.as_strided(base, (self.groups, rows, cols), (self.pitch, cols, 1), base.storage_offset() + off)

    def put(self, index, grads):
        """grads: a [G, *shape] tensor or a list of G tensors of the parameter's shape."""
        self.pending[index] = grads

    def can_direct(self, index):
        return self.pitch is not None

    def done(self, index):
        """The kernel writing `direct(index, …)` has been enqueued: nothing is left to scatter for this parameter."""
        self.pending[index] = self.DONE

    def flush(self, first_index=0):
        """Scatter the pending gradients of parameters `first_index…` (everything by default) into the rows; a partial
        flush marks them done so the final one skips them."""
        srcs, offs, sizes = [], [], []
        for index, (p, (off, n, shape)) in enumerate(zip(self.params, self.slices)):
            if index < first_index:
                continue
            grads = self.pending.get(index)
            if grads is self.DONE:
                continue
            if first_index:
                self.pending[index] = self.DONE
            for g in range(self.groups):
                row = self.rows[g]
                if grads is None:
                    row[off:off + n].zero_()
                    continue
                t = grads[g]
                if t.shape != p.shape:
                    t = t.reshape(p.shape)
                if self.contiguous_rows and t.is_cuda and t.dtype == torch.float32 and _same_layout(t, p):
                    srcs.append(t.data_ptr()); offs.append(self.row_offsets[g] + off); sizes.append(n)
                else:
                    self.view_of(row, off, n, shape).copy_(t)
        if srcs:
            k = len(srcs)
            base = self.rows[0]
            _lib.check(_lib.require().gb_multi_copy((ctypes.c_void_p * k)(*srcs), (ctypes.c_uint64 * k)(*offs), (ctypes.c_uint32 * k)(*sizes), k,
                                                    base.data_ptr(), _lib.stream_ptr(base.device)))
            _lib.count((k + 111) // 112)
        if not first_index:
            self.pending = {}


class ScatterSink:
    """Gradient sink for the coordinate-sharded server: the gradient of worker g (arena row `first_row + g`) ends up,
    slice by slice, in the arena of the rank that AGGREGATES that coordinate range — a reduce-scatter by construction.
    `arenas[r]` is rank r's peer-mapped arena ([rows, per] floats, row pitch `per`), `bounds[r]` the coordinate range
    [lo, hi) rank r owns, `rank` this process's rank.

    Library kernels (the batched weight-gradient GEMMs) only ever write THIS GPU's memory: `direct()` hands them a view of
    the local arena when this rank owns the whole parameter, otherwise a persistent local staging tensor. `done()` then
    moves the staged slices to their owners with our own multi-tensor copy kernel (plain `st.global` on the peer-mapped
    pointers), on the stream that produced them, so the NVLink transfer overlaps the rest of the backward pass.
    BatchNorm parameters, biases and anything else `put()` here follow in one multi-tensor copy per destination rank."""

    DONE = _GradSink.DONE

    def __init__(self, params, slices, view_of, groups, arenas, first_row, per, bounds, rank=0):
        self.params, self.slices, self.view_of, self.groups = params, slices, view_of, groups
        self.arenas, self.first_row, self.per, self.bounds, self.rank = arenas, first_row, per, bounds, rank
        self.pending = {}
        self._pieces = [self._owners(off, n) for off, n, _ in slices]
        self._staging = {}

    def _owners(self, off, n):
        pieces = []
        for r, (lo, hi) in enumerate(self.bounds):
            a, b = max(off, lo), min(off + n, hi)
            if a < b:
                pieces.append((r, a, b))
        return pieces

    def put(self, index, grads):
        self.pending[index] = grads

    def is_local(self, tensor):
        """True when `tensor` lies in memory of THIS GPU (the local arena or an ordinary torch allocation) — what every
        operand of a library kernel must satisfy."""
        if not tensor.is_cuda:
            return True
        ptr = tensor.data_ptr()
        for r, arena in enumerate(self.arenas):
            if r != self.rank and arena.is_cuda and arena.data_ptr() <= ptr < arena.data_ptr() + 4 * arena.numel():
                return False
        return True

    def can_direct(self, index):
        return True

    def direct(self, index, rows, cols):
        """A LOCAL [G, rows, cols] destination for the per-worker gradients of parameter `index`: the arena rows
        themselves when this rank aggregates the whole parameter, else a staging tensor that `done()` pushes."""
        pieces = self._pieces[index]
        if len(pieces) == 1 and pieces[0][0] == self.rank:
            a = pieces[0][1]
            base = self.arenas[self.rank]
            out = torch.as_strided(base, (self.groups, rows, cols), (self.per, cols, 1),
                                   base.storage_offset() + self.first_row * self.per + (a - self.bounds[self.rank][0]))
        else:
            out = self._staging.get(index)
            if out is None or out.shape != (self.groups, rows, cols):
                out = self._staging[index] = torch.empty((self.groups, rows, cols), device=self.params[index].device)
        assert self.is_local(out), "library kernels must only write this GPU's memory"
        return out

    def done(self, index):
        """The GEMM writing `direct(index, …)` has been enqueued on the current stream: push staged slices to the ranks
        that aggregate them (same stream) and mark the parameter complete."""
        staged = self._staging.get(index)
        pieces = self._pieces[index]
        if staged is not None and not (len(pieces) == 1 and pieces[0][0] == self.rank):
            off, n, _ = self.slices[index]
            table = {}
            for g in range(self.groups if not staged.is_cuda else 0):      # host arenas (unit tests of the slicing logic)
                for r, a, b in pieces:
                    at = (self.first_row + g) * self.per + (a - self.bounds[r][0])
                    self.arenas[r][at:at + (b - a)].copy_(staged[g].reshape(-1)[a - off:b - off])
            for g in range(self.groups if staged.is_cuda else 0):
                src = staged.data_ptr() + 4 * g * n
                for r, a, b in pieces:
                    entry = table.setdefault(r, ([], [], []))
                    entry[0].append(src + 4 * (a - off))
                    entry[1].append((self.first_row + g) * self.per + (a - self.bounds[r][0]))
                    entry[2].append(b - a)
            self._push(table)
        self.pending[index] = self.DONE

    def _push(self, table):
        """One launch of our multi-tensor copy kernel per destination rank (its table lives in the kernel parameters)."""
        lib = _lib.require() if table else None
        for r, (srcs, offs, sizes) in table.items():
            k = len(srcs)
            base = self.arenas[r]
            _lib.check(lib.gb_multi_copy((ctypes.c_void_p * k)(*srcs), (ctypes.c_uint64 * k)(*offs), (ctypes.c_uint32 * k)(*sizes), k,
                                         base.data_ptr(), _lib.stream_ptr(base.device)))
            _lib.count((k + 111) // 112)

    def flush(self):
        table = {}
        keep = []          # staging copies must outlive the launches below
        for index, (p, (off, n, shape)) in enumerate(zip(self.params, self.slices)):
            grads = self.pending.get(index)
            if grads is self.DONE:
                continue
            for g in range(self.groups):
                src = None
                if grads is not None:
                    t = grads[g]
                    if t.shape != p.shape:
                        t = t.reshape(p.shape)
                    if not (t.dtype == torch.float32 and _same_layout(t, p) and t.is_cuda):
                        staged = torch.empty(n, device=p.device)           # the parameter's storage order
                        self.view_of(staged, 0, n, shape).copy_(t)
                        t = staged
                    keep.append(t)
                    src = t.data_ptr()
                for r, a, b in self._pieces[index]:
                    dst_off = (self.first_row + g) * self.per + (a - self.bounds[r][0])
                    if not self.arenas[r].is_cuda:      # host arenas (unit tests of the slicing logic)
                        if src is None:
                            self.arenas[r][dst_off:dst_off + (b - a)].zero_()
                        else:
                            self.arenas[r][dst_off:dst_off + (b - a)].copy_(t.reshape(-1)[a - off:b - off])
                    else:
                        if src is None:                  # no gradient: push zeros with the same kernel
                            zeros = self._zeros(b - a, p.device)
                            keep.append(zeros)
                            entry = table.setdefault(r, ([], [], []))
                            entry[0].append(zeros.data_ptr()); entry[1].append(dst_off); entry[2].append(b - a)
                            continue
                        entry = table.setdefault(r, ([], [], []))
                        entry[0].append(src + 4 * (a - off)); entry[1].append(dst_off); entry[2].append(b - a)
        self._push(table)
        self.pending = {}

    def _zeros(self, n, device):
        z = getattr(self, "_zero_buf", None)
        if z is None or z.numel() < n:
            z = self._zero_buf = torch.zeros(n, device=device)
        return z[:n]


class _GroupedConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, owner, conv, widx, bidx):
        ctx.save_for_backward(x, w)
        ctx.cfg = (owner, conv, widx, bidx)
        if (POINTWISE_GEMM[0] and conv.kernel_size == (1, 1) and conv.stride == (1, 1) and conv.padding == (0, 0) and conv.groups == 1
                and x.is_cuda and x.dtype == torch.float32 and x.shape[1] % 4 == 0 and torch.backends.cudnn.allow_tf32
                and x.is_contiguous(memory_format=torch.channels_last) and _lib.available()):
            # a stride-1 1×1 convolution over NHWC activations IS the GEMM y[M, Cout] = x[M, Cin] · W[Cout, Cin]ᵀ:
            # run it on our persistent tcgen05 kernel (TF32, like the library convolution it replaces)
            n, cin, h, wd = x.shape
            ym = gemm_nt(x.permute(0, 2, 3, 1).reshape(-1, cin), w.reshape(w.shape[0], cin), bias=b)
            return ym.view(n, h, wd, w.shape[0]).permute(0, 3, 1, 2)
        return F.conv2d(x, w, b, conv.stride, conv.padding, conv.dilation, conv.groups)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        owner, conv, widx, bidx = ctx.cfg
        G = owner.groups
        B = x.shape[0] // G
        stride, padding, dilation = list(conv.stride), list(conv.padding), list(conv.dilation)
        pointwise = (conv.kernel_size == (1, 1) and conv.groups == 1 and conv.padding == (0, 0) and x.is_cuda
                     and x.is_contiguous(memory_format=torch.channels_last))
        gx = None
        if pointwise:
            # per-worker weight gradients of a 1×1 convolution = one batched GEMM over the NHWC views, written straight
            # into the gradient rows, on a side stream (off the data-gradient chain)
            if not gy.is_contiguous(memory_format=torch.channels_last):
                gy = gy.contiguous(memory_format=torch.channels_last)
            strided = conv.stride != (1, 1)
            xs = x[:, :, ::conv.stride[0], ::conv.stride[1]].contiguous(memory_format=torch.channels_last) if strided else x
            cin, cout = xs.shape[1], gy.shape[1]
            xm = xs.permute(0, 2, 3, 1).reshape(G, -1, cin)
            gm = gy.permute(0, 2, 3, 1).reshape(G, -1, cout)
            side = owner.next_side()
            with owner.tf32_matmul():
                if side is not None:
                    side.wait_stream(torch.cuda.current_stream(x.device))
                    xs.record_stream(side); gy.record_stream(side)
                with torch.cuda.stream(side) if side is not None else _NoStream():
                    out = owner.sink.direct(widx, cout, cin)
                    if out is not None:
                        torch.bmm(gm.transpose(1, 2), xm, out=out)
                        owner.sink.done(widx)
                    else:
                        owner.sink.put(widx, torch.bmm(gm.transpose(1, 2), xm).unsqueeze(-1).unsqueeze(-1))
                if strided and ctx.needs_input_grad[0]:
                    # data gradient of a strided 1×1 convolution: one GEMM scattered onto the sampled pixels (the library's
                    # strided-dgrad kernels take 50-150 µs on these shapes)
                    dxs = gm.reshape(-1, cout) @ w.reshape(cout, cin)
                    gx = torch.zeros_like(x)
                    gx[:, :, ::conv.stride[0], ::conv.stride[1]] = dxs.view(xs.shape[0], xs.shape[2], xs.shape[3], cin).permute(0, 3, 1, 2)
        elif (x.is_cuda and owner.replica.channels_last and conv.groups == 1 and conv.dilation == (1, 1) and x.shape[1] % 4 == 0
              and x.is_contiguous(memory_format=torch.channels_last) and owner.sink.can_direct(widx)):
            # k×k convolution: gather the patches once (own im2col kernel, (kh, kw, c) order), then ONE batched GEMM
            # dW[g] = dY[g]ᵀ · cols[g] writes all workers' gradients in OHWI order straight into their rows
            if not gy.is_contiguous(memory_format=torch.channels_last):
                gy = gy.contiguous(memory_format=torch.channels_last)
            cout, k = gy.shape[1], conv.kernel_size[0] * conv.kernel_size[1] * x.shape[1]
            side = owner.next_side()
            if side is not None:
                side.wait_stream(torch.cuda.current_stream(x.device))
                x.record_stream(side); gy.record_stream(side)
            with owner.tf32_matmul(), (torch.cuda.stream(side) if side is not None else _NoStream()):
                cols = im2col_nhwc(x, conv.kernel_size, conv.stride, conv.padding).view(G, -1, k)
                gm = gy.permute(0, 2, 3, 1).reshape(G, -1, cout)
                torch.bmm(gm.transpose(1, 2), cols, out=owner.sink.direct(widx, cout, k))
                owner.sink.done(widx)
        else:
            sides = owner.all_sides()
            cur = torch.cuda.current_stream(x.device) if sides else None
            if sides:
                for s in sides:
                    s.wait_stream(cur)
                for s in sides:
                    gy.record_stream(s); x.record_stream(s)
            grads = []
            for g in range(G):
                xs, gs = x[g * B:(g + 1) * B], gy[g * B:(g + 1) * B]
                if sides:
                    with torch.cuda.stream(sides[g % len(sides)]):
                        grads.append(_aten.convolution_backward(gs, xs, w, None, stride, padding, dilation, False, [0, 0], conv.groups,
                                                                [False, True, False])[1])
                else:
                    grads.append(_aten.convolution_backward(gs, xs, w, None, stride, padding, dilation, False, [0, 0], conv.groups,
                                                            [False, True, False])[1])
            owner.sink.put(widx, grads)
        if bidx is not None:
            owner.sink.put(bidx, gy.reshape(G, B, *gy.shape[1:]).sum((1, 3, 4)))
        if gx is None and ctx.needs_input_grad[0]:
            if (x.is_cuda and max(conv.stride) > 1 and conv.dilation == (1, 1) and conv.groups == 1
                    and all(2 * p == k - 1 for p, k in zip(conv.padding, conv.kernel_size))):
                # strided k×k data gradient = stride-1 data gradient of the zero-upsampled output gradient ("same"
                # padding keeps the sizes equal); the library's strided-dgrad kernels are 5-10x slower on these shapes
                up = torch.empty((gy.shape[0], gy.shape[1], x.shape[2], x.shape[3]), device=gy.device, dtype=gy.dtype,
                                 memory_format=torch.channels_last if x.is_contiguous(memory_format=torch.channels_last) else torch.contiguous_format).zero_()
                up[:, :, ::conv.stride[0], ::conv.stride[1]] = gy
                gx = _aten.convolution_backward(up, x, w, None, [1, 1], padding, dilation, False, [0, 0], 1, [True, False, False])[0]
            else:
                gx = _aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0], conv.groups, [True, False, False])[0]
        return gx, None, None, None, None, None, None


class _GroupedLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, owner, widx, bidx):
        ctx.save_for_backward(x, w)
        ctx.cfg = (owner, widx, bidx)
        return F.linear(x, w, b)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        owner, widx, bidx = ctx.cfg
        G = owner.groups
        xm = x.reshape(G, -1, x.shape[-1])
        gm = gy.reshape(G, -1, gy.shape[-1])
        out = owner.sink.direct(widx, gm.shape[-1], xm.shape[-1])
        if out is not None:
            torch.bmm(gm.transpose(1, 2), xm, out=out)
            owner.sink.done(widx)
        else:
            owner.sink.put(widx, torch.bmm(gm.transpose(1, 2), xm))
        if bidx is not None:
            owner.sink.put(bidx, gm.sum(1))
        gx = gy @ w if ctx.needs_input_grad[0] else None
        return gx, None, None, None, None, None


class _TcConv(torch.autograd.Function):
    """Conv2d of the fused pass on the tcgen05 implicit-GEMM kernels: forward, data gradient (once over the concatenated
    batch) and the G per-worker weight gradients (one launch, written straight into the workers' gradient rows)."""

    @staticmethod
    def forward(ctx, x, w, b, owner, conv, widx, bidx, fork=False):
        """`fork`: also return the input itself (the identity branch of a residual block) — its gradient then arrives here and is
        summed inside the data-gradient kernel's epilogue instead of by a separate add kernel."""
        skip = x.view_as(x) if fork else None
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        ctx.save_for_backward(x, w)
        ctx.cfg = (owner, conv, widx, bidx)
        ctx.fork = fork
        cache = {"w2": getattr(conv, "_gb_w2", None)}
        ho, wo = convops.out_size(x.shape[2], x.shape[3], conv.kernel_size, conv.stride, conv.padding)
        # the output is allocated here in its logical NCHW shape (channels-last strides) so that autograd sees a fresh tensor,
        # not a view — layers after it may work in place
        y = torch.empty((x.shape[0], w.shape[0], ho, wo), device=x.device, memory_format=torch.channels_last)
        convops.conv_fprop(x.permute(0, 2, 3, 1), w.permute(0, 2, 3, 1), b, conv.stride, conv.padding, out=y.permute(0, 2, 3, 1), cache=cache)
        ctx.cols = cache.get("cols")       # explicit patch matrix (RGB stem): reused by the weight gradient
        return (y, skip) if fork else y

    @staticmethod
    def backward(ctx, gy, gskip=None):
        x, w = ctx.saved_tensors
        owner, conv, widx, bidx = ctx.cfg
        G = owner.groups
        if not gy.is_contiguous(memory_format=torch.channels_last):
            gy = gy.contiguous(memory_format=torch.channels_last)
        xn, gyn, wn = x.permute(0, 2, 3, 1), gy.permute(0, 2, 3, 1), w.permute(0, 2, 3, 1)
        cout, kh, kw, cin = wn.shape
        side = owner.next_side()
        if side is not None:
            side.wait_stream(torch.cuda.current_stream(x.device))
            x.record_stream(side); gy.record_stream(side)
        with torch.cuda.stream(side) if side is not None else _NoStream():
            if owner.sink.can_direct(widx):
                convops.conv_wgrad(xn, gyn, G, (kh, kw), conv.stride, conv.padding, out=owner.sink.direct(widx, cout, kh * kw * cin), cols=ctx.cols)
                owner.sink.done(widx)
            else:
                dw = convops.conv_wgrad(xn, gyn, G, (kh, kw), conv.stride, conv.padding, cols=ctx.cols)
                owner.sink.put(widx, dw.view(G, cout, kh, kw, cin).permute(0, 1, 4, 2, 3))
            if bidx is not None:
                owner.sink.put(bidx, convops.colsum_grouped(gyn.reshape(-1, cout), G))
        gx = None
        if ctx.needs_input_grad[0]:
            addend = None
            if gskip is not None:
                if not gskip.is_contiguous(memory_format=torch.channels_last):
                    gskip = gskip.contiguous(memory_format=torch.channels_last)
                addend = gskip.permute(0, 2, 3, 1)
            gx = convops.conv_dgrad(gyn, wn, (x.shape[2], x.shape[3]), conv.stride, conv.padding, addend=addend).permute(0, 3, 1, 2)
        return gx, None, None, None, None, None, None, None


class _TcSplit(torch.autograd.Function):
    """The point where a residual block's input forks (first convolution / identity): two views forward; backward sums the two
    gradients with our own kernel, which takes part in the programmatic launch chain (autograd's accumulation does not)."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x), x.view_as(x)

    @staticmethod
    def backward(ctx, ga, gb):
        if ga is None or gb is None:
            return ga if gb is None else gb
        same = ga.shape == gb.shape and ga.stride() == gb.stride() and ga.dtype == gb.dtype == torch.float32 and ga.is_cuda
        dense = same and (ga.is_contiguous() or (ga.dim() == 4 and ga.is_contiguous(memory_format=torch.channels_last)))
        return convops.add(ga, gb) if dense else ga + gb


class _TcLinear(torch.autograd.Function):
    """Linear layer of the fused pass as a 1×1-geometry GEMM on the same kernels."""

    @staticmethod
    def forward(ctx, x, w, b, owner, widx, bidx):
        ctx.save_for_backward(x, w)
        ctx.cfg = (owner, widx, bidx)
        y = torch.empty((x.shape[0], w.shape[0]), device=x.device)
        convops.conv_fprop(x.unsqueeze(1).unsqueeze(1), w.unsqueeze(1).unsqueeze(1), b, out=y.unsqueeze(1).unsqueeze(1))
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        owner, widx, bidx = ctx.cfg
        G = owner.groups
        xn, gyn, wn = x.unsqueeze(1).unsqueeze(1), gy.unsqueeze(1).unsqueeze(1), w.unsqueeze(1).unsqueeze(1)
        nout, nin = w.shape
        if owner.sink.can_direct(widx):
            convops.conv_wgrad(xn, gyn, G, (1, 1), out=owner.sink.direct(widx, nout, nin))
            owner.sink.done(widx)
        else:
            owner.sink.put(widx, convops.conv_wgrad(xn, gyn, G, (1, 1)))
        if bidx is not None:
            owner.sink.put(bidx, convops.colsum_grouped(gy, G, cols=nout))
        gx = convops.conv_dgrad(gyn, wn, (1, 1)).view(x.shape) if ctx.needs_input_grad[0] else None
        return gx, None, None, None, None, None


class _TcReLU(torch.autograd.Function):
    """ReLU modules of the fused pass (networks without BatchNorm to fuse it into: CifarNet, VGG, the MLPs)."""

    @staticmethod
    def forward(ctx, x):
        dense = x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))
        if not dense:
            x = x.contiguous()
        y = convops.relu(x)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, gy):
        y, = ctx.saved_tensors
        if gy.stride() != y.stride():
            gy = gy.contiguous(memory_format=torch.channels_last) if (y.dim() == 4 and not y.is_contiguous()) else gy.contiguous()
            if gy.stride() != y.stride():
                return torch.where(y > 0, gy, torch.zeros_like(gy))
        return convops.relu_backward(gy, y)


class _TcMaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p):
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        ho, wo = convops.out_size(x.shape[2], x.shape[3], k, s, p)
        y = torch.empty((x.shape[0], x.shape[1], ho, wo), device=x.device, memory_format=torch.channels_last)
        _, arg = convops.max_pool_nhwc(x.permute(0, 2, 3, 1), k, s, p, out=y.permute(0, 2, 3, 1))
        ctx.save_for_backward(arg)
        ctx.cfg = ((x.shape[0], x.shape[2], x.shape[3], x.shape[1]), k, s, p)
        return y

    @staticmethod
    def backward(ctx, gy):
        arg, = ctx.saved_tensors
        shape, k, s, p = ctx.cfg
        if not gy.is_contiguous(memory_format=torch.channels_last):
            gy = gy.contiguous(memory_format=torch.channels_last)
        return convops.max_pool_nhwc_backward(gy.permute(0, 2, 3, 1), arg, shape, k, s, p).permute(0, 3, 1, 2), None, None, None


class _TcAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        ctx.hw = (x.shape[2], x.shape[3])
        y = torch.empty((x.shape[0], x.shape[1], 1, 1), device=x.device)
        convops.avg_pool_nhwc(x.permute(0, 2, 3, 1), out=y.view(x.shape[0], x.shape[1]))
        return y

    @staticmethod
    def backward(ctx, gy):
        return convops.avg_pool_nhwc_backward(gy.reshape(gy.shape[0], gy.shape[1]), ctx.hw).permute(0, 3, 1, 2)


class _NoStream:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


class _Tf32:
    def __init__(self, enable):
        self.enable = enable

    def __enter__(self):
        self.prev = torch.backends.cuda.matmul.allow_tf32
        if self.enable:
            torch.backends.cuda.matmul.allow_tf32 = True

    def __exit__(self, *exc):
        torch.backends.cuda.matmul.allow_tf32 = self.prev


class GroupedWorkers:
    """Gradients of G co-located workers from one pass over `replica` (a FlatModel aliasing the shared parameters).

    rows[g] receives worker g's flat gradient; `data`/`target` are the concatenated [G·B, …] inputs; `losses` ([G])
    receives the per-worker mean losses. BatchNorm running statistics are kept per worker in `[G, C]` tables
    (`running_tables[name]`), so every worker's replica can alias its own row."""

    _BATCHNORM = (torch.nn.BatchNorm1d, torch.nn.BatchNorm2d)

    def __init__(self, replica, rows, loss_fn, side_streams=4, sink=None):
        """`sink`: a factory `(params, slices, view_of) -> sink` replacing the default rows-of-one-arena destination
        (e.g. `ScatterSink` for the coordinate-sharded server); `rows` is then only used for its length."""
        self.replica, self.module, self.rows = replica, replica.module, list(rows)
        self.groups = len(self.rows)
        self.loss_kind = self._loss_kind(loss_fn)
        reason = self.unsupported_reason(self.module, loss_fn)
        if reason:
            raise ValueError("worker fusion does not cover this model: " + reason)
        params = list(self.module.parameters())
        index = {id(p): i for i, p in enumerate(params)}
        self.sink = sink(params, replica.slices, replica._view) if sink is not None else _GradSink(params, replica.slices, self.rows, replica._view)
        device = params[0].device
        self.side_streams = [torch.cuda.Stream(device=device) for _ in range(side_streams)] if device.type == "cuda" and side_streams else []
        self._used_sides = set()
        self.running_tables = {}
        for name, m in self.module.named_modules():
            if isinstance(m, torch.nn.Conv2d):
                fwd = self._conv_forward(index[id(m.weight)], None if m.bias is None else index[id(m.bias)], m)
                m.forward = types.MethodType(fwd, m)
                m._gb_fork = fwd.fork
                if OWN_KERNELS[0] and device.type == "cuda" and os.environ.get("GARFIELD_OWN_ADD", "0") == "1":   # opt-in: no gain measured
                    m._gb_split = _TcSplit.apply
            elif isinstance(m, torch.nn.Linear):
                m.forward = types.MethodType(self._linear_forward(index[id(m.weight)], None if m.bias is None else index[id(m.bias)]), m)
            elif isinstance(m, self._BATCHNORM):
                tables = None
                if m.track_running_stats:
                    tables = (m.running_mean.detach().unsqueeze(0).repeat(self.groups, 1).contiguous(),
                              m.running_var.detach().unsqueeze(0).repeat(self.groups, 1).contiguous())
                    self.running_tables[name] = tables
                m._gb_group = _BnGroup(self, index[id(m.weight)], index[id(m.bias)], tables)
                m.forward = types.MethodType(lambda bn, x: bn._gb_group(bn, x, None, False), m)
            elif (isinstance(m, torch.nn.MaxPool2d) and isinstance(m.kernel_size, int) and isinstance(m.padding, int) and m.dilation == 1
                  and not m.ceil_mode and not m.return_indices and isinstance(m.stride if m.stride is not None else m.kernel_size, int)):
                m.forward = types.MethodType(self._maxpool_forward(type(m).forward), m)
            elif isinstance(m, torch.nn.ReLU):
                m.forward = types.MethodType(self._relu_forward(type(m).forward), m)
            elif isinstance(m, torch.nn.AdaptiveAvgPool2d) and m.output_size in (1, (1, 1)):
                m.forward = types.MethodType(self._avgpool_forward(type(m).forward), m)

    # -- what the fusion covers ---------------------------------------------------------------------------------
    @staticmethod
    def _loss_kind(loss_fn):
        for cls, kind in ((torch.nn.CrossEntropyLoss, "ce"), (torch.nn.NLLLoss, "nll"), (torch.nn.BCELoss, "bce")):
            if isinstance(loss_fn, cls):
                return kind if getattr(loss_fn, "reduction", "mean") == "mean" and getattr(loss_fn, "weight", None) is None else None
        return None

    @classmethod
    def unsupported_reason(cls, module, loss_fn):
        """None when every parameter of `module` belongs to a layer the fusion knows how to differentiate per worker."""
        if cls._loss_kind(loss_fn) is None:
            return "loss %s" % type(loss_fn).__name__
        covered = set()
        for m in module.modules():
            own = list(m.parameters(recurse=False))
            if not own:
                continue
            if isinstance(m, torch.nn.Conv2d) and m.padding_mode == "zeros" and not isinstance(m.padding, str):
                pass
            elif isinstance(m, torch.nn.Linear):
                pass
            elif isinstance(m, cls._BATCHNORM) and m.affine and m.momentum is not None:
                pass
            else:
                return "layer %s" % type(m).__name__
            covered.update(id(p) for p in own)
        if any(id(p) not in covered for p in module.parameters()):
            return "parameters outside Conv2d / Linear / BatchNorm layers"
        if len({id(p) for p in module.parameters()}) != len(list(module.parameters())):
            return "shared parameters"
        return None

    # -- patched layer forwards -----------------------------------------------------------------------------------
    def _conv_forward(self, widx, bidx, conv_module=None):
        owner = self

        def forward(conv, x):
            if owner.own_kernels(x) and conv.groups == 1 and conv.dilation == (1, 1) and x.dim() == 4:
                return _TcConv.apply(x, conv.weight, conv.bias, owner, conv, widx, bidx)
            return _GroupedConv.apply(x, conv.weight, conv.bias, owner, conv, widx, bidx)

        def fork(x):
            """(conv(x), x): the block's identity branch leaves through the convolution's autograd node (see _TcConv.forward)."""
            if owner.own_kernels(x) and conv_module.groups == 1 and conv_module.dilation == (1, 1) and x.dim() == 4 and x.requires_grad:
                return _TcConv.apply(x, conv_module.weight, conv_module.bias, owner, conv_module, widx, bidx, True)
            return conv_module(x), x
        forward.fork = fork
        return forward

    def _linear_forward(self, widx, bidx):
        owner = self

        def forward(lin, x):
            if x.dim() != 2:
                raise ValueError("worker fusion expects 2-D inputs to Linear layers")
            if owner.own_kernels(x) and x.is_contiguous() and lin.in_features % 4 == 0:
                return _TcLinear.apply(x, lin.weight, lin.bias, owner, widx, bidx)
            return _GroupedLinear.apply(x, lin.weight, lin.bias, owner, widx, bidx)
        return forward

    def own_kernels(self, x):
        """Run this layer on the library's own sm_100a kernels? (CUDA fp32 activations, channels-last parameter storage)"""
        return (OWN_KERNELS[0] and x.is_cuda and x.dtype == torch.float32 and getattr(self.replica, "channels_last", False)
                and _lib.available())

    def _relu_forward(self, stock):
        owner = self

        def forward(act, x):
            if owner.own_kernels(x) and x.data_ptr() % 16 == 0 and x.requires_grad:
                return _TcReLU.apply(x)
            return stock(act, x)
        return forward

    def _maxpool_forward(self, stock):
        owner = self

        def forward(pool, x):
            if owner.own_kernels(x) and x.dim() == 4:
                return _TcMaxPool.apply(x, pool.kernel_size, pool.stride if pool.stride is not None else pool.kernel_size, pool.padding)
            return stock(pool, x)
        return forward

    def _avgpool_forward(self, stock):
        owner = self

        def forward(pool, x):
            if owner.own_kernels(x) and x.dim() == 4:
                if x.shape[2] == 1 and x.shape[3] == 1:
                    return x                      # a 1×1 plane is its own average
                return _TcAvgPool.apply(x)
            return stock(pool, x)
        return forward

    def add_milestone(self, module, callback):
        """Call `callback()` during the backward pass as soon as the gradients of `module` and of everything after it in
        the forward order have been produced (their small gradients are flushed into the rows first). Used to start
        work on that coordinate range — e.g. pairwise distances — while the rest of the backward pass runs."""
        params = list(self.module.parameters())
        first = min(i for i, p in enumerate(params) if any(p is q for q in module.parameters()))
        self.milestone_offset = self.replica.slices[first][0]

        def hook(_module, _grad_input, _grad_output):
            self.sink.flush(first_index=first)
            callback()
        module.register_full_backward_hook(hook)
        return self.milestone_offset

    def join_side_streams(self, stream):
        """Make `stream` wait for the side streams that carried work in this pass (waiting on an untouched stream would
        tie a graph capture to uncaptured work)."""
        for k in sorted(self._used_sides):
            stream.wait_stream(self.side_streams[k])

    def next_side(self):
        if not self.side_streams:
            return None
        self._side_rr = (getattr(self, "_side_rr", 0) + 1) % len(self.side_streams)
        self._used_sides.add(self._side_rr)
        return self.side_streams[self._side_rr]

    def all_sides(self):
        """Every side stream, marked as used (the per-worker weight-gradient kernels are spread over all of them)."""
        self._used_sides.update(range(len(self.side_streams)))
        return self.side_streams

    def tf32_matmul(self):
        return _Tf32(torch.backends.cudnn.allow_tf32)   # weight-gradient GEMMs at the precision of the convolutions they replace

    # -- one step ------------------------------------------------------------------------------------------------
    def forward_backward(self, data, target, losses):
        G = self.groups
        self._used_sides = set()
        self.module.train()
        out = self.module(data)
        if self.loss_kind == "ce" and self.own_kernels(out) and out.dim() == 2 and out.is_contiguous():
            # fused grouped cross-entropy: per-row losses and d(Σ_g mean_g)/d(logits) in one launch, the backward pass starts
            # from the (16-byte padded) logit gradients
            row_loss, dlogits = convops.softmax_ce_grouped(out.detach(), target, G)
            out.backward(dlogits[:, :out.shape[1]])
            convops.colsum_grouped(row_loss.view(-1, 1), G, scale=1.0 / (out.shape[0] // G), out=losses.view(G, 1))
            if self.side_streams:
                self.join_side_streams(torch.cuda.current_stream(data.device))
            self.sink.flush()
            return
        if self.loss_kind == "ce":
            per_sample = F.cross_entropy(out, target, reduction="none")
        elif self.loss_kind == "nll":
            per_sample = F.nll_loss(out, target, reduction="none")
        else:
            per_sample = F.binary_cross_entropy(out, target, reduction="none")
        per_worker = per_sample.reshape(G, -1).mean(1)
        per_worker.sum().backward()
        losses.copy_(per_worker.detach())
        if self.side_streams:
            self.join_side_streams(torch.cuda.current_stream(data.device))
        self.sink.flush()


class _BnGroup:
    """Grouped BatchNorm call for one layer (also reached from the fused ResNet block forwards)."""

    def __init__(self, owner, widx, bidx, tables):
        self.owner, self.widx, self.bidx, self.tables = owner, widx, bidx, tables

    def __call__(self, bn, x, residual, relu):
        if not bn.training:
            y = F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)
            if residual is not None:
                y = y + residual
            return torch.relu(y) if relu else y
        sink = self.owner.sink
        rm, rv = self.tables if self.tables is not None else (None, None)
        return grouped_batch_norm(x, bn.weight, bn.bias, rm, rv, self.owner.groups, bn.eps, bn.momentum, residual, relu,
                                  lambda dg, db: (sink.put(self.widx, dg), sink.put(self.bidx, db)))
