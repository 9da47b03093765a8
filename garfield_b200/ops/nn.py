"""Hand-written sm_100a neural-network ops with autograd support (no cuBLAS / cuDNN on these paths).

linear(x, weight, bias)   y = x·Wᵀ + b — forward, dgrad and wgrad are all `gemm_nt` launches of the
                          tcgen05/TMEM/TMA GEMM (csrc/cuda/gemm_tc.cu); `relu=True` fuses the
                          activation into the forward epilogue.
gemm_nt(a, b, ...)        C = A·Bᵀ with both operands K-major; fp32 operands run as TF32 (cuDNN/cuBLAS
                          "allow_tf32" numerics), bf16 operands as kind::f16.

Reference counterpart: `nn.Linear` / ATen + cuBLAS.
"""

from __future__ import annotations

import torch

from . import _lib

__all__ = ["gemm_nt", "linear", "Linear", "available"]


def available():
    return _lib.available()


def _pad_k(t, multiple):
    """Return a contiguous [rows, K'] copy whose row stride is a multiple of 16 bytes."""
    k = t.shape[1]
    if k % multiple == 0 and t.is_contiguous() and t.data_ptr() % 16 == 0:
        return t
    kp = (k + multiple - 1) // multiple * multiple
    out = torch.zeros(t.shape[0], kp, dtype=t.dtype, device=t.device)
    out[:, :k] = t
    return out


def gemm_nt(a, b, bias=None, relu=False, out=None):
    """C[M,N] = A[M,K] · B[N,K]ᵀ (+ bias) (ReLU); fp32 output. A and B: both float32 or both bfloat16."""
    if a.dim() != 2 or b.dim() != 2 or a.shape[1] != b.shape[1] or a.dtype != b.dtype:
        raise ValueError("gemm_nt expects A[M,K] and B[N,K] of the same dtype")
    if a.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("gemm_nt supports float32 (TF32) and bfloat16 operands")
    lib = _lib.require()
    mult = 4 if a.dtype == torch.float32 else 8
    a2, b2 = _pad_k(a, mult), _pad_k(b, mult)
    m, n, k = a.shape[0], b.shape[0], a2.shape[1]
    if out is None:
        out = torch.empty(m, n, dtype=torch.float32, device=a.device)
    if bias is not None:
        bias = bias.to(torch.float32).contiguous()
    with torch.cuda.device(a.device):
        _lib.check(lib.gb_gemm_nt(a2.data_ptr(), b2.data_ptr(), out.data_ptr(), bias.data_ptr() if bias is not None else None,
                                  m, n, k, a2.stride(0), b2.stride(0), out.stride(0), 0 if a.dtype == torch.float32 else 1,
                                  int(bool(relu)), _lib.stream_ptr(a.device)))
    _lib.count()
    return out


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu):
        x2 = x.reshape(-1, x.shape[-1])
        y = gemm_nt(x2, weight, bias=bias, relu=relu)
        ctx.save_for_backward(x2, weight, y if relu else None)
        ctx.has_bias, ctx.relu, ctx.in_shape = bias is not None, relu, x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, grad_out):
        x2, weight, y = ctx.saved_tensors
        g = grad_out.reshape(-1, weight.shape[0]).to(torch.float32)
        if ctx.relu:
            g = g * (y > 0)
        grad_x = grad_w = grad_b = None
        if ctx.needs_input_grad[0]:
            grad_x = gemm_nt(g.contiguous(), weight.t().contiguous()).reshape(ctx.in_shape)   # dx = g · W
        if ctx.needs_input_grad[1]:
            grad_w = gemm_nt(g.t().contiguous(), x2.t().contiguous())                          # dW = gᵀ · x
        if ctx.has_bias and ctx.needs_input_grad[2]:
            grad_b = g.sum(dim=0)
        return grad_x, grad_w, grad_b, None


def linear(x, weight, bias=None, relu=False):
    """`F.linear` (optionally fused with ReLU) on the tcgen05 GEMM; falls back to torch off-GPU."""
    if not (x.is_cuda and _lib.available()):
        y = torch.nn.functional.linear(x, weight, bias)
        return torch.relu(y) if relu else y
    return _LinearFn.apply(x, weight, bias, relu)


class Linear(torch.nn.Linear):
    """Drop-in `nn.Linear` whose forward/backward run on the hand-written GEMM."""

    def forward(self, x):
        return linear(x, self.weight, self.bias)


# --------------------------------------------------------------------------- small-CNN layer ops ---

def _ptr(t):
    return t.data_ptr() if t is not None else None


def _on_device(*tensors):
    return all(t is None or t.is_cuda for t in tensors) and _lib.available()


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, relu):
        x, weight = x.contiguous().float(), weight.contiguous().float()
        n, cin, h, w = x.shape
        cout, _, kh, kw = weight.shape
        ho, wo = (h + 2 * padding - kh) // stride + 1, (w + 2 * padding - kw) // stride + 1
        y = torch.empty(n, cout, ho, wo, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_conv2d_forward(_ptr(x), _ptr(weight), _ptr(bias), _ptr(y), n, cin, h, w, cout, kh, kw, stride,
                                                        padding, int(relu), _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.cfg = (stride, padding, relu, bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        stride, padding, relu, has_bias = ctx.cfg
        dy = dy.contiguous().float()
        if relu:
            dy = dy * (y > 0)
        n, cin, h, w = x.shape
        cout, _, kh, kw = weight.shape
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw = torch.empty_like(weight) if ctx.needs_input_grad[1] else None
        db = torch.empty(cout, device=x.device) if (has_bias and ctx.needs_input_grad[2] and dw is not None) else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_conv2d_backward(_ptr(x), _ptr(weight), _ptr(dy), _ptr(dx), _ptr(dw), _ptr(db), n, cin, h, w, cout,
                                                         kh, kw, stride, padding, _lib.stream_ptr(x.device)))
        _lib.count(2)
        if has_bias and ctx.needs_input_grad[2] and db is None:
            db = dy.sum(dim=(0, 2, 3))
        return dx, dw, db, None, None, None


def conv2d(x, weight, bias=None, stride=1, padding=0, relu=False):
    """Direct NCHW convolution (optionally fused with ReLU) with hand-written forward/dgrad/wgrad kernels."""
    if not _on_device(x, weight, bias):
        y = torch.nn.functional.conv2d(x, weight, bias, stride, padding)
        return torch.relu(y) if relu else y
    return _Conv2dFn.apply(x, weight, bias, int(stride), int(padding), bool(relu))


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, stride):
        x = x.contiguous().float()
        n, c, h, w = x.shape
        ho, wo = (h - kernel) // stride + 1, (w - kernel) // stride + 1
        y = torch.empty(n, c, ho, wo, device=x.device)
        arg = torch.empty(n, c, ho, wo, dtype=torch.int32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_maxpool2d_forward(_ptr(x), _ptr(y), _ptr(arg), n * c, h, w, kernel, stride, _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.save_for_backward(arg)
        ctx.shape = (n, c, h, w, ho, wo)
        return y

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        n, c, h, w, ho, wo = ctx.shape
        dy = dy.contiguous().float()
        dx = torch.zeros(n, c, h, w, device=dy.device)
        with torch.cuda.device(dy.device):
            _lib.check(_lib.require().gb_maxpool2d_backward(_ptr(dy), _ptr(arg), _ptr(dx), n * c, h * w, ho * wo, _lib.stream_ptr(dy.device)))
        _lib.count()
        return dx, None, None


def max_pool2d(x, kernel_size, stride=None):
    stride = kernel_size if stride is None else stride
    if not _on_device(x):
        return torch.nn.functional.max_pool2d(x, kernel_size, stride)
    return _MaxPoolFn.apply(x, int(kernel_size), int(stride))


class _SoftmaxLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, target, mode):
        inp = inp.contiguous().float()
        rows, classes = inp.shape
        row_loss = torch.empty(rows, device=inp.device)
        grad = torch.empty_like(inp)
        target = target.contiguous().long()
        with torch.cuda.device(inp.device):
            _lib.check(_lib.require().gb_softmax_loss(_ptr(inp), _ptr(target), _ptr(row_loss), _ptr(grad), rows, classes, mode,
                                                      _lib.stream_ptr(inp.device)))
        _lib.count()
        ctx.save_for_backward(grad)
        return row_loss.mean()

    @staticmethod
    def backward(ctx, dloss):
        (grad,) = ctx.saved_tensors
        return grad * dloss, None, None


def cross_entropy(logits, target):
    """Mean cross-entropy of raw logits (log-softmax + NLL fused, gradient produced in the same kernel)."""
    if not _on_device(logits):
        return torch.nn.functional.cross_entropy(logits, target)
    return _SoftmaxLossFn.apply(logits, target, 1)


def nll_loss(log_probs, target):
    """Mean negative log-likelihood of log-probabilities."""
    if not _on_device(log_probs):
        return torch.nn.functional.nll_loss(log_probs, target)
    return _SoftmaxLossFn.apply(log_probs, target, 0)


class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, eps, momentum, relu):
        x = x.contiguous().float()
        n, c = x.shape[0], x.shape[1]
        hw = x.numel() // (n * c)
        y = torch.empty_like(x)
        mean, invstd = torch.empty(c, device=x.device), torch.empty(c, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_batchnorm_forward(_ptr(x), _ptr(gamma), _ptr(beta), _ptr(y), _ptr(mean), _ptr(invstd),
                                                           _ptr(running_mean), _ptr(running_var), n, c, hw, float(eps), float(momentum),
                                                           int(relu), _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.save_for_backward(x, gamma, mean, invstd, y if relu else None)
        ctx.relu = relu
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, mean, invstd, y = ctx.saved_tensors
        dy = dy.contiguous().float()
        if ctx.relu:
            dy = dy * (y > 0)
        n, c = x.shape[0], x.shape[1]
        hw = x.numel() // (n * c)
        dx, dgamma, dbeta = torch.empty_like(x), torch.empty(c, device=x.device), torch.empty(c, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_batchnorm_backward(_ptr(x), _ptr(dy), _ptr(gamma), _ptr(mean), _ptr(invstd), _ptr(dx), _ptr(dgamma),
                                                            _ptr(dbeta), n, c, hw, _lib.stream_ptr(x.device)))
        _lib.count()
        return dx, (dgamma if gamma is not None else None), (dbeta if gamma is not None else None), None, None, None, None, None


def batch_norm_train(x, gamma, beta, running_mean=None, running_var=None, eps=1e-5, momentum=0.1, relu=False):
    """Training-mode BatchNorm2d (batch statistics, running-stat update), optionally fused with ReLU."""
    if not _on_device(x, gamma):
        y = torch.nn.functional.batch_norm(x, running_mean, running_var, gamma, beta, True, momentum, eps)
        return torch.relu(y) if relu else y
    return _BatchNormFn.apply(x, gamma, beta, running_mean, running_var, eps, momentum, relu)


class _GlobalAvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous().float()
        n, c = x.shape[0], x.shape[1]
        hw = x.numel() // (n * c)
        y = torch.empty(n, c, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_global_avgpool_forward(_ptr(x), _ptr(y), n * c, hw, _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.shape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous().float()
        dx = torch.empty(ctx.shape, device=dy.device)
        n, c = ctx.shape[0], ctx.shape[1]
        with torch.cuda.device(dy.device):
            _lib.check(_lib.require().gb_global_avgpool_backward(_ptr(dy), _ptr(dx), n * c, dx.numel() // (n * c), _lib.stream_ptr(dy.device)))
        _lib.count()
        return dx


def global_avg_pool(x):
    """Mean over the spatial dimensions: [N, C, H, W] → [N, C]."""
    if not _on_device(x):
        return x.mean(dim=tuple(range(2, x.dim())))
    return _GlobalAvgPoolFn.apply(x)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        x = x.contiguous().float()
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_dropout(_ptr(x), _ptr(y), x.numel(), float(p), int(seed), _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.cfg = (float(p), int(seed))
        return y

    @staticmethod
    def backward(ctx, dy):
        p, seed = ctx.cfg
        dy = dy.contiguous().float()
        dx = torch.empty_like(dy)
        with torch.cuda.device(dy.device):
            _lib.check(_lib.require().gb_dropout(_ptr(dy), _ptr(dx), dy.numel(), p, seed, _lib.stream_ptr(dy.device)))
        _lib.count()
        return dx, None, None


_dropout_calls = [0]


def dropout(x, p=0.5, training=True, seed=None):
    """Inverted dropout whose mask is regenerated (not stored) in the backward pass."""
    if not training or p == 0.0:
        return x
    if not _on_device(x):
        return torch.nn.functional.dropout(x, p, True)
    if seed is None:
        _dropout_calls[0] += 1
        seed = (torch.initial_seed() * 0x9E3779B97F4A7C15 + _dropout_calls[0]) & 0x7FFFFFFFFFFFFFFF
    return _DropoutFn.apply(x, p, seed)


class _BceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prob, target):
        prob, target = prob.contiguous().float(), target.contiguous().float()
        loss, grad = torch.empty_like(prob), torch.empty_like(prob)
        with torch.cuda.device(prob.device):
            _lib.check(_lib.require().gb_bce_loss(_ptr(prob), _ptr(target), _ptr(loss), _ptr(grad), prob.numel(), _lib.stream_ptr(prob.device)))
        _lib.count()
        ctx.save_for_backward(grad)
        return loss.mean()

    @staticmethod
    def backward(ctx, dloss):
        (grad,) = ctx.saved_tensors
        return grad * dloss, None


def binary_cross_entropy(prob, target):
    """nn.BCELoss (mean reduction) on probabilities."""
    if not _on_device(prob):
        return torch.nn.functional.binary_cross_entropy(prob, target)
    return _BceFn.apply(prob, target)


# ------------------------------------------------------------ fused channels-last BatchNorm (+ add) (+ ReLU) ---

class _FusedBnFn(torch.autograd.Function):
    """groups > 1: the batch holds `groups` equal statistics groups (co-located workers' batches) — per-group
    mean / variance / running buffers ([groups, C]) and per-group parameter gradients, which go to `sink(dgamma, dbeta)`
    (both [groups, C]) when a sink is given; autograd always sees their sum over the groups."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu, groups=1, sink=None):
        n, c, h, w = x.shape
        rows = n // groups * h * w
        y = torch.empty_like(x)                      # keeps the channels-last strides of x
        mean, invstd = torch.empty(groups, c, device=x.device), torch.empty(groups, c, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_bn_nhwc_forward_grouped(_ptr(x), _ptr(residual), _ptr(gamma), _ptr(beta), _ptr(y), _ptr(mean), _ptr(invstd),
                                                                 _ptr(running_mean), _ptr(running_var), groups, rows, c, float(eps), float(momentum),
                                                                 int(relu), _lib.stream_ptr(x.device)))
        _lib.count()
        ctx.save_for_backward(x, gamma, mean, invstd, y if relu else None)
        ctx.cfg = (groups, rows, c, relu, residual is not None, sink)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, mean, invstd, y = ctx.saved_tensors
        groups, rows, c, relu, has_res, sink = ctx.cfg
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        dgamma, dbeta = torch.empty(groups, c, device=x.device), torch.empty(groups, c, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.require().gb_bn_nhwc_backward_grouped(_ptr(x), _ptr(y), _ptr(dy), _ptr(gamma), _ptr(mean), _ptr(invstd), _ptr(dx),
                                                                  _ptr(dres), _ptr(dgamma), _ptr(dbeta), groups, rows, c, int(relu),
                                                                  _lib.stream_ptr(x.device)))
        _lib.count()
        if sink is not None:
            sink(dgamma, dbeta)
            return dx, None, None, dres, None, None, None, None, None, None, None
        if groups == 1:
            return dx, dgamma[0], dbeta[0], dres, None, None, None, None, None, None, None
        return dx, dgamma.sum(0), dbeta.sum(0), dres, None, None, None, None, None, None, None


def fused_bn_usable(x, bn, residual=None):
    """True when `fused_batch_norm` can take the hand-written kernel for this call."""
    return (bn.training and x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and bn.affine and bn.momentum is not None
            and x.shape[1] % 4 == 0 and x.is_contiguous(memory_format=torch.channels_last) and _lib.available()
            and (residual is None or (residual.shape == x.shape and residual.is_contiguous(memory_format=torch.channels_last))))


def fused_batch_norm(x, bn, residual=None, relu=True):
    """`[relu](bn(x) [+ residual])` for a `BatchNorm2d` module `bn` in ONE kernel (training mode,
    channels-last fp32 CUDA activations); anything else takes the equivalent torch ops. A layer owned by a worker
    group (`parallel.grouped`) is normalised per worker."""
    group = getattr(bn, "_gb_group", None)
    if group is not None:
        return group(bn, x, residual, relu)
    if not fused_bn_usable(x, bn, residual):
        y = torch.nn.functional.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training or bn.running_mean is None,
                                           bn.momentum if bn.momentum is not None else 0.1, bn.eps)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if relu else y
    return _FusedBnFn.apply(x, bn.weight, bn.bias, residual, bn.running_mean if bn.track_running_stats else None,
                            bn.running_var if bn.track_running_stats else None, bn.eps, bn.momentum, relu)


def grouped_batch_norm(x, weight, bias, running_mean, running_var, groups, eps=1e-5, momentum=0.1, residual=None, relu=False, sink=None):
    """Training-mode BatchNorm over `groups` equal slices of the batch, each with its own statistics
    (running_mean / running_var: [groups, C] or None). ONE kernel on channels-last CUDA activations; elsewhere the
    same arithmetic in torch. `sink(dgamma, dbeta)` receives the per-group parameter gradients ([groups, C])."""
    n, c = x.shape[0], x.shape[1]
    if n % groups:
        raise ValueError("batch %d is not divisible into %d groups" % (n, groups))
    if (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and c % 4 == 0 and x.is_contiguous(memory_format=torch.channels_last)
            and _lib.available() and (residual is None or (residual.shape == x.shape and residual.is_contiguous(memory_format=torch.channels_last)))):
        return _FusedBnFn.apply(x, weight, bias, residual, running_mean, running_var, eps, momentum, relu, groups, sink)
    xg = x.reshape(groups, n // groups, *x.shape[1:])
    dims = [1] + list(range(3, xg.dim()))
    mean = xg.mean(dims, keepdim=True)
    var = xg.var(dims, unbiased=False, keepdim=True)
    if running_mean is not None:
        with torch.no_grad():
            count = xg.numel() // (groups * c)
            running_mean.mul_(1 - momentum).add_(momentum * mean.reshape(groups, c))
            running_var.mul_(1 - momentum).add_(momentum * var.reshape(groups, c) * (count / max(count - 1, 1)))
    shape = [groups, 1, c] + [1] * (xg.dim() - 3)
    gam, bet = _per_group(weight, bias, groups, sink)
    y = ((xg - mean) * torch.rsqrt(var + eps) * gam.reshape(shape) + bet.reshape(shape)).reshape(x.shape)
    if residual is not None:
        y = y + residual
    return torch.relu(y) if relu else y


class _PerGroupFn(torch.autograd.Function):
    """Replicate (weight, bias) for every group; the per-group gradients go to the sink instead of being summed."""

    @staticmethod
    def forward(ctx, weight, bias, groups, sink):
        ctx.sink = sink
        return weight.unsqueeze(0).repeat(groups, *[1] * weight.dim()), bias.unsqueeze(0).repeat(groups, *[1] * bias.dim())

    @staticmethod
    def backward(ctx, dw, db):
        ctx.sink(dw, db)
        return None, None, None, None


def _per_group(weight, bias, groups, sink):
    if sink is None:
        return weight.unsqueeze(0).expand(groups, *weight.shape), bias.unsqueeze(0).expand(groups, *bias.shape)
    return _PerGroupFn.apply(weight, bias, groups, sink)


def im2col_nhwc(x, kernel_size, stride, padding):
    """Patch matrix [N·Ho·Wo, KH·KW·C] in (kh, kw, c) order of a channels-last CUDA activation (C % 4 == 0)."""
    n, c, h, w = x.shape
    (kh, kw), (sh, sw), (ph, pw) = kernel_size, stride, padding
    ho, wo = (h + 2 * ph - kh) // sh + 1, (w + 2 * pw - kw) // sw + 1
    if not (x.is_cuda and x.dtype == torch.float32 and c % 4 == 0 and x.is_contiguous(memory_format=torch.channels_last)):
        cols = torch.nn.functional.unfold(x, (kh, kw), 1, (ph, pw), (sh, sw))                 # [N, C·KH·KW, L], (c, kh, kw) order
        return cols.view(n, c, kh * kw, ho * wo).permute(0, 3, 2, 1).reshape(n * ho * wo, kh * kw * c)
    cols = torch.empty(n * ho * wo, kh * kw * c, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.require().gb_im2col_nhwc(_ptr(x), _ptr(cols), n, h, w, c, kh, kw, sh, sw, ph, pw, _lib.stream_ptr(x.device)))
    _lib.count()
    return cols
