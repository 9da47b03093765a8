"""ctypes binding of libgarfield_b200_cuda.so (declared in include/garfield_b200.h).

The library is loaded once per process. On a machine with a CUDA device a missing/unloadable
library is a hard error (`require()`), never a silent PyTorch fallback.
"""

from __future__ import annotations

import ctypes

import torch

from .. import _build

_handle = None
_error = None

vp, i32, i64, u64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_float
pp = ctypes.POINTER(ctypes.c_void_p)


class FusedUpdate(ctypes.Structure):
    """Mirror of `gb_fused_update`."""
    _fields_ = [("mode", i32), ("nesterov", i32), ("first_step", i32), ("n_mirrors", i32),
                ("lr", f32), ("momentum", f32), ("dampening", f32), ("weight_decay", f32),
                ("param", vp), ("mom", vp), ("mirror", vp * 8), ("multicast", vp)]


def _declare(h):
    h.gb_cuda_last_error.restype = ctypes.c_char_p
    h.gb_coordwise.argtypes = [pp, i32, i64, i32, i32, i32, f32, u64, vp, i32, vp, ctypes.POINTER(FusedUpdate), vp]
    h.gb_pairwise_scratch_bytes.restype = i64
    h.gb_pairwise_scratch_bytes.argtypes = [i32]
    h.gb_pairwise_select.argtypes = [pp, i32, i64, i32, i32, i32, i32, i32, vp, i64, vp, vp, vp, vp]
    h.gb_pairwise_scratch_bytes_f64.restype = i64
    h.gb_pairwise_scratch_bytes_f64.argtypes = [i32]
    h.gb_pairwise_select_f64.argtypes = [pp, i32, i64, i32, i32, i32, i32, i32, vp, i64, vp, vp, vp, vp]
    h.gb_coordwise_f64.argtypes = [pp, i32, i64, i32, i32, i32, vp, i32, vp, vp]
    h.gb_gram_scratch_bytes.restype = i64
    h.gb_gram_scratch_bytes.argtypes = [i32]
    h.gb_gram_select_tc.argtypes = [pp, i32, i64, i32, i32, i32, i32, i32, vp, i64, vp, vp, vp, vp]
    h.gb_pairwise_cta_limit.argtypes = [i32]
    h.gb_pairwise_plan.argtypes = [i32, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.c_char_p]
    h.gb_select_only.argtypes = [vp, i32, i32, i32, i32, i32, i32, vp, vp, vp]
    h.gb_gemm_nt.argtypes = [vp, vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, i32, vp]
    h.gb_conv_fprop_tc.argtypes = [vp, vp, vp, vp] + [i32] * 14 + [vp]
    h.gb_conv_dgrad_tc.argtypes = [vp, vp, vp] + [i32] * 11 + [vp, vp]
    h.gb_conv_wgrad_tc.argtypes = [vp, vp, vp] + [i32] * 15 + [i64, i64, vp]
    h.gb_im2col_pad_nhwc.argtypes = [vp, vp] + [i32] * 12 + [vp]
    h.gb_pad_rows.argtypes = [vp, vp, i64, i32, i64, i32, vp]
    h.gb_col2im_nhwc.argtypes = [vp, vp] + [i32] * 12 + [vp]
    h.gb_upsample_zero_nhwc.argtypes = [vp, vp] + [i32] * 8 + [vp]
    h.gb_maxpool_nhwc_forward.argtypes = [vp, vp, vp] + [i32] * 7 + [vp]
    h.gb_maxpool_nhwc_backward.argtypes = [vp, vp, vp] + [i32] * 7 + [vp]
    h.gb_softmax_ce_grouped.argtypes = [vp, i32, vp, vp, vp, i32, i32, i32, i32, vp]
    h.gb_colsum_grouped.argtypes = [vp, i64, vp, i32, i32, i32, f32, vp]
    h.gb_relu_forward.argtypes = [vp, vp, i64, vp]
    h.gb_relu_backward.argtypes = [vp, vp, vp, i64, vp]
    h.gb_add.argtypes = [vp, vp, vp, i64, vp]
    h.gb_avgpool_nhwc_forward.argtypes = [vp, vp, i64, i32, i32, vp]
    h.gb_avgpool_nhwc_backward.argtypes = [vp, vp, i64, i32, i32, vp]
    h.gb_conv2d_forward.argtypes = [vp, vp, vp, vp] + [i32] * 10 + [vp]
    h.gb_conv2d_backward.argtypes = [vp, vp, vp, vp, vp, vp] + [i32] * 9 + [vp]
    h.gb_maxpool2d_forward.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, vp]
    h.gb_maxpool2d_backward.argtypes = [vp, vp, vp, i32, i32, i32, vp]
    h.gb_softmax_loss.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp]
    h.gb_batchnorm_forward.argtypes = [vp] * 8 + [i32, i32, i32, f32, f32, i32, vp]
    h.gb_batchnorm_backward.argtypes = [vp] * 8 + [i32, i32, i32, vp]
    h.gb_flag_signal.argtypes = [vp, ctypes.c_uint32, vp]
    h.gb_flag_wait.argtypes = [pp, i32, ctypes.c_uint32, u64, vp, vp]
    h.gb_epoch_bump.argtypes = [vp, vp]
    h.gb_flag_signal_ep.argtypes = [vp, vp, vp]
    h.gb_flag_wait_ep.argtypes = [pp, i32, vp, ctypes.c_uint32, u64, vp, vp, vp]
    h.gb_reduce_slots.argtypes = [pp, i32, i32, vp, vp]
    h.gb_quorum_decide.argtypes = [pp, i32, vp, i32, u64, pp, pp, i32, vp, vp]
    h.gb_gossip_rounds.argtypes = [pp, pp, i32, i64, i32, i32, i32, pp, pp, i32, i32, i32, vp, vp, pp, i32, vp, u64, vp, vp]
    h.gb_pull_median_operand.argtypes = [pp, i32, i32, i32, i32, vp, vp, vp]
    h.gb_multimem_average.argtypes = [vp, vp, i32, i64, i64, i64, f32, vp]
    h.gb_multi_copy.argtypes = [pp, vp, vp, i32, vp, vp]
    h.gb_global_avgpool_forward.argtypes = [vp, vp, i64, i32, vp]
    h.gb_global_avgpool_backward.argtypes = [vp, vp, i64, i32, vp]
    h.gb_dropout.argtypes = [vp, vp, i64, f32, u64, vp]
    h.gb_bce_loss.argtypes = [vp, vp, vp, vp, i64, vp]
    h.gb_bn_nhwc_forward.argtypes = [vp] * 9 + [i32, i32, f32, f32, i32, vp]
    h.gb_bn_nhwc_backward.argtypes = [vp] * 10 + [i32, i32, i32, vp]
    h.gb_im2col_nhwc.argtypes = [vp, vp] + [i32] * 10 + [vp]
    h.gb_bn_nhwc_forward_grouped.argtypes = [vp] * 9 + [i32, i32, i32, f32, f32, i32, vp]
    h.gb_bn_nhwc_backward_grouped.argtypes = [vp] * 10 + [i32, i32, i32, i32, vp]
    h.gb_sgd_step.argtypes = [vp, vp, vp, i64, f32, f32, f32, f32, i32, i32, pp, i32, vp]
    h.gb_adam_step.argtypes = [vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i32, i32, vp]
    h.gb_rmsprop_step.argtypes = [vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, vp]
    h.gb_adagrad_step.argtypes = [vp, vp, vp, i64, f32, f32, f32, f32, i32, vp]
    h.gb_attack_random.argtypes = [vp, i64, u64, vp]
    h.gb_scale.argtypes = [vp, vp, i64, f32, vp]
    h.gb_attack_drop.argtypes = [vp, vp, i64, f32, u64, vp]
    h.gb_attack_collude.argtypes = [pp, i32, i64, i32, f32, i32, vp, vp]
    h.gb_axpby.argtypes = [vp, vp, i64, f32, f32, vp]
    h.gb_broadcast_copy.argtypes = [vp, pp, i32, i64, vp]
    h.gb_ipc_handle_size.restype = i32
    h.gb_arena_alloc.argtypes = [i64, pp]
    h.gb_arena_free.argtypes = [vp]
    h.gb_ipc_export.argtypes = [vp, vp]
    h.gb_ipc_open.argtypes = [vp, pp]
    h.gb_ipc_close.argtypes = [vp]
    h.gb_enable_peer.argtypes = [i32]
    for name in ("gb_gemm_bf16", "gb_gemm_workspace_bytes"):
        if hasattr(h, name):
            pass


def load():
    """Return the ctypes handle, or None when the library cannot be loaded (error kept)."""
    global _handle, _error
    if _handle is None and _error is None:
        try:
            h = _build.load_cuda()
            _declare(h)
            _handle = h
        except Exception as err:
            _error = err
    return _handle


def available():
    """True when a CUDA device is present AND the sm_100a library is loaded."""
    if not torch.cuda.is_available():
        return False
    if load() is None:
        raise RuntimeError("CUDA device present but libgarfield_b200_cuda.so failed to load: %r" % (_error,))
    return True


def require():
    h = load()
    if h is None:
        raise RuntimeError("libgarfield_b200_cuda.so unavailable: %r" % (_error,))
    return h


def check(rc):
    if rc != 0:
        raise RuntimeError(require().gb_cuda_last_error().decode())


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def row_table(tensors):
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


launches = [0]  # number of kernels of this library launched by this process (bench accounting)


def count(n=1):
    launches[0] += n
