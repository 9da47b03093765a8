"""Python face of the native multi-replica register (csrc/cpu/multiregister.cpp).

`MultiRegister(replicas, nbytes, consume=False)`: writers publish fixed-size byte payloads
(e.g. a flat fp32 tensor), readers get the newest one (or, in consume mode, each payload exactly
once, oldest first). Blocking with optional timeouts; `close()` wakes everyone.
Role parity: reference native/include/multiregister.hpp:232-552 and the TF
`py_multibuffer` ctypes interface (tensorflow_impl/rsrcs/native/py_multibuffer).
"""

from __future__ import annotations

import ctypes

import torch

from .. import native as _native

__all__ = ["MultiRegister", "RegisterClosed", "RegisterTimeout"]


class RegisterClosed(Exception):
    pass


class RegisterTimeout(Exception):
    pass


class MultiRegister:
    def __init__(self, replicas, nbytes, consume=False):
        self._lib = _native.lib()
        self._nbytes = int(nbytes)
        self._handle = self._lib.gb_mreg_create(int(replicas), self._nbytes, 1 if consume else 0)
        if not self._handle:
            raise ValueError("invalid register geometry")

    @staticmethod
    def _ns(timeout):
        return -1 if timeout is None else int(timeout * 1e9)

    def _check(self, rc):
        if rc == 1:
            raise RegisterClosed()
        if rc == 2:
            raise RegisterTimeout()

    def write(self, tensor, timeout=None):
        """Publish the bytes of a contiguous CPU tensor of exactly `nbytes` bytes."""
        tensor = tensor.contiguous()
        if tensor.numel() * tensor.element_size() != self._nbytes or tensor.device.type != "cpu":
            raise ValueError("payload must be a CPU tensor of %d bytes" % self._nbytes)
        self._check(self._lib.gb_mreg_write(self._handle, tensor.data_ptr(), self._ns(timeout)))

    def read(self, out, timeout=None):
        """Fill `out` (CPU tensor of `nbytes` bytes) and return the payload's sequence number."""
        if out.numel() * out.element_size() != self._nbytes or not out.is_contiguous():
            raise ValueError("destination must be a contiguous CPU tensor of %d bytes" % self._nbytes)
        seq = ctypes.c_uint64(0)
        self._check(self._lib.gb_mreg_read(self._handle, out.data_ptr(), ctypes.byref(seq), self._ns(timeout)))
        return seq.value

    def close(self):
        if self._handle:
            self._lib.gb_mreg_close(self._handle)

    def __del__(self):
        handle, self._handle = getattr(self, "_handle", None), None
        if handle:
            self._lib.gb_mreg_destroy(handle)
