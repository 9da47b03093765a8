"""Peer-memory arenas: device buffers of one rank that every other rank of the node can read and
write directly from inside a kernel (NVLink / NVSwitch P2P), shared through CUDA IPC.

This replaces the reference's transport of CPU tensors through torch RPC / gloo collectives
(/root/reference/pytorch_impl/libs/garfieldpp/server.py:112-184,
 applications/Garfield_CC/trainer.py:84-196): gradients and parameters never leave device
memory, and the aggregation kernels take peer pointers in their row tables (fused gather + GAR).
`torch.distributed` is used only to exchange the 64-byte IPC handles (bootstrap).
"""

from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import _lib

__all__ = ["DeviceBuffer", "PeerArena", "SymmArena", "make_arena", "tensor_from_ptr"]


class _CudaArray:
    """Minimal __cuda_array_interface__ holder so torch can wrap a raw device pointer."""

    def __init__(self, ptr, nelem, owner=None):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (nelem,), "typestr": "<f4", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def tensor_from_ptr(ptr, nelem, device, owner=None):
    """float32 tensor view of `nelem` elements at device address `ptr` (no ownership)."""
    with torch.cuda.device(device):
        return torch.as_tensor(_CudaArray(ptr, nelem, owner), device=torch.device("cuda", torch.device(device).index))


class DeviceBuffer:
    """A cudaMalloc'd float32 buffer owned by this process (IPC-exportable, unlike memory carved
    out of the caching allocator)."""

    def __init__(self, nelem, device):
        self.lib = _lib.require()
        self.device = torch.device(device)
        self.nelem = int(nelem)
        raw = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.gb_arena_alloc(self.nelem * 4, ctypes.byref(raw)))
        self.ptr = raw.value
        self.tensor = tensor_from_ptr(self.ptr, self.nelem, self.device, owner=self)
        self.tensor.zero_()

    def ipc_handle(self):
        size = self.lib.gb_ipc_handle_size()
        buf = ctypes.create_string_buffer(size)
        _lib.check(self.lib.gb_ipc_export(ctypes.c_void_p(self.ptr), buf))
        return bytes(buf.raw)

    def free(self):
        if self.ptr:
            with torch.cuda.device(self.device):
                self.lib.gb_arena_free(ctypes.c_void_p(self.ptr))
            self.ptr = None


class PeerArena:
    """One float32 buffer of `nelem` elements per rank, mapped into every rank of `group`.

    `local` is this rank's tensor; `peer(r)` is a tensor aliasing rank r's buffer (valid for
    loads and stores from kernels and copies running on this rank's device).
    """

    def __init__(self, nelem, device, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device)
        self.nelem = int(nelem)
        self._own = DeviceBuffer(self.nelem, self.device)
        self.local = self._own.tensor
        self._peers = {self.rank: self.local}
        self._opened = []
        self.multicast_ptr = 0
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, self._own.ipc_handle(), group=group)
            lib = _lib.require()
            for r, handle in enumerate(handles):
                if r == self.rank:
                    continue
                raw = ctypes.c_void_p()
                with torch.cuda.device(self.device):
                    _lib.check(lib.gb_ipc_open(ctypes.create_string_buffer(handle, len(handle)), ctypes.byref(raw)))
                self._opened.append(raw.value)
                self._peers[r] = tensor_from_ptr(raw.value, self.nelem, self.device, owner=self)
            dist.barrier(group=group)

    def peer(self, rank):
        return self._peers[rank]

    def close(self):
        lib = _lib.require()
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        with torch.cuda.device(self.device):
            for ptr in self._opened:
                lib.gb_ipc_close(ctypes.c_void_p(ptr))
        self._opened = []
        self._own.free()


class SymmArena:
    """PeerArena on torch symmetric memory: the same per-rank buffer mapped into every rank, plus — where the fabric supports
    NVLS — ONE multicast address bound to all ranks' buffers (`multicast_ptr`, 0 when unavailable). Kernels of this library
    use it with `multimem.st` (a store that lands in every GPU's copy) and `multimem.ld_reduce` (a load that returns the sum
    of all copies, added inside the NVSwitch). torch only allocates and exchanges handles; every access is ours."""

    def __init__(self, nelem, device, group=None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = torch.device(device)
        self.nelem = int(nelem)
        with torch.cuda.device(self.device):
            self._tensor = symm.empty(self.nelem, dtype=torch.float32, device=self.device)
            self._handle = symm.rendezvous(self._tensor, self.group)
        self.local = self._tensor
        self.local.zero_()
        ptrs = list(self._handle.buffer_ptrs)
        self._peers = {r: (self.local if r == self.rank else tensor_from_ptr(ptrs[r], self.nelem, self.device, owner=self)) for r in range(self.world)}
        self.multicast_ptr = int(getattr(self._handle, "multicast_ptr", 0) or 0)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=group)

    def peer(self, rank):
        return self._peers[rank]

    def close(self):
        if dist.is_initialized():
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
        self._peers = {}
        self._handle = None
        self._tensor = None


def make_arena(nelem, device, group=None, multicast=False):
    """A peer arena; with `multicast=True` and several ranks, one on symmetric memory with an NVLS multicast address when
    the platform provides it (falls back to the CUDA-IPC arena, whose `multicast_ptr` is 0)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if multicast and world > 1:
        try:
            arena = SymmArena(nelem, device, group)
            ok = torch.tensor([1 if arena.multicast_ptr else 0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()):
                return arena
            arena.close()
        except Exception:
            pass
    arena = PeerArena(nelem, device, group)
    arena.multicast_ptr = 0
    return arena
