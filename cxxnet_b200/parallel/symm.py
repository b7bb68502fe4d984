"""Symmetric (peer-mapped) device memory for the one-process-per-GPU model.

Every rank allocates one block with ``cudaMalloc`` (through the kernel library), publishes its
CUDA IPC handle over the ``torch.distributed`` process group (plumbing only), and maps every
other rank's block with ``cudaIpcOpenMemHandle``.  Over NVLink / NVSwitch those mappings are
plain global addresses, so a kernel can load and store peer memory directly.  This replaces
the reference's host-staged mshadow-ps buffers.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from .. import ops


class _RawCudaBuffer:
    """zero-copy view of a raw device allocation for ``torch.as_tensor``."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class SymmetricBlock:
    def __init__(self, nbytes: int, device: torch.device, group=None):
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device
        self.nbytes = ops.round_up(nbytes, 1 << 21)
        L = ops.lib()
        L.cxn_ipc_alloc.argtypes = [ctypes.c_ulonglong, ctypes.POINTER(ctypes.c_ulonglong), ctypes.c_char_p]
        L.cxn_ipc_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_ulonglong)]
        L.cxn_ipc_close.argtypes = [ctypes.c_ulonglong]
        L.cxn_ipc_free.argtypes = [ctypes.c_ulonglong]
        ptr = ctypes.c_ulonglong(0)
        handle = ctypes.create_string_buffer(64)
        torch.cuda.synchronize(device)
        ops.check(L.cxn_ipc_alloc(self.nbytes, ctypes.byref(ptr), handle), "cxn_ipc_alloc")
        self.local_ptr = int(ptr.value)
        handles: List[bytes] = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self.peer_ptrs: List[int] = []
        for r in range(self.world):
            if r == self.rank:
                self.peer_ptrs.append(self.local_ptr)
            else:
                p = ctypes.c_ulonglong(0)
                ops.check(L.cxn_ipc_open(handles[r], ctypes.byref(p)), "cxn_ipc_open(rank %d)" % r)
                self.peer_ptrs.append(int(p.value))
        self.peers_c = (ctypes.c_ulonglong * self.world)(*self.peer_ptrs)
        self._holder = _RawCudaBuffer(self.local_ptr, self.nbytes)
        self.bytes_t = torch.as_tensor(self._holder, device=device)
        self._top = 0
        dist.barrier(group=group)

    def carve(self, nbytes: int, align: int = 256):
        """bump-allocate ``nbytes``; returns (byte offset, uint8 tensor view)."""
        off = ops.round_up(self._top, align)
        if off + nbytes > self.nbytes:
            raise MemoryError("symmetric block exhausted")
        self._top = off + nbytes
        return off, self.bytes_t[off:off + nbytes]

    def alloc(self, n: int, dtype: torch.dtype):
        esz = torch.empty((), dtype=dtype).element_size()
        off, view = self.carve(n * esz)
        return off, view.view(dtype)

    def close(self) -> None:
        L = ops.lib()
        torch.cuda.synchronize(self.device)
        for r, p in enumerate(self.peer_ptrs):
            if r != self.rank:
                L.cxn_ipc_close(p)
        self.peer_ptrs = []
