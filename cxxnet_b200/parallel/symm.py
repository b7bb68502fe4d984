"""Symmetric (peer-mapped) device memory for the one-process-per-GPU model.

Two ways to get it, same interface (``local_ptr``, ``peer_ptrs``, ``multicast_ptr``, ``carve`` / ``alloc``):

``ipc``    every rank allocates one block with ``cudaMalloc`` (through the kernel library), publishes its CUDA IPC
           handle over the ``torch.distributed`` process group (plumbing only), and maps every other rank's block
           with ``cudaIpcOpenMemHandle``.  Unicast peer addresses only.
``nvls``   the block comes from ``torch.distributed._symmetric_memory`` (CUDA virtual memory management: every
           rank's physical allocation mapped into every process, exchanged by torch's rendezvous - plumbing), which
           also binds the blocks to an NVSwitch MULTICAST object where the fabric supports it.  ``multicast_ptr`` is
           then an address whose stores reach every rank's block (``multimem.st``) and whose load-reduce sums all
           ranks' copies inside the switch (``multimem.ld_reduce``): the exchange kernels use it for the gradient
           reduce-scatter, the weight all-gather and the activation gathers (csrc/kernels/fused_exchange.cu,
           tp_collectives.cu).  Selected with ``CXN_SYMM=nvls`` (default when available: ``CXN_SYMM=auto``).

Over NVLink / NVSwitch the peer mappings are plain global addresses, so a kernel can load and store peer memory
directly.  This replaces the reference's host-staged mshadow-ps buffers.  Every peer range
and the multicast range are reported to the kernel library (``cxn_note_peer_range``): the cluster-multicast TMA
routes of the GEMM / convolution take only operands in this GPU's own memory.
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
        self.multicast_ptr = 0
        self.kind = "ipc"
        self._symm_tensor = None
        L = ops.lib()
        import os
        want = os.environ.get("CXN_SYMM", "auto")
        if want in ("auto", "nvls") and self._try_nvls(want == "nvls"):
            return
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

    def _try_nvls(self, required: bool) -> bool:
        """allocate through torch's symmetric memory; True when the block (with or without a multicast address) is up"""
        import torch.distributed as dist
        L = ops.lib()
        try:
            import torch.distributed._symmetric_memory as symm_mem
            group = self.group if self.group is not None else dist.group.WORLD
            t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            hdl = symm_mem.rendezvous(t, group=group)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            ok = len(ptrs) == self.world and ptrs[self.rank] == t.data_ptr()
        except Exception as exc:        # no VMM / fabric support, or an older torch: the IPC path works everywhere
            if required:
                raise
            ok, exc_text = False, repr(exc)
        # every rank must take the same path
        flag = torch.tensor([1 if ok else 0], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            if required:
                raise RuntimeError("CXN_SYMM=nvls: torch symmetric memory is not available on every rank")
            return False
        t.zero_()
        torch.cuda.synchronize(self.device)
        self._symm_tensor, self._symm_handle = t, hdl
        self.kind = "nvls" if mc else "vmm"
        self.local_ptr = ptrs[self.rank]
        self.peer_ptrs = ptrs
        self.multicast_ptr = mc
        L.cxn_note_peer_range.argtypes = [ctypes.c_void_p, ctypes.c_ulonglong]
        for r, p in enumerate(ptrs):
            if r != self.rank:
                L.cxn_note_peer_range(p, self.nbytes)
        if mc:
            L.cxn_note_peer_range(mc, self.nbytes)
        self.peers_c = (ctypes.c_ulonglong * self.world)(*self.peer_ptrs)
        self.bytes_t = t
        self._top = 0
        dist.barrier(group=self.group)
        return True

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
        if self._symm_tensor is not None:
            self.peer_ptrs = []
            return
        for r, p in enumerate(self.peer_ptrs):
            if r != self.rank:
                L.cxn_ipc_close(p)
        self.peer_ptrs = []
