"""Symmetric peer-memory heap for the GPUs of one NVSwitch box.

Every rank of a subset allocates one heap with ``cudaMalloc`` (through the native library), publishes its CUDA IPC
handle over the bootstrap process group (NCCL/gloo is used for rendezvous only) and maps all peers' heaps into its
own address space. After that:

  * ``heap.view(offset, shape, dtype)``  a torch tensor aliasing the local heap (kernels write their outputs here);
  * ``heap.peer_ptr(rank, offset)``      the address of the same offset in a peer's heap — valid inside kernels
                                         (TMA tensor maps, ld/st) because the mapping goes over NVLink;
  * ``heap.barrier()``                   a one-warp kernel on the current stream: release-publish an epoch into every
                                         peer's flag array, acquire-wait for all peers. No host synchronisation.

Offsets are identical on every rank ("symmetric"), so a layout computed once is valid everywhere.
``torch.distributed._symmetric_memory`` is PyTorch's analogue; this implementation is ~100 lines, has no
allocator of its own to fight, and keeps the flag protocol in our own kernels.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..ops import native

_FLAG_BYTES = 4096          # start of the heap: barrier flags (8 x u32 used), rest reserved


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can alias raw device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def _lib():
    lib = native.load()
    if not hasattr(lib, "_peer_sigs"):
        vp = C.c_void_p
        lib.ab_heap_alloc.restype = C.c_int
        lib.ab_heap_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        lib.ab_heap_free.restype = C.c_int
        lib.ab_heap_free.argtypes = [vp]
        lib.ab_ipc_get_handle.restype = C.c_int
        lib.ab_ipc_get_handle.argtypes = [vp, C.c_char_p]
        lib.ab_ipc_open_handle.restype = C.c_int
        lib.ab_ipc_open_handle.argtypes = [C.c_char_p, C.POINTER(vp)]
        lib.ab_ipc_close_handle.restype = C.c_int
        lib.ab_ipc_close_handle.argtypes = [vp]
        lib.ab_peer_barrier.restype = C.c_int
        lib.ab_peer_barrier.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_uint, vp]
        lib.ab_peer_barrier_auto.restype = C.c_int
        lib.ab_peer_barrier_auto.argtypes = [C.POINTER(vp), C.c_int, C.c_int, vp]
        lib._peer_sigs = True
    return lib


class PeerHeap:
    def __init__(self, nbytes: int, group: Optional[dist.ProcessGroup] = None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        assert self.world <= 8
        lib = _lib()
        self.nbytes = int(nbytes) + _FLAG_BYTES
        base = C.c_void_p()
        native.check(lib.ab_heap_alloc(C.byref(base), self.nbytes), "ab_heap_alloc")
        self.base = base.value
        self._bytes = torch.as_tensor(_CudaArray(self.base, self.nbytes), device="cuda")
        self.ptrs: List[int] = [0] * self.world
        self.ptrs[self.rank] = self.base
        self._opened: List[int] = []
        if self.world > 1:
            h = C.create_string_buffer(64)
            native.check(lib.ab_ipc_get_handle(self.base, h), "ab_ipc_get_handle")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, h.raw, group=group)
            for r, hb in enumerate(handles):
                if r == self.rank:
                    continue
                p = C.c_void_p()
                native.check(lib.ab_ipc_open_handle(hb, C.byref(p)), "ab_ipc_open_handle")
                self.ptrs[r] = p.value
                self._opened.append(p.value)
            dist.barrier(group=group)
        self._epoch = 0
        self._cursor = _FLAG_BYTES

    # ---- layout -------------------------------------------------------------------------------
    def alloc(self, shape: Sequence[int], dtype: torch.dtype, align: int = 256):
        """Bump-allocate the same region on every rank; returns (local tensor view, byte offset)."""
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + align - 1) // align * align
        if off + n > self.nbytes:
            raise MemoryError(f"peer heap exhausted: need {off + n} of {self.nbytes} bytes")
        self._cursor = off + n
        return self.view(off, shape, dtype), off

    def view(self, offset: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        return self._bytes[offset:offset + n].view(dtype).view(*shape)

    def peer_ptr(self, rank: int, offset: int) -> int:
        return self.ptrs[rank] + offset

    def reset(self) -> None:
        self._cursor = _FLAG_BYTES

    # ---- synchronisation ------------------------------------------------------------------------
    def barrier(self) -> None:
        """Device-side barrier across the subset on the current stream."""
        if self.world == 1:
            return
        self._epoch += 1
        arr = (C.c_void_p * self.world)(*self.ptrs)
        native.check(_lib().ab_peer_barrier(arr, self.rank, self.world, self._epoch, native.stream_ptr()),
                     "ab_peer_barrier")

    dry = False          # set by ``dry_run``: barriers are skipped so that no kernel of this process spins

    def barrier_replayable(self) -> None:
        """Device-side barrier whose epoch lives in device memory: safe to capture in a CUDA graph and replay."""
        if self.world == 1 or self.dry:
            return
        arr = (C.c_void_p * self.world)(*self.ptrs)
        native.check(_lib().ab_peer_barrier_auto(arr, self.rank, self.world, native.stream_ptr()), "ab_peer_barrier_auto")

    def dry_run(self, fn) -> None:
        """In-process groups only (one thread per GPU): run ``fn`` once with the barriers switched off, then meet the
        other ranks' threads on the host. The first launch of a kernel loads its module (and sets its attributes) under
        a process-wide lock and waits for the launching GPU to go idle; with two GPUs of one process spinning in a peer
        barrier for each other, two such first launches deadlock. After the dry pass every kernel variant, allocation
        and attribute the real pass needs exists, so nothing blocks while a peer spins. Separate processes have separate
        loaders and do not need this (no-op)."""
        if not getattr(self, "_local", False):
            return
        self.dry = True
        try:
            fn()
            # the barrier kernel itself: one self-only launch loads it (every rank does this once, so the device-side
            # epochs stay equal; the value it leaves in slot 0 is below every later epoch)
            arr = (C.c_void_p * 1)(self.base)
            native.check(_lib().ab_peer_barrier_auto(arr, 0, 1, native.stream_ptr()), "ab_peer_barrier_auto")
            torch.cuda.synchronize()
        finally:
            self.dry = False
        self._tbar.wait(600)

    @classmethod
    def local_group(cls, nbytes: int, devices: Sequence[int]) -> List["PeerHeap"]:
        """One heap per device of THIS process (the miner's in-process GPU pool): no IPC handles, the peers'
        allocations are addressed directly once peer access is enabled. Returns the heaps in ``devices`` order; each
        must be used from a thread whose current device is its own."""
        lib = _lib()
        heaps = []
        for r, d in enumerate(devices):
            h = cls.__new__(cls)
            h.group, h.rank, h.world, h._local, h.device = None, r, len(devices), True, int(d)
            h.nbytes = int(nbytes) + _FLAG_BYTES
            with torch.cuda.device(d):
                base = C.c_void_p()
                native.check(lib.ab_heap_alloc(C.byref(base), h.nbytes), "ab_heap_alloc")
                h.base = base.value
                h._bytes = torch.as_tensor(_CudaArray(h.base, h.nbytes), device=f"cuda:{d}")
            h._opened, h._epoch, h._cursor = [], 0, _FLAG_BYTES
            heaps.append(h)
        for a in devices:                      # torch enables peer access lazily on the first cross-device copy
            for b in devices:
                if a != b:
                    assert torch.cuda.can_device_access_peer(a, b), f"no peer access {a} -> {b}"
                    torch.empty(1, device=f"cuda:{a}").copy_(torch.empty(1, device=f"cuda:{b}"))
        import threading
        tbar = threading.Barrier(len(devices))
        for h in heaps:
            h.ptrs = [x.base for x in heaps]
            h._tbar = tbar
        torch.cuda.synchronize()
        return heaps

    def close(self) -> None:
        lib = _lib()
        if getattr(self, "_local", False):
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                if self.base:
                    self._bytes = None
                    lib.ab_heap_free(self.base)
                    self.base = 0
            return
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)
        for p in self._opened:
            lib.ab_ipc_close_handle(p)
        self._opened = []
        if self.base:
            self._bytes = None
            lib.ab_heap_free(self.base)
            self.base = 0
