"""NVLS multicast buffers: one store, every GPU of the subset receives it (SURVEY §5.9 item 2, §8 row R1 "NVLS multicast
for broadcast only"). No counterpart in the reference (single GPU, no collectives).

A ``MulticastBuffer`` is ``nbytes`` of physical memory on every rank (CUDA VMM: ``cuMemCreate``), all bound to one
multicast object (``cuMulticastCreate`` / ``cuMulticastBindMem``) and mapped twice into this rank's address space:

  ``local``   the rank's own copy (ordinary loads and stores; what consumers read)
  ``mc_ptr``  the multicast address: a ``multimem.st`` to it is replicated by the NVSwitch into EVERY rank's copy, so a
              producer sends each byte over its links once instead of once per peer

Only stores go through the switch here (``multimem.st``): values are *copied*, never reduced in the switch, because an
in-switch reduction has a hardware-defined order and every exchange in this repository feeds the CID (canonical-order
rule). The handles travel between the ranks' processes as POSIX file descriptors over a unix socket
(rendezvous and barriers use the torch.distributed group, as for ``PeerHeap``).

``MulticastBuffer.supported()`` says whether the driver offers multicast on this box (it needs NVSwitch + fabric
manager); callers fall back to per-peer stores (``PatchContext.copy``) when it does not.
"""
from __future__ import annotations

import array
import ctypes as C
import os
import socket
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import native


def _cu():
    from cuda.bindings import driver as cu
    return cu


def _ok(res, what: str):
    cu = _cu()
    err = res[0]
    if err != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(f"{what}: {err}")
    return res[1] if len(res) == 2 else res[1:]


def _send_fd(sock: socket.socket, fd: int) -> None:
    sock.sendmsg([b"h"], [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))])


def _recv_fd(sock: socket.socket) -> int:
    fds = array.array("i")
    _, anc, _, _ = sock.recvmsg(1, socket.CMSG_LEN(fds.itemsize))
    for level, typ, data in anc:
        if level == socket.SOL_SOCKET and typ == socket.SCM_RIGHTS:
            fds.frombytes(data[:fds.itemsize])
            return fds[0]
    raise RuntimeError("no file descriptor received")


class MulticastBuffer:
    @staticmethod
    def supported(device: Optional[int] = None) -> bool:
        try:
            cu = _cu()
            dev = torch.cuda.current_device() if device is None else device
            torch.cuda.init()
            v = _ok(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "attr")
            return bool(v)
        except Exception:  # noqa: BLE001 - no driver, old driver, no cuda-python: not supported
            return False

    def __init__(self, nbytes: int, group: Optional[dist.ProcessGroup] = None):
        cu = _cu()
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        dev = torch.cuda.current_device()
        torch.zeros(1, device="cuda")                      # make sure the primary context exists and is current
        FD = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
        mprop = cu.CUmulticastObjectProp()
        mprop.numDevices = self.world
        mprop.handleTypes = FD
        mprop.flags = 0
        mprop.size = int(nbytes)
        gran = _ok(cu.cuMulticastGetGranularity(mprop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED),
                   "cuMulticastGetGranularity")
        self.nbytes = (int(nbytes) + gran - 1) // gran * gran
        mprop.size = self.nbytes
        # ---- the multicast object: rank 0 creates it and hands every peer a file descriptor ----
        path = f"/tmp/arbius_b200_nvls_{os.environ.get('MASTER_PORT', '0')}_{os.getpid() if self.world == 1 else ''}"
        if self.world > 1:
            obj = [None]
            if self.rank == 0:
                obj[0] = f"/tmp/arbius_b200_nvls_{os.getpid()}_{id(self) & 0xffff}"
            dist.broadcast_object_list(obj, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            path = obj[0]
        if self.rank == 0:
            self.mc = _ok(cu.cuMulticastCreate(mprop), "cuMulticastCreate")
            if self.world > 1:
                fd = _ok(cu.cuMemExportToShareableHandle(self.mc, FD, 0), "export multicast handle")
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                if os.path.exists(path):
                    os.unlink(path)
                srv.bind(path)
                srv.listen(self.world)
                dist.barrier(group=group)                  # the socket is listening
                for _ in range(self.world - 1):
                    conn, _ = srv.accept()
                    _send_fd(conn, int(fd))
                    conn.close()
                srv.close()
                os.unlink(path)
                os.close(int(fd))
        else:
            dist.barrier(group=group)
            cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            cli.connect(path)
            fd = _recv_fd(cli)
            cli.close()
            self.mc = _ok(cu.cuMemImportFromShareableHandle(fd, FD), "import multicast handle")
            os.close(fd)
        _ok(cu.cuMulticastAddDevice(self.mc, dev), "cuMulticastAddDevice")
        if self.world > 1:
            dist.barrier(group=group)                      # every device has been added before anyone binds memory
        # ---- this rank's physical memory, bound to the object and mapped for ordinary access ----
        aprop = cu.CUmemAllocationProp()
        aprop.type = cu.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        aprop.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        aprop.location.id = dev
        aprop.requestedHandleTypes = FD
        self.mem = _ok(cu.cuMemCreate(self.nbytes, aprop, 0), "cuMemCreate")
        _ok(cu.cuMulticastBindMem(self.mc, 0, self.mem, 0, self.nbytes, 0), "cuMulticastBindMem")
        acc = cu.CUmemAccessDesc()
        acc.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        acc.location.id = dev
        acc.flags = cu.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
        self.local_ptr = int(_ok(cu.cuMemAddressReserve(self.nbytes, gran, 0, 0), "reserve local"))
        _ok(cu.cuMemMap(self.local_ptr, self.nbytes, 0, self.mem, 0), "map local")
        _ok(cu.cuMemSetAccess(self.local_ptr, self.nbytes, [acc], 1), "access local")
        self.mc_ptr = int(_ok(cu.cuMemAddressReserve(self.nbytes, gran, 0, 0), "reserve multicast"))
        _ok(cu.cuMemMap(self.mc_ptr, self.nbytes, 0, self.mc, 0), "map multicast")
        _ok(cu.cuMemSetAccess(self.mc_ptr, self.nbytes, [acc], 1), "access multicast")
        from .symm import _CudaArray
        self.local = torch.as_tensor(_CudaArray(self.local_ptr, self.nbytes), device="cuda")
        self.local.zero_()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=group)

    def broadcast(self, src: torch.Tensor, offset: int = 0) -> None:
        """Store ``src`` (contiguous, a multiple of 16 bytes) at ``offset`` of EVERY rank's copy with multimem.st."""
        lib = native.load()
        if not hasattr(lib, "_nvls_sig"):
            lib.ab_multimem_copy.restype = C.c_int
            lib.ab_multimem_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
            lib._nvls_sig = True
        n = src.numel() * src.element_size()
        assert src.is_contiguous() and n % 16 == 0 and offset % 16 == 0 and offset + n <= self.nbytes
        native.check(lib.ab_multimem_copy(self.mc_ptr + offset, src.data_ptr(), n, native.stream_ptr()), "ab_multimem_copy")

    def close(self) -> None:
        cu = _cu()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)
        self.local = None
        for ptr in (self.mc_ptr, self.local_ptr):
            cu.cuMemUnmap(ptr, self.nbytes)
            cu.cuMemAddressFree(ptr, self.nbytes)
        cu.cuMulticastUnbind(self.mc, torch.cuda.current_device(), 0, self.nbytes)
        cu.cuMemRelease(self.mem)
        cu.cuMemRelease(self.mc)
