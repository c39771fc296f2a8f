"""Host side of the communicator: rendezvous of CUDA IPC handles over an existing
``torch.distributed`` process group, and tensor views of the symmetric heap.

Stands where ``dist.init_process_group("nccl")`` + ProcessGroupNCCL stand for the
gradient path of the reference (01_torch_distributor/01_basic_torch_distributor.py:269,291):
the process group (gloo or NCCL, whatever the launcher set up) is only used once,
to all-gather 128-byte handle blobs; after that the ranks talk through peer
loads/stores issued by the kernels in csrc/.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch
import torch.distributed as dist

from . import _lib


class _RawCudaBuffer:
    """Exposes a raw device allocation through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner  # keeps the communicator (and so the allocation) alive
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }


class Communicator:
    """One per process and use: owns ``heap_bytes`` of peer-visible device memory.

    ``process_group`` (default: the global group) must already be initialised
    when world size > 1; with no initialised group the communicator is a
    single-rank one and nothing is exchanged.
    """

    def __init__(self, process_group=None, heap_bytes: int = 0, device: Optional[torch.device] = None,
                 multicast: Optional[bool] = False):
        """``multicast``: False (default) = the heap is a cudaMalloc block shared through CUDA IPC; None = use
        an NVLS (multicast) mapping of the heap when the platform offers one; True = require it.  With a
        multicast mapping the two-shot bucket kernel reduces in the NVSwitch (csrc/update.cu)."""
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(process_group)
            self.world = dist.get_world_size(process_group)
        else:
            self.rank, self.world = 0, 1
        if self.world not in (1, 2, 4, 8):
            raise _lib.B200DPError(f"world size {self.world} is not supported (1, 2, 4 or 8 GPUs of one node)")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.process_group = process_group
        lib = _lib.lib()
        self._handle = ctypes.c_void_p()
        self._symm = None
        with torch.cuda.device(self.device):
            symm = (self._symmetric_heap(int(heap_bytes), process_group)
                    if (multicast is not False and self.world > 1 and heap_bytes > 0) else None)
            if symm is not None:
                tensor, hdl, nbytes = symm
                off = int(getattr(hdl, "offset", 0) or 0)  # the tensor's place inside the exchanged block
                ptrs = (ctypes.c_void_p * self.world)(*[int(p) + off for p in hdl.buffer_ptrs])
                _lib.check(lib.b200dp_comm_create_external(self.rank, self.world, nbytes, ptrs, int(hdl.multicast_ptr) + off,
                                                           ctypes.byref(self._handle)), "b200dp_comm_create_external")
                self._symm = (tensor, hdl)  # the allocation lives as long as the communicator
            else:
                if multicast is True and self.world > 1 and heap_bytes > 0:
                    raise _lib.B200DPError("no multicast (NVLS) mapping is available for the heap on this platform")
                _lib.check(lib.b200dp_comm_create(self.rank, self.world, int(heap_bytes), ctypes.byref(self._handle)),
                           "b200dp_comm_create")
            if self.world > 1:
                blob = ctypes.create_string_buffer(_lib.COMM_HANDLE_BYTES)
                _lib.check(lib.b200dp_comm_export(self._handle, blob), "b200dp_comm_export")
                gathered = [None] * self.world
                dist.all_gather_object(gathered, bytes(blob.raw), group=process_group)
                allb = ctypes.create_string_buffer(b"".join(gathered), _lib.COMM_HANDLE_BYTES * self.world)
                _lib.check(lib.b200dp_comm_open_peers(self._handle, allb), "b200dp_comm_open_peers")
                # nobody may touch a peer's flags before every rank has mapped them
                dist.barrier(group=process_group)
        self.multicast = bool(lib.b200dp_comm_has_multicast(self._handle))
        self.heap_bytes = int(lib.b200dp_comm_heap_bytes(self._handle))
        self._heap_ptr = int(lib.b200dp_comm_heap(self._handle, self.rank))
        self._heap_u8 = torch.as_tensor(_RawCudaBuffer(self._heap_ptr, self.heap_bytes, self), device=self.device)

    def _symmetric_heap(self, heap_bytes: int, process_group):
        """The heap as a torch symmetric-memory allocation (cuMemCreate + handle exchange + multicast
        binding done by torch.distributed._symmetric_memory), or None when that is not available here or
        gives no multicast address — every rank takes the same decision (it is agreed on with an all-reduce)."""
        ok, out = 1, None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            nbytes = max(1, (heap_bytes + (2 << 20) - 1) >> 21) << 21
            group = process_group if process_group is not None else dist.group.WORLD
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
            hdl = symm_mem.rendezvous(t, group)
            off = int(getattr(hdl, "offset", 0) or 0)
            if (not int(hdl.multicast_ptr) or len(hdl.buffer_ptrs) != self.world
                    or int(hdl.buffer_ptrs[self.rank]) + off != t.data_ptr()):
                ok = 0
            else:
                t.zero_()
                out = (t, hdl, nbytes)
        except Exception:  # noqa: BLE001  (no driver support, no NVSwitch, an older torch, ...)
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=process_group)
        torch.cuda.synchronize(self.device)
        return out if int(flag) == 1 else None

    # -- C handle ------------------------------------------------------------------
    @property
    def handle(self) -> ctypes.c_void_p:
        if not self._handle:
            raise _lib.B200DPError("communicator was destroyed")
        return self._handle

    def next_seq(self) -> int:
        return int(_lib.lib().b200dp_comm_next_seq(self.handle))

    def barrier(self) -> None:
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200dp_comm_barrier(self.handle, _lib.current_stream()), "b200dp_comm_barrier")

    def all_gather_scalars(self, values: torch.Tensor) -> torch.Tensor:
        """``[world, k]`` tensor whose row r holds rank r's ``values`` (1-D, float32 or int32, on this
        communicator's device).  One small kernel over peer memory; what the reference does with three
        ``accelerator.gather(torch.tensor(x).to(device))`` calls per epoch
        (04_accelerate/01_cifar_accelerate.ipynb:638-640).  Collective."""
        if values.dim() != 1 or values.dtype not in (torch.float32, torch.int32) or not values.is_cuda:
            raise ValueError("expected a 1-D float32 or int32 CUDA tensor")
        values = values.contiguous()
        out = torch.empty((self.world, values.numel()), dtype=values.dtype, device=values.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200dp_comm_all_gather32(self.handle, values.data_ptr(), values.numel(), out.data_ptr(),
                                                           _lib.current_stream()), "b200dp_comm_all_gather32")
        return out

    def check_error(self) -> None:
        """Raise if a kernel on this rank timed out waiting for a peer."""
        rc = _lib.lib().b200dp_comm_error(self.handle)
        if rc != 0:
            raise _lib.B200DPError("a libb200dp kernel gave up waiting for a peer rank (peer dead or out of step)")

    # -- heap views ----------------------------------------------------------------
    def heap_view(self, byte_offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """A tensor over ``numel`` elements of the LOCAL heap starting at ``byte_offset``."""
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        if byte_offset % 16 or byte_offset + nbytes > self.heap_bytes:
            raise _lib.B200DPError("heap view out of range or misaligned")
        return self._heap_u8[byte_offset:byte_offset + nbytes].view(dtype)

    def destroy(self) -> None:
        if self._handle:
            self._heap_u8 = None
            _lib.lib().b200dp_comm_destroy(self._handle)
            self._handle = ctypes.c_void_p()
            self._symm = None


_stats_comms = {}


def stats_communicator(process_group=None) -> Communicator:
    """Process-wide communicator used by SyncBatchNorm layers (no heap needed)."""
    key = (id(process_group) if process_group is not None else None, torch.cuda.current_device())
    comm = _stats_comms.get(key)
    if comm is None:
        comm = Communicator(process_group, heap_bytes=0)
        _stats_comms[key] = comm
    return comm


def all_gather_scalars(values: torch.Tensor, process_group=None) -> torch.Tensor:
    """``accelerator.gather`` for a handful of per-rank metric values: ``[world, k]`` from a 1-D float32 /
    int32 CUDA tensor, through the process-wide communicator the SyncBatchNorm layers use.  A single
    process (no process group, or world size 1) gets its own values back as ``[1, k]``."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(process_group) == 1:
        if values.dim() != 1:
            raise ValueError("expected a 1-D tensor")
        return values.unsqueeze(0).clone()
    return stats_communicator(process_group).all_gather_scalars(values)


def reset_stats_communicators() -> None:
    for comm in _stats_comms.values():
        comm.destroy()
    _stats_comms.clear()
