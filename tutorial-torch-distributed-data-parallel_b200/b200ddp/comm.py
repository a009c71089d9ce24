"""Native NCCL communicator owned by libb200ddp.so, bootstrapped over torch.distributed.

The reference creates its process group with dist.init_process_group (multi-GPU-training-torch.py:34-37)
and lets c10d drive NCCL. We keep that rendezvous (host side stays PyTorch) and use it once to hand
the ncclUniqueId from rank 0 to everybody; after that every collective on the hot path is issued by
the C library on its own communication stream (include/b200ddp.h, R4).
"""
from __future__ import annotations

import atexit
import os
import ctypes
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import _build, _lib

_COMMS: Dict[tuple, "NativeComm"] = {}


class NativeComm:
    def __init__(self, handle: int, rank: int, world: int, device: torch.device) -> None:
        self.handle = ctypes.c_void_p(handle)
        self.rank = rank
        self.world = world
        self.device = device

    # thin wrappers; all stream ordered
    def allreduce_(self, t: torch.Tensor, premul: float = 1.0, stream: Optional[int] = None) -> None:
        L = _lib.load()
        s = stream if stream is not None else _lib.cur_stream(t.device)
        _lib.check(L.b200ddp_comm_allreduce(self.handle, t.data_ptr(), t.numel(), premul, s), "comm_allreduce")

    def allgather(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        L = _lib.load()
        assert recv.numel() == send.numel() * self.world
        _lib.check(L.b200ddp_comm_allgather(self.handle, send.data_ptr(), recv.data_ptr(), send.numel(),
                                            _lib.cur_stream(send.device)), "comm_allgather")

    def broadcast_(self, t: torch.Tensor, root: int = 0) -> None:
        L = _lib.load()
        _lib.check(L.b200ddp_comm_broadcast(self.handle, t.data_ptr(), t.numel() * t.element_size(), root,
                                            _lib.cur_stream(t.device)), "comm_broadcast")

    def bucket_reduce_launch(self, ptr: int, n: int, premul: float, device) -> None:
        L = _lib.load()
        _lib.check(L.b200ddp_bucket_reduce_launch(self.handle, ptr, n, premul, _lib.cur_stream(device)),
                   "bucket_reduce_launch")

    def bucket_reduce_join(self, device) -> None:
        L = _lib.load()
        _lib.check(L.b200ddp_bucket_reduce_join(self.handle, _lib.cur_stream(device)), "bucket_reduce_join")

    def destroy(self) -> None:
        if self.handle:
            try:
                _lib.load().b200ddp_comm_destroy(self.handle)
            finally:
                self.handle = ctypes.c_void_p(None)


def get_comm(device: torch.device, process_group=None) -> Optional[NativeComm]:
    """Communicator for (process_group, device); None when the job has a single rank."""
    if not (dist.is_available() and dist.is_initialized()):
        return None
    world = dist.get_world_size(process_group)
    if world == 1:
        return None
    if device.type != "cuda":
        raise _lib.B200Error("the native communicator needs a CUDA device")
    key = (id(process_group) if process_group is not None else 0, device.index)
    if key in _COMMS:
        return _COMMS[key]
    L = _lib.load()
    nccl_path = _build.nccl_library_path()
    _lib.check(L.b200ddp_nccl_load(nccl_path.encode() if nccl_path else None), "nccl_load")
    rank = dist.get_rank(process_group)
    buf = ctypes.create_string_buffer(128)
    if rank == 0:
        _lib.check(L.b200ddp_comm_unique_id(buf, 128), "comm_unique_id")
    box = [bytes(buf.raw)]
    src = dist.get_global_rank(process_group, 0) if process_group is not None else 0
    dist.broadcast_object_list(box, src=src, group=process_group)
    uid = ctypes.create_string_buffer(box[0], 128)
    handle = ctypes.c_void_p()
    dev_index = device.index if device.index is not None else torch.cuda.current_device()
    _lib.check(L.b200ddp_comm_create(ctypes.byref(handle), uid, rank, world, dev_index), "comm_create")
    comm = NativeComm(handle.value, rank, world, device)
    _COMMS[key] = comm
    # leave the SMs of the collective's channels out of the one-wave grid plans (see b200ddp/__init__.py)
    try:
        channels = int(os.environ.get("NCCL_MAX_NCHANNELS", "16"))
    except ValueError:
        channels = 16
    L.b200ddp_set_sm_budget(148 - max(0, min(channels, 32)))
    return comm


def destroy_all() -> None:
    for c in list(_COMMS.values()):
        c.destroy()
    _COMMS.clear()


atexit.register(destroy_all)
