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
        self.arena_mode = 0
        self._small_ok = None
        self.algo = "ring"
        self._arenas = []

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

    def small_collective(self, send: torch.Tensor, out: torch.Tensor, sum_mode: bool, scale: float = 1.0) -> bool:
        """SyncBatchNorm-sized all-gather (out [world, n]) or rank-ordered sum (out [n]) as one kernel over peer memory
        (b200ddp_comm_small_collective). False = peer path unavailable, nothing was launched: use NCCL."""
        # B200DDP_SMALL_COLLECTIVES=nccl selects NCCL's collectives instead (at 2 ranks the peer kernel
        # is the faster one, 3.91 -> 3.72 ms/step of AlexNet-BN; at 8 ranks that model's step time is set by host-side
        # stragglers and varies more from run to run than the two implementations differ)
        want = os.environ.get("B200DDP_SMALL_COLLECTIVES", "own")
        if self._small_ok is False or send.numel() > 2048 or want != "own":
            return False
        if send.data_ptr() % 16:
            return False
        rc = _lib.load().b200ddp_comm_small_collective(self.handle, send.data_ptr(), out.data_ptr(), send.numel(),
                                                       int(bool(sum_mode)), float(scale), _lib.cur_stream(send.device))
        if rc == 2:                       # kNotInitialised: no symmetric memory on this system
            self._small_ok = False
            return False
        _lib.check(rc, "comm_small_collective")
        self._small_ok = True
        return True

    def bucket_wait_last(self, stream: int) -> None:
        """`stream` (a raw cudaStream_t) waits for the bucket collective launched last."""
        _lib.check(_lib.load().b200ddp_bucket_wait_last(self.handle, stream), "bucket_wait_last")

    # ---- registered / peer-mapped gradient arena (R4) -----------------------------------------------------
    def arena_alloc(self, numel: int) -> torch.Tensor:
        """COLLECTIVE. fp32 tensor of `numel` zeros in memory from ncclMemAlloc, registered with the communicator as
        a symmetric window (b200ddp_comm_arena_alloc). `self.arena_mode`: 2 = peers + NVSwitch multicast mapped,
        1 = peers mapped, 0 = registered for NCCL's own collectives only. The communicator owns the memory."""
        ptr, mode = ctypes.c_void_p(), ctypes.c_int(0)
        _lib.check(_lib.load().b200ddp_comm_arena_alloc(self.handle, int(numel) * 4, ctypes.byref(ptr), ctypes.byref(mode)),
                   "comm_arena_alloc")
        self.arena_mode = int(mode.value)

        class _Foreign:      # torch.as_tensor wraps (does not copy) an object with __cuda_array_interface__
            pass
        holder = _Foreign()
        holder.__cuda_array_interface__ = {"shape": (int(numel),), "typestr": "<f4", "data": (int(ptr.value), False),
                                           "version": 2, "strides": None}
        t = torch.as_tensor(holder, device=self.device)
        assert t.data_ptr() == ptr.value and t.dtype == torch.float32
        self._arenas.append((holder, t))
        return t

    def set_algo(self, algo: str, ctas: int = 0) -> bool:
        """'ring' (NCCL PreMulSum), 'own' (this library's peer-memory kernel) or 'sum' (NCCL ncclSum, pre-scaled data).
        Returns whether the peer path exists on this system."""
        code = {"ring": 0, "premul": 0, "own": 1, "peer": 1, "sum": 2}[algo]
        rc = int(_lib.load().b200ddp_comm_set_algo(self.handle, code, int(ctas)))
        if rc < 0:
            raise _lib.B200Error("comm_set_algo: bad argument")
        self.algo = algo
        return rc > 0

    def set_timing(self, on: bool) -> None:
        _lib.check(_lib.load().b200ddp_comm_set_timing(self.handle, int(bool(on))), "comm_set_timing")

    def bucket_times(self):
        """[(ms, elements)] of the bucket collectives of the last joined backward (synchronise first)."""
        ms = (ctypes.c_float * 64)()
        n = (ctypes.c_longlong * 64)()
        k = int(_lib.load().b200ddp_comm_bucket_times(self.handle, ms, n, 64))
        return [(float(ms[i]), int(n[i])) for i in range(max(k, 0))]

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
    reserve_sms_for_nccl(True)
    return comm


def reserve_sms_for_nccl(on: bool) -> None:
    """NCCL's persistent kernels own one SM per channel while a collective runs beside the backward pass: leave those
    SMs out of the one-wave grid plans (see b200ddp/__init__.py). Our own all-reduce kernel is built to co-reside with
    the convolution CTAs instead (peer_allreduce.cu), so with it the compute kernels plan for the whole device."""
    L = _lib.load()
    if not on:
        L.b200ddp_set_sm_budget(0)
        return
    try:
        channels = int(os.environ.get("NCCL_MAX_NCHANNELS", "16"))
    except ValueError:
        channels = 16
    n = ctypes.c_int(148)
    L.b200ddp_device_info(torch.cuda.current_device(), ctypes.byref(n), None, None)
    L.b200ddp_set_sm_budget(max(16, n.value - max(0, min(channels, 32))))


def destroy_all() -> None:
    for c in list(_COMMS.values()):
        c.destroy()
    _COMMS.clear()


atexit.register(destroy_all)
