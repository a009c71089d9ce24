"""ctypes binding of libb200ddp.so (the C ABI declared in include/b200ddp.h).

There is no CPU implementation behind these calls: if the shared library is missing, or a call is
made for tensors that are not on a CUDA device, the caller gets an exception (no
fallback for kernels).
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import (POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_ulonglong, c_void_p)

import torch

from . import _build

_LIB = None
_LOCK = threading.Lock()


class B200Error(RuntimeError):
    pass


_P = c_void_p  # every device pointer travels as void*

class WeightPrepJob(ctypes.Structure):
    """b200ddp_weight_prep_job (include/b200ddp.h)."""
    _fields_ = [("w", c_void_p), ("out", c_void_p), ("Cout", c_int), ("Cin", c_int), ("KH", c_int), ("KW", c_int),
                ("ld_out", c_longlong), ("mode", c_int)]


class AlexNetDesc(ctypes.Structure):
    """b200ddp_alexnet_desc (include/b200ddp.h)."""
    _fields_ = [("B", c_int), ("hidden", c_int), ("num_classes", c_int), ("dropout_p", c_float),
                ("w", c_void_p * 8), ("b", c_void_p * 8), ("conv1_ktab", c_void_p), ("conv1_order", c_int), ("workspace", c_void_p),
                ("workspace_bytes", c_longlong), ("device", c_int)]


READY_FN = ctypes.CFUNCTYPE(None, c_void_p, c_int)      # b200ddp_ready_fn

_PROTOS = {
    "b200ddp_last_error": (c_char_p, []),
    "b200ddp_version": (c_int, []),
    "b200ddp_launch_count": (c_ulonglong, []),
    "b200ddp_set_sm_budget": (c_int, [c_int]),
    "b200ddp_device_info": (c_int, [c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "b200ddp_adam_step_flat": (c_int, [_P, _P, _P, _P, c_longlong, c_double, c_double, c_double, c_double, c_double,
                                       c_longlong, _P, c_int, _P]),
    "b200ddp_plan_buckets": (c_int, [POINTER(c_longlong), POINTER(c_int), c_int, POINTER(c_longlong), c_int,
                                     POINTER(c_int), POINTER(c_int)]),
    "b200ddp_pack_scale_clip": (c_int, [_P, _P, c_longlong, c_float, c_float, c_int, c_int, _P]),
    "b200ddp_scale_by_device_scalar": (c_int, [_P, c_longlong, _P, c_int, c_int, _P]),
    "b200ddp_sqnorm_scratch_bytes": (c_longlong, []),
    "b200ddp_sqnorm": (c_int, [_P, c_longlong, _P, c_int, _P, c_longlong, c_int, _P]),
    "b200ddp_clip_coef": (c_int, [_P, c_float, c_int, _P, _P, c_int, _P]),
    "b200ddp_nccl_load": (c_int, [c_char_p]),
    "b200ddp_nccl_version": (c_int, [POINTER(c_int)]),
    "b200ddp_comm_unique_id": (c_int, [_P, c_int]),
    "b200ddp_comm_create": (c_int, [POINTER(c_void_p), _P, c_int, c_int, c_int]),
    "b200ddp_comm_destroy": (c_int, [_P]),
    "b200ddp_comm_rank": (c_int, [_P, POINTER(c_int), POINTER(c_int)]),
    "b200ddp_comm_allreduce": (c_int, [_P, _P, c_longlong, c_float, _P]),
    "b200ddp_comm_allgather": (c_int, [_P, _P, _P, c_longlong, _P]),
    "b200ddp_comm_broadcast": (c_int, [_P, _P, c_longlong, c_int, _P]),
    "b200ddp_bucket_reduce_launch": (c_int, [_P, _P, c_longlong, c_float, _P]),
    "b200ddp_bucket_reduce_join": (c_int, [_P, _P]),
    "b200ddp_bucket_wait_last": (c_int, [_P, _P]),
    "b200ddp_comm_arena_alloc": (c_int, [_P, c_longlong, POINTER(c_void_p), POINTER(c_int)]),
    "b200ddp_comm_set_algo": (c_int, [_P, c_int, c_int]),
    "b200ddp_comm_small_collective": (c_int, [_P, _P, _P, c_int, c_int, c_float, _P]),
    "b200ddp_comm_set_cta_shape": (c_int, [_P, c_int]),
    "b200ddp_comm_set_timing": (c_int, [_P, c_int]),
    "b200ddp_comm_bucket_times": (c_int, [_P, POINTER(c_float), POINTER(c_longlong), c_int]),
    "b200ddp_linear_workspace_bytes": (c_longlong, [c_int, c_int, c_int]),
    "b200ddp_linear_fwd": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, c_ulonglong, c_int, _P,
                                   c_longlong, c_int, _P]),
    "b200ddp_linear_bwd_dx": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, _P, c_longlong, c_int, _P]),
    "b200ddp_linear_bwd_dw": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_gate_bias_bwd": (c_int, [_P, _P, c_float, _P, _P, c_int, c_int, c_int, _P]),
    "b200ddp_softmax_ce": (c_int, [_P, _P, _P, _P, c_int, c_int, _P, c_int, _P]),
    "b200ddp_conv2d_fwd": (c_int, [_P, _P, c_longlong, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                   c_int, c_int, c_int, c_int, _P]),
    "b200ddp_alexnet_workspace_bytes": (c_longlong, [c_int, c_int, c_int]),
    "b200ddp_alexnet_plan_create": (c_int, [POINTER(c_void_p), POINTER(AlexNetDesc)]),
    "b200ddp_alexnet_plan_destroy": (c_int, [_P]),
    "b200ddp_alexnet_forward": (c_int, [_P, _P, _P, c_int, c_ulonglong, c_ulonglong, c_int, _P]),
    "b200ddp_alexnet_backward": (c_int, [_P, _P, _P, POINTER(c_void_p), POINTER(c_void_p), c_float, READY_FN, _P, c_int, _P]),
    "b200ddp_eval_scratch_bytes": (c_longlong, [c_int]),
    "b200ddp_head_eval": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, _P, _P, c_longlong, c_int, _P]),
    "b200ddp_logits_eval": (c_int, [_P, _P, c_int, c_int, _P, _P, c_longlong, c_int, _P]),
    "b200ddp_resize_coeffs": (c_int, [c_int, c_int, POINTER(c_int)]),
    "b200ddp_image_transform": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, POINTER(c_float),
                                        POINTER(c_float), c_int, _P]),
    "b200ddp_conv2d_fwd_fp32": (c_int, [_P, _P, _P, _P] + [c_int] * 11 + [_P]),
    "b200ddp_conv2d_dgrad_fp32": (c_int, [_P, _P, _P] + [c_int] * 10 + [_P]),
    "b200ddp_conv2d_wgrad_fp32": (c_int, [_P, _P, _P] + [c_int] * 10 + [_P]),
    "b200ddp_conv2d_wgrad_workspace_bytes": (c_longlong, [c_int] * 9),
    "b200ddp_conv2d_wgrad": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                     _P, c_longlong, c_int, _P]),
    "b200ddp_conv2d_weight_prep": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_longlong, c_int, c_int, _P]),
    "b200ddp_space_to_depth": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_conv2d_weight_prep_multi": (c_int, [POINTER(WeightPrepJob), c_int, c_int, _P]),
    "b200ddp_channel_sum_scratch_bytes": (c_longlong, [c_int, c_int]),
    "b200ddp_channel_sum": (c_int, [_P, _P, c_int, c_int, c_int, _P, c_longlong, c_int, _P]),
    "b200ddp_channel_sum_static": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_relu_maxpool_fwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_relu_maxpool_bwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_relu_fwd": (c_int, [_P, c_longlong, c_int, _P]),
    "b200ddp_bias_grad_scratch_bytes": (c_longlong, [c_int, c_int]),
    "b200ddp_relu_maxpool_bwd_bias": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, _P, c_longlong, c_int, _P]),
    "b200ddp_relu_bwd_bias": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P, c_longlong, c_int, _P]),
    "b200ddp_relu_bwd": (c_int, [_P, _P, _P, c_longlong, c_int, _P]),
    "b200ddp_dropout_fwd": (c_int, [_P, _P, _P, c_longlong, c_float, c_ulonglong, c_int, _P]),
    "b200ddp_bn_scratch_bytes": (c_longlong, [c_int, c_int]),
    "b200ddp_bn_stats": (c_int, [_P, c_int, c_int, c_int, c_float, _P, _P, _P, c_longlong, c_int, _P]),
    "b200ddp_bn_gather_stats": (c_int, [_P, c_int, c_int, c_float, c_float, _P, _P, _P, _P, c_int, _P]),
    "b200ddp_bn_elemt": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ddp_bn_bwd_reduce": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, _P, _P, _P, _P, _P, c_longlong, c_int,
                                      _P]),
    "b200ddp_bn_bwd_elemt": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, c_float, _P, c_int, c_int, c_int, c_int, _P]),
}


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = False):
    """Load libb200ddp.so (once). Raises B200Error when it is absent: there is no CPU fallback."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = lib_path()
        if not os.path.exists(path):
            if build_if_missing:
                _build.build()
            else:
                raise B200Error(
                    f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                    "b200ddp has no CPU/eager fallback for its CUDA kernels.")
        l = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(l, name)  # AttributeError here means the .so is stale w.r.t. the header
            fn.restype = res
            fn.argtypes = args
        _LIB = l
    return _LIB


def is_loaded() -> bool:
    return _LIB is not None


def last_error() -> str:
    return load().b200ddp_last_error().decode(errors="replace")


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200Error(f"libb200ddp {what} failed (code {rc}): {last_error()}")


def launch_count() -> int:
    return int(load().b200ddp_launch_count()) if _LIB is not None else 0


def ptr(t) -> int | None:
    """Device pointer of a tensor (None passes NULL)."""
    return None if t is None else t.data_ptr()


def cur_stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def dev_index(t) -> int:
    if not t.is_cuda:
        raise B200Error("b200ddp kernels need CUDA tensors (no CPU fallback); got a tensor on " + str(t.device))
    return t.device.index if t.device.index is not None else torch.cuda.current_device()


def f32c(t, name: str = "tensor"):
    """Contiguous fp32 CUDA tensor or an exception (the kernels do not convert)."""
    if t.dtype != torch.float32:
        raise B200Error(f"{name}: expected float32, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


_WORKSPACES: dict = {}


def workspace(device, nbytes: int, tag: str = "ws", zero: bool = False, stream=None):
    """Per-(device, tag) scratch tensor, grown on demand and reused (the C ABI never allocates).
    A tag is used from ONE stream only. `stream` (a torch.cuda.Stream) = that stream when it is not the current one:
    the buffer is then allocated under it, so that when a larger request replaces the buffer the caching allocator
    reuses the old block in that stream's order (kernels still reading it there are not overtaken)."""
    key = (str(device), tag)
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < nbytes:
        n = max(int(nbytes), 256)
        make = torch.zeros if zero else torch.empty
        if stream is not None:
            with torch.cuda.stream(stream):
                buf = make(n, dtype=torch.uint8, device=device)
        else:
            buf = make(n, dtype=torch.uint8, device=device)
        _WORKSPACES[key] = buf
    return buf


# ---- optional per-entry-point timing (used by bench.py for the roofline of the dominant kernel) ----------------
_TIMING = None   # list of (entry point, int args, start event, end event) while enabled


class _TimedLib:
    """Proxy around the ctypes library: records a CUDA event before and after every kernel-launching entry point,
    on the stream the call is issued on. Only bench.py turns it on; the events cost ~2 us of host time per call."""

    _SKIP = ("_bytes", "last_error", "version", "launch_count", "device_info", "plan_buckets", "nccl_", "comm_create",
             "comm_destroy", "comm_rank", "comm_unique_id")

    def __init__(self, lib):
        object.__setattr__(self, "_lib", lib)

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if _TIMING is None or any(t in name for t in self._SKIP):
            return fn

        def timed(*args):
            # plain integer arguments (sizes, flags) identify the call; pointers / stream handles are far larger
            # (NCCL-allocated arenas sit at virtual addresses of a few tens of GB, so the cut is at 2^31)
            ints = tuple(a for a in args if isinstance(a, int) and not isinstance(a, bool) and 0 <= a < (1 << 31))
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args)
            e1.record()
            _TIMING.append((name, ints, e0, e1))
            return rc
        return timed


_PROXY = None


def enable_timing() -> None:
    """Start recording per-call CUDA events (calls made through _lib.load() after this point)."""
    global _TIMING, _PROXY, _LIB
    lib = load()
    _TIMING = []
    if not isinstance(lib, _TimedLib):
        _PROXY = _TimedLib(lib)
        _LIB = _PROXY


def disable_timing() -> dict:
    """Stop recording; returns [(entry_point, int_args, ms), ...] in call order (synchronises the device)."""
    global _TIMING, _LIB
    torch.cuda.synchronize()
    out = [(name, ints, a.elapsed_time(b)) for name, ints, a, b in (_TIMING or [])]
    _TIMING = None
    if isinstance(_LIB, _TimedLib):
        _LIB = object.__getattribute__(_LIB, "_lib")
    return out
