"""R1-R4 — gradient bucketing, clip-before-aggregation and overlapped all-reduce.

Plays the role of c10d::Reducer (torch reducer.hpp:45-584) behind the DistributedDataParallel
wrapper that reference multi-GPU-training-torch.py:245 constructs:

* bucket plan: parameters in reverse registration order (the order their gradients become ready,
  SURVEY 3.2), greedy byte caps [1 MiB, 25 MiB] — computed by the native planner
  (b200ddp_plan_buckets) which restates compute_bucket_assignment_by_size. Because the gradient arena
  uses the same order, every bucket is one contiguous range reduced in place.
* per-parameter post-accumulate-grad hooks mark readiness; a bucket whose parameters are all ready
  is (clipped and) all-reduced on the communication stream while backward keeps running; buckets are
  launched strictly in index order so all ranks issue identical NCCL sequences.
* averaging is folded into the collective (PreMulSum 1/world_size) — no separate divide pass.
* clipping happens on the LOCAL gradients before aggregation (reference README.md:92-95):
  value mode clamps each bucket in place right before its all-reduce (overlap preserved);
  norm mode accumulates per-bucket sums of squares during backward, computes
  coef = min(1, c/(||g||+1e-6)) on the device when backward ends, scales and then reduces.

With CPU tensors (gloo) the same host logic runs with torch ops in place of the kernels — that
"plumbing mode" exists only so the control flow can be tested without a GPU.
"""
from __future__ import annotations

import ctypes
import os
import weakref
from typing import List, Optional

import torch
import torch.distributed as dist
from torch.autograd import Variable

from . import _lib
from .arena import FlatArena

MiB = 1 << 20


def plan_buckets(nbytes: List[int], caps: List[int], keys: Optional[List[int]] = None) -> List[List[int]]:
    """Bucket index lists for tensors of `nbytes` (native planner, host only)."""
    L = _lib.load()
    n = len(nbytes)
    a = (ctypes.c_longlong * n)(*nbytes)
    c = (ctypes.c_longlong * len(caps))(*caps)
    k = (ctypes.c_int * n)(*keys) if keys is not None else None
    out = (ctypes.c_int * max(n, 1))()
    nb = ctypes.c_int(0)
    _lib.check(L.b200ddp_plan_buckets(a, k, n, c, len(caps), out, ctypes.byref(nb)), "plan_buckets")
    buckets: List[List[int]] = [[] for _ in range(nb.value)]
    for i in range(n):
        buckets[out[i]].append(i)
    return buckets


class _Bucket:
    __slots__ = ("index", "members", "start", "end", "pending", "ready", "launched", "work")

    def __init__(self, index: int, members: List[int], start: int, end: int) -> None:
        self.index = index
        self.members = members
        self.start = start
        self.end = end
        self.pending = len(members)
        self.ready = False
        self.launched = False
        self.work = None


class Reducer:
    def __init__(self, arena: FlatArena, process_group=None, bucket_cap_mb: float = 25.0,
                 first_bucket_cap_mb: float = 1.0, clip_grad_value: Optional[float] = None,
                 clip_grad_norm: Optional[float] = None, nonfinite: str = "propagate", comm=None) -> None:
        if clip_grad_value is not None and clip_grad_norm is not None:
            raise ValueError("choose clip_grad_value or clip_grad_norm, not both")
        if nonfinite not in ("propagate", "zero"):
            raise ValueError("nonfinite must be 'propagate' or 'zero'")
        self.arena = arena
        self.pg = process_group
        self.on_gpu = arena.device.type == "cuda"
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.comm = comm
        self.clip_value = None if clip_grad_value is None else float(clip_grad_value)
        self.clip_norm = None if clip_grad_norm is None else float(clip_grad_norm)
        # Which all-reduce runs (B200DDP_ALLREDUCE):
        #   own  (default) this library's kernel over the registered arena's NVSwitch multicast / peer mappings
        #        (peer_allreduce.cu), 1/world folded in; needs comm.arena_mode >= 1, else falls back to ring;
        #   ring NCCL all-reduce with a PreMulSum(1/world) operator (NCCL runs user-defined operators as RING);
        #   sum  NCCL ncclSum on gradients pre-scaled by 1/world at the loss (see DistributedDataParallel.forward), which
        #        NCCL may run as NVLS; the clip-before-aggregation modes never pre-scale (thresholds act on raw gradients).
        self.algo = "ring"
        if self.on_gpu and self.world > 1 and comm is not None:
            from . import comm as _comm_mod
            want = os.environ.get("B200DDP_ALLREDUCE", "own")
            in_arena = getattr(comm, "arena_mode", 0) >= 1 and any(t.data_ptr() == arena.g_flat.data_ptr()
                                                                    for _, t in getattr(comm, "_arenas", []))
            # Shape of the own kernel, measured on 2 and 8 B200s (profiles/r02_allreduce_sweep.md): a few 1024-thread CTAs
            # that own their SM (the compute grids are planned around them) beat many small co-resident ones for the step
            # time, although the small ones reach a higher bus bandwidth: 8 CTAs at 2 ranks (each rank moves half of every
            # bucket through its SMs), 4 at 8 ranks (an eighth).
            big = os.environ.get("B200DDP_AR_BIG", "1") != "0"
            if big:
                dflt = 8 if self.world <= 2 else (6 if self.world <= 4 else 4)
            else:
                dflt = 64 if self.world < 4 else (32 if arena.numel * 4 > (400 << 20) else 16)
            ctas = int(os.environ.get("B200DDP_AR_CTAS", str(dflt)))
            if want in ("own", "peer") and in_arena and comm.set_algo("own", ctas):
                self.algo = "own"
                _lib.load().b200ddp_comm_set_cta_shape(comm.handle, 1 if big else 0)
            elif want == "sum":
                comm.set_algo("sum")
                self.algo = "sum"
            else:
                comm.set_algo("ring")
            _comm_mod.reserve_sms_for_nccl(self.algo != "own")
            if self.algo == "own" and big:       # 1024-thread CTAs own their SM: plan the compute grids around them
                n = ctypes.c_int(0)
                _lib.check(_lib.load().b200ddp_device_info(_lib.dev_index(arena.g_flat), ctypes.byref(n), None, None),
                           "device_info")
                _lib.load().b200ddp_set_sm_budget(max(16, n.value - ctas))
        self._chunk_elems = max(1, int(os.environ.get("B200DDP_AR_CHUNK_MB", "64"))) << 18      # MB -> fp32 elements
        self.pass_prescaled = False
        self.nonfinite_zero = nonfinite == "zero"
        self.prescale = False
        self._update_prescale()
        self.require_sync = True
        # optimizer step per bucket during backward (set by b200ddp.optim.Adam, see _fused_adam)
        self.fused_opt = None
        self.fresh_grads = False
        self._fused_now = None
        self._fused_t = 0
        self.fused_applied = False
        self.opt_stream = torch.cuda.Stream(arena.device) if self.on_gpu else None
        # measurement knob (bench.py): run the step with every collective launch skipped - same kernels, same streams,
        # no communication - so that "exposed communication" = step time - dry step time. Results are then local.
        self.dry_run_collectives = False
        arena.reducer = weakref.ref(self)
        self.last_local_norm: Optional[torch.Tensor] = None

        params = arena.params
        sizes = [p.numel() * 4 for p in params]
        caps = [int(first_bucket_cap_mb * MiB), int(bucket_cap_mb * MiB)]
        if self.on_gpu or _lib_available():
            groups = plan_buckets(sizes, caps)
        else:  # pragma: no cover - only when the .so has not been built on a CPU-only box
            raise _lib.B200Error("libb200ddp.so is required for the bucket planner")
        self.buckets: List[_Bucket] = []
        for bi, members in enumerate(groups):
            assert members == list(range(members[0], members[-1] + 1)), "buckets must be contiguous in arena order"
            start, end = arena.span(members[0], members[-1])
            self.buckets.append(_Bucket(bi, members, start, end))
        self.bucket_of = [0] * len(params)
        for b in self.buckets:
            for i in b.members:
                self.bucket_of[i] = b.index
        self.next_bucket = 0
        self._callback_queued = False
        self._native_done = set()
        self._ready_order: List[int] = []
        self.first_ready_order: Optional[List[int]] = None
        # Early launch inside the LAST bucket: its final members (the first layer's weight and bias) become ready only
        # when the whole backward pass is over, so a bucket-granular launch leaves the all-reduce of everything else in
        # that bucket (the other conv layers, ~10 MB) exposed after the last kernel. The bucket stays one bucket (same
        # plan and sizes as the reference's reducer); its all-reduce is issued as two ranges: the prefix as soon as its
        # members are ready - it then runs under the first layer's weight-gradient kernel - and the tail at the end.
        self._tail_prefix = None           # (bucket index, set of prefix members, prefix end offset)
        self._tail_prefix_pending = 0
        self._tail_early = False
        self._update_tail_prefix()
        if self.on_gpu:
            dev = arena.device
            self._sq = torch.zeros(1, dtype=torch.float32, device=dev)
            self._coef = torch.ones(1, dtype=torch.float32, device=dev)
            self._norm = torch.zeros(1, dtype=torch.float32, device=dev)
            nbytes = int(_lib.load().b200ddp_sqnorm_scratch_bytes())
            self._sq_scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        self._hooks = [p.register_post_accumulate_grad_hook(self._make_hook(i)) for i, p in enumerate(params)]

    def _update_prescale(self) -> None:
        self.prescale = (self.on_gpu and self.clip_value is None and self.clip_norm is None and self.algo == "sum")

    def _update_tail_prefix(self) -> None:
        self._tail_prefix = None
        arena = self.arena
        if (self.on_gpu and self.world > 1 and self.clip_value is None and self.clip_norm is None and self.buckets
                and os.environ.get("B200DDP_TAIL_SPLIT", "1") != "0"):
            last = self.buckets[-1]
            if len(last.members) > 2:
                prefix = last.members[:-2]
                p_start, p_end = arena.span(prefix[0], prefix[-1])
                if (p_end - p_start) * 4 >= (1 << 20):
                    self._tail_prefix = (last.index, frozenset(prefix), p_end)

    def configure_clipping(self, clip_grad_value=None, clip_grad_norm=None, nonfinite: str = "propagate") -> None:
        if clip_grad_value is not None and clip_grad_norm is not None:
            raise ValueError("choose clip_grad_value or clip_grad_norm, not both")
        if nonfinite not in ("propagate", "zero"):
            raise ValueError("nonfinite must be 'propagate' or 'zero'")
        self.clip_value = None if clip_grad_value is None else float(clip_grad_value)
        self.clip_norm = None if clip_grad_norm is None else float(clip_grad_norm)
        self.nonfinite_zero = nonfinite == "zero"
        self._update_prescale()
        self._update_tail_prefix()

    # ---- optimizer step inside backward (the idea behind upstream's _register_fused_optim,
    #      torch/nn/parallel/distributed.py:2097) ------------------------------------------------------------------
    def _fused_adam(self, start: int, end: int) -> None:
        """Adam on arena range [start, end) as soon as that range's averaged gradient exists: on a second stream, after
        the range's all-reduce (or, single rank, after its producers), while backward goes on computing the earlier
        layers' gradients. Same kernel and arithmetic as the one-launch step: parameters come out bit-identical."""
        opt = self._fused_now
        if opt is None or end <= start:
            return
        st = self.opt_stream
        if self.world > 1 and not self.dry_run_collectives:
            self.comm.bucket_wait_last(st.cuda_stream)
        else:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.arena.device))
            st.wait_event(ev)
        opt._launch_range(self.arena, start, end, self._fused_t, st.cuda_stream)

    # ---- public ------------------------------------------------------------------------------------
    def bucket_sizes_bytes(self) -> List[int]:
        return [sum(self.arena.params[i].numel() * 4 for i in b.members) for b in self.buckets]

    def bucket_param_indices(self) -> List[List[int]]:
        return [list(b.members) for b in self.buckets]

    def remove_hooks(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []

    # ---- backward-time machinery ---------------------------------------------------------------------
    def _make_hook(self, i: int):
        def hook(param):
            self._on_grad_ready(i, param)
        return hook

    def _reset(self) -> None:
        for b in self.buckets:
            b.pending = len(b.members)
            b.ready = False
            b.launched = False
            b.work = None
        self.next_bucket = 0
        self._callback_queued = False
        self._ready_order = []
        self._native_done = set()
        self._tail_early = False
        self._tail_prefix_pending = len(self._tail_prefix[1]) if self._tail_prefix is not None else 0
        # fused optimizer: only for a backward pass that starts from empty gradients (no accumulation pending)
        opt = self.fused_opt() if self.fused_opt is not None else None
        self._fused_now = None
        self.fused_applied = False
        if opt is not None and self.on_gpu and self.fresh_grads and opt._can_fuse(self.arena):
            self._fused_now = opt
            self._fused_t = opt._next_step_count(self.arena)
        self.fresh_grads = False

    def mark_ready_native(self, indices) -> None:
        """Called from inside the native backward (b200ddp.engine): the gradients of these arena parameters have just been
        enqueued, written straight into their arena slices. Their autograd hooks, which fire later in the same backward
        pass, are then ignored."""
        for i in indices:
            self._on_grad_ready(i, None, native=True)

    def _on_grad_ready(self, i: int, param, native: bool = False) -> None:
        if not self.require_sync:
            return
        if self._callback_queued and i in self._native_done:
            return
        if not self._callback_queued:
            self._reset()
            self._callback_queued = True
            if self.on_gpu and self.clip_norm is not None:
                self._sq.zero_()
            Variable._execution_engine.queue_callback(self._finalize)
        self._ready_order.append(i)
        arena = self.arena
        if native:
            self._native_done.add(i)
        elif not arena.grad_in_arena(i):
            # gradient was produced outside the arena (e.g. by a cuDNN wgrad): pack it into its slice
            g = param.grad
            view = arena.grad_view(i)
            if self.on_gpu:
                L = _lib.load()
                src = g if g.is_contiguous() else g.contiguous()
                _lib.check(L.b200ddp_pack_scale_clip(view.data_ptr(), src.data_ptr(), src.numel(), 1.0, 0.0, 0,
                                                     _lib.dev_index(view), _lib.cur_stream(view.device)),
                           "pack_scale_clip")
            else:
                view.copy_(g)
            param.grad = view
        b = self.buckets[self.bucket_of[i]]
        b.pending -= 1
        if self._tail_prefix is not None and b.index == self._tail_prefix[0] and i in self._tail_prefix[1]:
            self._tail_prefix_pending -= 1
            if (self._tail_prefix_pending == 0 and not self._tail_early and b.pending > 0
                    and self.next_bucket == b.index):      # every earlier bucket is already on the comm stream
                from .nn import join_side_work
                join_side_work(self.arena.device)
                self._tail_early = True
                if not self.dry_run_collectives:
                    self.comm.bucket_reduce_launch(self.arena.g_flat[b.start:self._tail_prefix[2]].data_ptr(),
                                                   self._tail_prefix[2] - b.start,
                                                   1.0 if self.pass_prescaled else 1.0 / self.world, self.arena.device)
                self._fused_adam(b.start, self._tail_prefix[2])
        if b.pending == 0:
            b.ready = True
            while self.next_bucket < len(self.buckets) and self.buckets[self.next_bucket].ready:
                self._bucket_ready(self.buckets[self.next_bucket])
                self.next_bucket += 1

    def _bucket_tensor(self, b: _Bucket) -> torch.Tensor:
        return self.arena.g_flat[b.start:b.end]

    def _bucket_ready(self, b: _Bucket) -> None:
        t = self._bucket_tensor(b)
        n = b.end - b.start
        if self.on_gpu:
            from .nn import join_side_work
            join_side_work(t.device)       # weight gradients computed on the side stream must be complete
            L = _lib.load()
            dev, st = _lib.dev_index(t), _lib.cur_stream(t.device)
            if self.clip_value is not None:
                _lib.check(L.b200ddp_pack_scale_clip(t.data_ptr(), t.data_ptr(), n, 1.0, self.clip_value,
                                                     1 if self.nonfinite_zero else 0, dev, st), "clip_value")
            if self.clip_norm is not None:
                _lib.check(L.b200ddp_sqnorm(t.data_ptr(), n, self._sq.data_ptr(), 1, self._sq_scratch.data_ptr(),
                                            self._sq_scratch.numel(), dev, st), "sqnorm")
                return  # all-reduce deferred until the full local norm is known (see _finalize)
            self._launch_allreduce(b)
        else:
            if self.clip_value is not None:
                if self.nonfinite_zero:
                    torch.nan_to_num_(t, nan=0.0)
                t.clamp_(-self.clip_value, self.clip_value)
            if self.clip_norm is not None:
                return
            self._launch_allreduce(b)

    def _launch_allreduce(self, b: _Bucket) -> None:
        b.launched = True
        if self.world == 1:
            if self.on_gpu:
                self._fused_adam(b.start, b.end)
            return
        t = self._bucket_tensor(b)
        if self.on_gpu:
            premul = 1.0 if self.pass_prescaled else 1.0 / self.world
            start = b.start
            if self._tail_early and self._tail_prefix is not None and b.index == self._tail_prefix[0]:
                start = self._tail_prefix[2]                           # the prefix is already being reduced
                t = self.arena.g_flat[start:b.end]
            # Large buckets go out in chunks: the optimizer launch of chunk k (it waits for that chunk's collective only)
            # then overlaps the collective of chunk k+1, and reads the averaged gradients while they are still in L2.
            # Without this the Adam pass over a 300 MB bucket can only start when its whole all-reduce has finished.
            chunk = self._chunk_elems if self._fused_now is not None else (b.end - start)
            s = start
            while s < b.end:
                e = min(s + chunk, b.end)
                if not self.dry_run_collectives:
                    self.comm.bucket_reduce_launch(self.arena.g_flat.data_ptr() + 4 * s, e - s, premul, t.device)
                self._fused_adam(s, e)
                s = e
        else:
            t.mul_(1.0 / self.world)
            b.work = dist.all_reduce(t, group=self.pg, async_op=True)

    def _finalize(self) -> None:
        """Runs once when the autograd engine has finished this backward pass."""
        try:
            if not all(b.ready for b in self.buckets):
                missing = [i for b in self.buckets if not b.ready for i in b.members
                           if i not in self._ready_order]
                raise RuntimeError(
                    "b200ddp Reducer: backward finished but parameters "
                    f"{missing} produced no gradient (unused parameters are not supported)")
            if self.first_ready_order is None:
                self.first_ready_order = list(self._ready_order)
            if self.clip_norm is not None:
                self._clip_by_norm_then_reduce()
            if self.on_gpu and self._fused_now is not None:
                # every range has had its Adam launch: the compute stream joins the optimizer stream here, so that
                # parameters are final for whatever the caller enqueues next; optimizer.step() only counts the step
                done = torch.cuda.Event()
                done.record(self.opt_stream)
                torch.cuda.current_stream(self.arena.device).wait_event(done)
                self.fused_applied = True
                self._fused_now = None
            if self.world > 1:
                if self.on_gpu:
                    self.comm.bucket_reduce_join(self.arena.device)
                else:
                    for b in self.buckets:
                        if b.work is not None:
                            b.work.wait()
        finally:
            self._callback_queued = False

    def _clip_by_norm_then_reduce(self) -> None:
        g = self.arena.g_flat
        if self.on_gpu:
            L = _lib.load()
            dev, st = _lib.dev_index(g), _lib.cur_stream(g.device)
            _lib.check(L.b200ddp_clip_coef(self._sq.data_ptr(), self.clip_norm, 1 if self.nonfinite_zero else 0,
                                           self._coef.data_ptr(), self._norm.data_ptr(), dev, st), "clip_coef")
            for b in self.buckets:
                t = self._bucket_tensor(b)
                _lib.check(L.b200ddp_scale_by_device_scalar(t.data_ptr(), b.end - b.start, self._coef.data_ptr(),
                                                            1 if self.nonfinite_zero else 0, dev, st), "scale")
                self._launch_allreduce(b)
            self.last_local_norm = self._norm
        else:
            norm = torch.linalg.vector_norm(g)
            coef = self.clip_norm / (norm + 1e-6)
            if self.nonfinite_zero and not bool(torch.isfinite(norm)):
                g.zero_()
            else:
                g.mul_(torch.clamp(coef, max=1.0))
            self.last_local_norm = norm
            for b in self.buckets:
                self._launch_allreduce(b)


def _lib_available() -> bool:
    try:
        _lib.load()
        return True
    except Exception:
        return False
