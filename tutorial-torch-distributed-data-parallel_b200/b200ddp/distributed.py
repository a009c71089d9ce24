"""D1 — DistributedDataParallel-compatible wrapper (drop-in for torch.nn.parallel.DistributedDataParallel).

Call site in the reference: `ddp_model = DDP(model, device_ids=[rank])`
(multi-GPU-training-torch.py:245), then `optim.Adam(ddp_model.parameters(), lr=0.001)` (:249),
`model(inputs)` (:121), `loss.backward()` (:125) and `torch.save(model.state_dict(), ...)` on rank 0
(:217-221). Contract kept from torch/nn/parallel/distributed.py:
  * `.module`, `forward(*a, **kw)`, `.parameters()` yields the wrapped module's own Parameter objects,
    `.train()/.eval()` propagate, `state_dict()` keys carry the `module.` prefix (:797), the
    checkpoint loads with `torch.load(path, map_location=...)` + `load_state_dict`;
  * parameters (and buffers) are broadcast from rank 0 at construction (:883-890);
  * after `backward()` every `.grad` holds the MEAN over ranks; `no_sync()` (:1473-1500) accumulates
    locally and the first backward outside reduces the accumulated gradients;
  * forward under `torch.no_grad()` (evaluate(), reference :136-153) does not arm the reducer.
Extra keyword arguments place gradient clipping BEFORE aggregation as the reference README.md:92-95
prescribes: `clip_grad_value=c` or `clip_grad_norm=c`, and `nonfinite="zero"` to drop a rank whose
local gradient is not finite.
"""
from __future__ import annotations

import os
import warnings
from contextlib import contextmanager
from typing import Optional

import torch
import torch.distributed as dist
from torch import nn

from . import _lib
from .arena import FlatArena
from .comm import get_comm
from .reducer import Reducer

_warned_plumbing = False


def _warn_plumbing() -> None:
    global _warned_plumbing
    if not _warned_plumbing:
        _warned_plumbing = True
        warnings.warn("b200ddp: parameters are on the CPU -> host-logic plumbing mode (gloo, no CUDA kernels). "
                      "This mode exists for tests of the control flow only.", stacklevel=3)


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, process_group=None, bucket_cap_mb: float = 25.0,
                 find_unused_parameters: bool = False, check_reduction: bool = False,
                 gradient_as_bucket_view: bool = True, static_graph: bool = False,
                 clip_grad_value: Optional[float] = None, clip_grad_norm: Optional[float] = None,
                 nonfinite: str = "propagate") -> None:
        super().__init__()
        if find_unused_parameters:
            raise NotImplementedError("b200ddp DDP: find_unused_parameters is not supported (the toy model uses "
                                      "every parameter in every step)")
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter "
                               "that requires a gradient.")
        devices = {p.device for p in params}
        if len(devices) != 1:
            raise ValueError("b200ddp DDP supports single-device modules only; found " + str(devices))
        self.device = next(iter(devices))
        if device_ids is not None:
            if len(device_ids) != 1:
                raise ValueError("device_ids must hold exactly one device (one process per GPU)")
            want = torch.device(device_ids[0]) if not isinstance(device_ids[0], int) else torch.device(
                "cuda", device_ids[0])
            if self.device.type != "cuda" or (want.index is not None and want.index != self.device.index):
                raise ValueError(f"device_ids={device_ids} does not match the module's device {self.device}")
        self.module = module
        self.dim = dim
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.world_size = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if self.world_size > 1 else 0
        self.on_gpu = self.device.type == "cuda"
        if self.on_gpu:
            _lib.load()           # raises if the extension is missing: no fallback on a GPU
            torch.cuda.set_device(self.device)
            self.comm = get_comm(self.device, process_group)
        else:
            _warn_plumbing()
            self.comm = None
        # reverse registration order ~ the order in which gradients become ready.
        # With more than one rank the gradient arena lives in NCCL-registered, peer-mapped memory (arena_register,
        # SURVEY 8(b)): the all-reduce then runs as this library's own kernel over the NVSwitch (peer_allreduce.cu).
        grad_alloc = None
        if self.on_gpu and self.comm is not None and os.environ.get("B200DDP_ARENA", "registered") != "plain":
            grad_alloc = self.comm.arena_alloc
        self.arena = FlatArena(list(reversed(params)), grad_alloc=grad_alloc)
        self._sync_module_states()
        self.reducer = Reducer(self.arena, process_group, bucket_cap_mb=bucket_cap_mb,
                               clip_grad_value=clip_grad_value, clip_grad_norm=clip_grad_norm,
                               nonfinite=nonfinite, comm=self.comm)
        self.require_backward_grad_sync = True

    # ---- construction helpers ----------------------------------------------------------------------
    def _sync_module_states(self) -> None:
        if self.world_size == 1:
            return
        self._verify_shapes()
        with torch.no_grad():
            if self.on_gpu:
                self.comm.broadcast_(self.arena.p_flat, 0)
            else:
                dist.broadcast(self.arena.p_flat, src=self._src(), group=self.process_group)
            self._broadcast_buffers()

    def _src(self) -> int:
        return dist.get_global_rank(self.process_group, 0) if self.process_group is not None else 0

    def _verify_shapes(self) -> None:
        mine = [tuple(p.shape) for p in self.arena.params]
        box = [mine]
        dist.broadcast_object_list(box, src=self._src(), group=self.process_group)
        if box[0] != mine:
            raise RuntimeError("DDP: parameter shapes differ between rank 0 and rank %d" % self.rank)

    def _broadcast_buffers(self) -> None:
        for b in self.module.buffers():
            if self.on_gpu and b.is_floating_point() and b.dtype == torch.float32 and b.is_contiguous():
                self.comm.broadcast_(b, 0)
            else:
                dist.broadcast(b, src=self._src(), group=self.process_group)

    # ---- the step ----------------------------------------------------------------------------------------
    def forward(self, *inputs, **kwargs):
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            self.reducer.require_sync = True
            # no gradient pending anywhere: the backward pass that follows may apply the optimizer bucket by bucket
            self.reducer.fresh_grads = all(p.grad is None for p in self.arena.params)
            if self.broadcast_buffers and self.world_size > 1 and self.training and any(
                    True for _ in self.module.buffers()):
                with torch.no_grad():
                    self._broadcast_buffers()
        elif torch.is_grad_enabled():
            self.reducer.require_sync = False
        out = self.module(*inputs, **kwargs)
        self.reducer.pass_prescaled = False
        if (self.reducer.prescale and torch.is_grad_enabled() and self.require_backward_grad_sync
                and self.world_size > 1):
            # Mean over ranks = plain SUM of gradients that are already divided by the world size: scaling the
            # gradient that enters the model (one tiny multiply) lets the all-reduce use ncclSum, which NCCL can run
            # as an in-switch NVLS reduction on NVSwitch systems (user-defined PreMulSum falls back to RING).
            inv = 1.0 / self.world_size
            scaled = [False]

            def _scale(t):
                if torch.is_tensor(t) and t.requires_grad:
                    t.register_hook(lambda g: g * inv)
                    scaled[0] = True
                return t
            if torch.is_tensor(out):
                _scale(out)
            elif isinstance(out, (tuple, list)):
                for t in out:
                    _scale(t)
            elif isinstance(out, dict):
                for t in out.values():
                    _scale(t)
            self.reducer.pass_prescaled = scaled[0]
        return out

    @contextmanager
    def no_sync(self):
        """Backward passes inside accumulate locally; the next one outside reduces the sum."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        self.reducer.require_sync = False
        self.reducer.prescale = False      # locally accumulated gradients are unscaled: keep PreMulSum from here on
        try:
            yield
        finally:
            self.require_backward_grad_sync = old
            self.reducer.require_sync = old

    def set_clipping(self, clip_grad_value: Optional[float] = None, clip_grad_norm: Optional[float] = None,
                     nonfinite: str = "propagate") -> None:
        """Change the clip-before-aggregation mode (README.md:92-95) of an existing wrapper; takes effect at the next
        backward pass. Every rank must make the same call."""
        self.reducer.configure_clipping(clip_grad_value, clip_grad_norm, nonfinite)

    # ---- introspection used by tests / logging ---------------------------------------------------------
    def bucket_sizes(self):
        return self.reducer.bucket_sizes_bytes()

    def bucket_parameter_names(self):
        names = {id(p): n for n, p in self.module.named_parameters()}
        return [[names[id(self.arena.params[i])] for i in b] for b in self.reducer.bucket_param_indices()]
