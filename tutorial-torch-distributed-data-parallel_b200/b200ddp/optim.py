"""O1 — Adam with the step fused into one kernel over the flat arenas.

Drop-in for `optim.Adam(ddp_model.parameters(), lr=0.001)` (reference multi-GPU-training-torch.py:249,
multi-GPU-training-accelerate.py:126): same constructor defaults as torch/optim/adam.py:34-48, a
torch.optim.Optimizer subclass (param_groups, state_dict()/load_state_dict() with per-parameter
`step`, `exp_avg`, `exp_avg_sq`), `zero_grad()` and `step()`.

When the parameters live in a FlatArena (created by b200ddp's DistributedDataParallel or by
`flatten_parameters`) the update is ONE launch of b200ddp_adam_step_flat over the whole arena:
28 bytes per parameter instead of the ~72 bytes of the 7-kernel foreach chain.
Loose CUDA parameters get one launch each. CPU parameters run the same arithmetic with torch ops
(plumbing mode for the gloo tests; no kernels).
"""
from __future__ import annotations

import math
import os
import weakref
from typing import Optional

import torch
from torch.optim import Optimizer

from . import _lib
from .arena import FlatArena, arena_of


def flatten_parameters(module_or_params) -> FlatArena:
    """Move parameters into a FlatArena without wrapping in DDP (single-GPU use of the fused step)."""
    params = list(module_or_params.parameters()) if hasattr(module_or_params, "parameters") else list(module_or_params)
    params = [p for p in params if p.requires_grad]
    existing = arena_of(params)
    return existing if existing is not None else FlatArena(list(reversed(params)))


class Adam(Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 amsgrad: bool = False, *, maximize: bool = False, foreach=None, capturable: bool = False,
                 differentiable: bool = False, fused=None, overlap_backward: Optional[bool] = None) -> None:
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if not 0.0 <= weight_decay:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if amsgrad or maximize or capturable or differentiable:
            raise NotImplementedError("b200ddp Adam implements the reference's configuration only "
                                      "(amsgrad/maximize/capturable/differentiable are off in the tutorial)")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None)
        super().__init__(params, defaults)
        # Step inside backward: when the parameters are the arena of a b200ddp DistributedDataParallel, the wrapper's
        # reducer launches this optimizer's kernel per gradient bucket as each bucket's all-reduce completes, overlapped
        # with the rest of backward; `step()` then only counts the step. `.grad` is untouched, parameters are final
        # when backward() returns (so gradient surgery BETWEEN backward() and step() has no effect: switch it off with
        # overlap_backward=False or B200DDP_ADAM_IN_BACKWARD=0 if you do that). Identical parameters either way.
        self.overlap_backward = (os.environ.get("B200DDP_ADAM_IN_BACKWARD", "1") != "0" if overlap_backward is None
                                 else bool(overlap_backward))
        self._registered_with = None
        self._register_with_reducer()

    # ---- step inside backward ---------------------------------------------------------------------------------
    def _register_with_reducer(self) -> None:
        """Offer this optimizer to the reducer that owns the parameters' arena (if any). Called at construction and
        again from step(): reference multi-GPU-training-accelerate.py:126-129 builds the optimizer BEFORE prepare()
        wraps the model, i.e. before the arena exists."""
        if not self.overlap_backward or self._registered_with is not None or len(self.param_groups) != 1:
            return
        arena = arena_of(self.param_groups[0]["params"])
        ref = getattr(arena, "reducer", None) if arena is not None else None
        red = ref() if ref is not None else None
        if red is not None and arena.device.type == "cuda":
            red.fused_opt = weakref.ref(self)
            self._registered_with = ref

    def _can_fuse(self, arena: FlatArena) -> bool:
        if not self.overlap_backward or len(self.param_groups) != 1:
            return False
        return arena_of(self.param_groups[0]["params"]) is arena

    def _shared_step(self, arena: FlatArena):
        arena.ensure_moments()
        first = self._init_state(arena.params[0])
        for p in arena.params[1:]:      # one shared step tensor for all parameters of the arena (they advance together)
            st = self._init_state(p)
            st["step"] = first["step"]
        return first

    def _next_step_count(self, arena: FlatArena) -> int:
        return int(self._shared_step(arena)["step"].item()) + 1

    def _launch_range(self, arena: FlatArena, start: int, end: int, t: int, stream: int) -> None:
        """Adam on arena elements [start, end) with step count t, on `stream` (called by the reducer per bucket)."""
        g = self.param_groups[0]
        (b1, b2) = g["betas"]
        off = 4 * start
        _lib.check(_lib.load().b200ddp_adam_step_flat(arena.p_flat.data_ptr() + off, arena.g_flat.data_ptr() + off,
                                                      arena.m_flat.data_ptr() + off, arena.v_flat.data_ptr() + off,
                                                      end - start, float(g["lr"]), float(b1), float(b2), float(g["eps"]),
                                                      float(g["weight_decay"]), int(t), None, _lib.dev_index(arena.p_flat),
                                                      stream), "adam_step_flat(bucket)")

    # ---- state ------------------------------------------------------------------------------------------
    def _init_state(self, p) -> dict:
        st = self.state[p]
        if len(st) == 0:
            slot = getattr(p, "_b200_slot", None)
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            if slot is not None:
                st["exp_avg"], st["exp_avg_sq"] = slot[0].moment_views(slot[1])
            else:
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        # torch re-materialises the state tensors; move them back into the arenas the kernel updates
        with torch.no_grad():
            for p, st in self.state.items():
                slot = getattr(p, "_b200_slot", None)
                if "step" in st:
                    st["step"] = st["step"].detach().to("cpu", torch.float32)
                if slot is not None and "exp_avg" in st:
                    m, v = slot[0].moment_views(slot[1])
                    m.copy_(st["exp_avg"])
                    v.copy_(st["exp_avg_sq"])
                    st["exp_avg"], st["exp_avg_sq"] = m, v

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        from .nn import join_side_work
        join_side_work()
        self._register_with_reducer()
        red = self._registered_with() if self._registered_with is not None else None
        if red is not None and red.fused_applied:
            # the reducer already ran this step's update bucket by bucket during backward (and joined the streams)
            red.fused_applied = False
            self._shared_step(red.arena)["step"] += 1
            return loss
        for group in self.param_groups:
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            lr, (b1, b2), eps, wd = group["lr"], group["betas"], group["eps"], group["weight_decay"]
            arena = arena_of(group["params"])
            if (arena is not None and len(params) == len(arena.params) and arena.device.type == "cuda"
                    and all(arena.grad_in_arena(i) for i in range(len(arena.params)))):
                self._step_flat(arena, group, lr, b1, b2, eps, wd)
                continue
            for p in params:
                st = self._init_state(p)
                st["step"] += 1
                t = int(st["step"].item())
                g = p.grad
                if p.is_cuda:
                    self._step_tensor(p, g, st["exp_avg"], st["exp_avg_sq"], t, lr, b1, b2, eps, wd)
                else:
                    _adam_host(p, g, st["exp_avg"], st["exp_avg_sq"], t, lr, b1, b2, eps, wd)
        return loss

    def _step_flat(self, arena: FlatArena, group, lr, b1, b2, eps, wd) -> None:
        first = self._shared_step(arena)
        first["step"] += 1
        t = int(first["step"].item())
        L = _lib.load()
        _lib.check(L.b200ddp_adam_step_flat(arena.p_flat.data_ptr(), arena.g_flat.data_ptr(),
                                            arena.m_flat.data_ptr(), arena.v_flat.data_ptr(), arena.numel,
                                            float(lr), float(b1), float(b2), float(eps), float(wd), t, None,
                                            _lib.dev_index(arena.p_flat), _lib.cur_stream(arena.device)),
                   "adam_step_flat")

    @staticmethod
    def _step_tensor(p, g, m, v, t, lr, b1, b2, eps, wd) -> None:
        L = _lib.load()
        if not (p.is_contiguous() and m.is_contiguous() and v.is_contiguous()):
            raise _lib.B200Error("b200ddp Adam needs contiguous parameters")
        g = _lib.f32c(g, "grad")
        if any(x.data_ptr() % 16 for x in (p, g, m, v)):
            raise _lib.B200Error("b200ddp Adam: parameter storage must be 16-byte aligned")
        _lib.check(L.b200ddp_adam_step_flat(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(),
                                            float(lr), float(b1), float(b2), float(eps), float(wd), t, None,
                                            _lib.dev_index(p), _lib.cur_stream(p.device)), "adam_step_flat")


def _adam_host(p, g, m, v, t, lr, b1, b2, eps, wd) -> None:
    """Same arithmetic as the kernel, with torch ops (CPU plumbing mode only)."""
    if wd != 0:
        g = g.add(p, alpha=wd)
    m.lerp_(g, 1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** t
    bc2 = 1 - b2 ** t
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-(lr / bc1))
