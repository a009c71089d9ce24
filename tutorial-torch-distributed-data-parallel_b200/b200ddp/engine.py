"""T1 — the native step driver behind `model(inputs)` / `loss.backward()`.

The reference's train() (multi-GPU-training-torch.py:118-126) calls `model(inputs)` and `loss.backward()`; with the
per-layer modules of b200ddp.nn each of those walks ~25 Python autograd nodes and ~60 ctypes calls — about 2.0 ms of
host time per step against 2.3 ms of GPU time (measured on B200, bench.py `host_enqueue_ms_per_step`). For the
standard toy model this module replaces the walk by ONE autograd node whose forward and backward are one C call
each (csrc/engine.cu, b200ddp_alexnet_forward / _backward): the same kernels, in the same order, out of one
step-persistent workspace (no per-step tensor allocation). The DDP reducer is told about finished gradient groups
from inside the native backward (a ctypes callback), so the bucket all-reduces still overlap the rest of backward.

Used automatically by `b200ddp.models.AlexNet.forward` when it applies (CUDA, TF32 mode, 3x224x224 input, no
BatchNorm variant, every parameter trainable, no gradient wanted for the input); otherwise the module path runs.
`B200DDP_ENGINE=0` switches it off. Results are bit-identical either way (tests/test_gpu_engine.py).
"""
from __future__ import annotations

import ctypes
import os
import weakref
from typing import Optional

import torch

from . import _lib
from . import nn as bnn


def enabled() -> bool:
    return os.environ.get("B200DDP_ENGINE", "1") != "0"


class StepPlan:
    """Native plan for one (model, batch size): parameter pointers + a workspace holding every activation, gate and
    gradient temporary of a step (about 0.7 GB at batch 128)."""

    def __init__(self, model, batch: int) -> None:
        L = _lib.load()
        convs = [m for m in model.features if isinstance(m, bnn.Conv2d)]
        lins = [m for m in model.classifier if isinstance(m, bnn.Linear)]
        self.layers = convs + lins                       # conv1..conv5, fc1, fc2, head
        self.params = [p for m in self.layers for p in (m.weight, m.bias)]
        dev = self.params[0].device
        self.device = dev
        self.batch = int(batch)
        self.hidden = lins[0].out_features
        self.classes = lins[2].out_features
        self.dropout = float(lins[0].dropout)
        nbytes = int(L.b200ddp_alexnet_workspace_bytes(self.batch, self.hidden, self.classes))
        self.workspace = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
        base = (self.workspace.data_ptr() + 255) // 256 * 256
        self.ktab = bnn._conv_tables(dev, 3, 224, 224, 11, 11)
        d = _lib.AlexNetDesc()
        d.B, d.hidden, d.num_classes, d.dropout_p = self.batch, self.hidden, self.classes, self.dropout
        for i, m in enumerate(self.layers):
            d.w[i] = m.weight.data_ptr()
            d.b[i] = m.bias.data_ptr()
        d.conv1_ktab = self.ktab.data_ptr()
        self.conv1_order = 2 if bnn._use_slab(3, 11, 11, 4, 64, 224, 2) else 0
        d.conv1_order = self.conv1_order
        d.workspace, d.workspace_bytes = base, nbytes
        d.device = dev.index if dev.index is not None else torch.cuda.current_device()
        self.signature = tuple(p.data_ptr() for p in self.params)
        self.handle = ctypes.c_void_p()
        _lib.check(L.b200ddp_alexnet_plan_create(ctypes.byref(self.handle), ctypes.byref(d)), "alexnet_plan_create")
        self._live = None               # weakref to the autograd node of the last training forward, until its backward
        self._done = True
        self._dw = (ctypes.c_void_p * 8)()
        self._db = (ctypes.c_void_p * 8)()
        self._cb = None
        self._cb_error = None

    def busy(self) -> bool:
        """A training forward whose graph is still alive and has not run backward owns the workspace."""
        return (not self._done) and self._live is not None and self._live() is not None

    def __del__(self):
        try:
            if self.handle:
                _lib.load().b200ddp_alexnet_plan_destroy(self.handle)
                self.handle = ctypes.c_void_p()
        except Exception:
            pass


def applicable(model, x: torch.Tensor) -> bool:
    if not enabled() or getattr(model, "bn", False) or not x.is_cuda or x.dtype != torch.float32:
        return False
    if x.dim() != 4 or tuple(x.shape[1:]) != (3, 224, 224) or x.requires_grad:
        return False
    ok = getattr(model, "_engine_ok", None)
    if ok is None:
        convs = [m for m in model.features if isinstance(m, bnn.Conv2d)]
        lins = [m for m in model.classifier if isinstance(m, bnn.Linear)]
        geo = [(3, 64, 11, 4, 2), (64, 192, 5, 1, 2), (192, 384, 3, 1, 1), (384, 256, 3, 1, 1), (256, 256, 3, 1, 1)]
        ok = (len(convs) == 5 and len(lins) == 3 and lins[2].out_features <= 16 and lins[0].in_features == 9216
              and all(m.mode == bnn.MODE_TF32 for m in convs + lins)
              and all((c.in_channels, c.out_channels, c.kernel_size[0], c.stride[0], c.padding[0]) == g
                      for c, g in zip(convs, geo))
              and all(m.bias is not None for m in convs + lins) and lins[1].dropout == 0.0 and lins[2].dropout == 0.0
              and not bnn._use_s2d(3, 11, 11, 4))
        model._engine_ok = ok
    if not ok:
        return False
    return all(p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() for p in model.parameters())


def plan_for(model, batch: int) -> Optional[StepPlan]:
    plans = model.__dict__.setdefault("_engine_plans", {})
    plan = plans.get(batch)
    if plan is not None:
        if (plan.signature != tuple(p.data_ptr() for p in plan.params) or plan.dropout != float(plan.layers[5].dropout)
                or plan.conv1_order != (2 if bnn._use_slab(3, 11, 11, 4, 64, 224, 2) else 0)):
            plan = None                                   # parameters were re-pointed (DDP wrap, .to()): rebuild
    if plan is None:
        if len(plans) >= 3:                               # full batch, ragged last batch, evaluation batch
            plans.pop(next(iter(plans)))
        plan = plans[batch] = StepPlan(model, batch)
    return plan


StepPlan.__deepcopy__ = lambda self, memo: None          # a copied model builds its own plans


class _AlexNetFn(torch.autograd.Function):
    """The whole model as one autograd node. Inputs: plan, x, p_used, seeds, then the 16 parameters (so that autograd
    routes their gradients); output: logits."""

    @staticmethod
    def forward(ctx, plan, x, p_used, seed0, seed1, *params):
        L = _lib.load()
        out = torch.empty((plan.batch, plan.classes), dtype=torch.float32, device=x.device)
        st = _lib.cur_stream(x.device)
        train = any(ctx.needs_input_grad[5:])
        _lib.check(L.b200ddp_alexnet_forward(plan.handle, x.data_ptr(), out.data_ptr(), int(train), seed0, seed1, 0, st),
                   "alexnet_forward")
        if train:
            ctx.plan = plan
            ctx.p_used = p_used
            ctx.save_for_backward(x)
            plan._live = weakref.ref(ctx)
            plan._done = False
        return out

    @staticmethod
    def backward(ctx, dlogits):
        L = _lib.load()
        plan = ctx.plan
        (x,) = ctx.saved_tensors
        if plan._live is None or plan._live() is not ctx:
            raise _lib.B200Error("b200ddp engine: the activations of this forward pass were overwritten by a later "
                                 "forward pass of the same model and batch size")
        dlogits = _lib.f32c(dlogits, "grad_output")
        dev = x.device
        # where each gradient is written: the parameter's arena slice when autograd will adopt it, else a fresh buffer
        grads, adopted_all = [], True
        for i, p in enumerate(plan.params):
            t, adopted = bnn._grad_target(p, tuple(p.shape), dev)
            grads.append(t)
            adopted_all = adopted_all and adopted
            (plan._dw if i % 2 == 0 else plan._db)[i // 2] = t.data_ptr()
        # gradient groups -> the reducer (only when the arena slices ARE the gradients: nothing to accumulate afterwards)
        cb = _NO_CB
        arena = getattr(plan.params[0], "_b200_slot", (None,))[0]
        ref = getattr(arena, "reducer", None) if arena is not None else None
        red = ref() if ref is not None else None
        if adopted_all and red is not None and red.require_sync and red.on_gpu:
            if plan._cb is None or plan._cb[0] is not red:
                idx = [arena.index_of(p) for p in plan.params]
                groups = ([idx[14], idx[15], idx[12], idx[13]], [idx[10], idx[11]], idx[2:10], idx[0:2])

                def ready(_user, g, red=red, groups=groups, plan=plan):
                    try:                                  # an exception must not vanish inside the C callback
                        if plan._cb_error is None:
                            red.mark_ready_native(groups[g])
                    except BaseException as e:
                        plan._cb_error = e
                plan._cb = (red, _lib.READY_FN(ready))
            cb = plan._cb[1]
        use_side = 1 if (adopted_all and bnn._overlap_wgrad()) else 0
        _lib.check(L.b200ddp_alexnet_backward(plan.handle, x.data_ptr(), dlogits.data_ptr(), plan._dw, plan._db,
                                              float(ctx.p_used), cb, None, use_side, _lib.cur_stream(dev)),
                   "alexnet_backward")
        plan._done = True
        if plan._cb_error is not None:
            e, plan._cb_error = plan._cb_error, None
            raise e
        return (None, None, None, None, None, *grads)


_NO_CB = ctypes.cast(None, _lib.READY_FN)


def run(model, x: torch.Tensor, stop_at_hidden: bool = False) -> torch.Tensor:
    """Forward of `model` on `x` through the native plan; the caller has checked `applicable(model, x)`."""
    x = x if x.is_contiguous() else x.contiguous()
    plan = plan_for(model, int(x.shape[0]))
    grad_mode = torch.is_grad_enabled() and any(p.requires_grad for p in plan.params)
    if grad_mode and (stop_at_hidden or plan.busy() or not all(p.requires_grad for p in plan.params)):
        return None                                       # the module path handles it
    if not grad_mode:
        if plan.busy():
            return None                                   # evaluation between forward and backward: keep the activations
        L = _lib.load()
        n = plan.hidden if stop_at_hidden else plan.classes
        out = torch.empty((plan.batch, n), dtype=torch.float32, device=x.device)
        _lib.check(L.b200ddp_alexnet_forward(plan.handle, x.data_ptr(), out.data_ptr(), 0, 0, 0, int(stop_at_hidden),
                                             _lib.cur_stream(x.device)), "alexnet_forward")
        return out
    p = plan.dropout if model.training else 0.0
    seed0 = bnn._next_seed() if p > 0.0 else 0            # same draws, same order as the module path
    seed1 = bnn._next_seed() if p > 0.0 else 0
    return _AlexNetFn.apply(plan, x, p, seed0, seed1, *plan.params)
