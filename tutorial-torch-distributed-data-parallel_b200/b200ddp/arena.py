"""Flat parameter / gradient arenas (R2, O1).

Stock DDP copies every gradient into a bucket (x 1/world_size) and back and
foreach-Adam walks 16 separate tensors. Here every parameter's storage is a slice of ONE fp32 arena
and every gradient a slice of a second arena with the same layout, so that
  * a gradient bucket is a contiguous range that NCCL reduces in place,
  * wgrad kernels write straight into their slice (no pack copy),
  * the optimizer is one kernel launch over the whole arena.
Slices start at multiples of 64 elements (256 B) so TMA / float4 accesses stay aligned; the padding
is zero and stays zero. The Parameter objects are kept (only .data is re-pointed) so that
`optim.Adam(ddp.parameters())`, `state_dict()` and `load_state_dict()` behave as with stock torch.
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch

ALIGN_ELEMS = 64


class FlatArena:
    def __init__(self, params: Iterable[torch.nn.Parameter], grad_alloc=None) -> None:
        """`grad_alloc(numel) -> fp32 zeros tensor` (optional) supplies the gradient arena's memory, e.g. NCCL-registered
        peer-mapped memory from NativeComm.arena_alloc; default: a plain torch allocation."""
        self.params: List[torch.nn.Parameter] = list(params)
        if not self.params:
            raise ValueError("FlatArena needs at least one parameter")
        dev = self.params[0].device
        for p in self.params:
            if p.dtype != torch.float32:
                raise TypeError(f"b200ddp arenas are fp32; got {p.dtype}")
            if p.device != dev:
                raise ValueError("all parameters of an arena must live on one device")
        self.device = dev
        self.offsets: List[int] = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        self.numel = off
        self.p_flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.g_flat = grad_alloc(off) if grad_alloc is not None else torch.zeros(off, dtype=torch.float32, device=dev)
        if self.g_flat.numel() != off or self.g_flat.dtype != torch.float32 or self.g_flat.device != dev:
            raise ValueError("grad_alloc returned a buffer of the wrong size / dtype / device")
        # optimizer moments are created lazily by b200ddp.optim.Adam with the same layout
        self.m_flat: Optional[torch.Tensor] = None
        self.v_flat: Optional[torch.Tensor] = None
        self.opt_step = 0
        self._claimed = set()           # parameters whose arena gradient slice was handed out in this backward pass
        self._claim_reset_queued = False
        self._grad_views: List[torch.Tensor] = []
        with torch.no_grad():
            for i, p in enumerate(self.params):
                o, n = self.offsets[i], p.numel()
                dst = self.p_flat[o:o + n].view(p.shape)
                dst.copy_(p.data)
                p.data = dst
                self._grad_views.append(self.g_flat[o:o + n].view(p.shape))
                p._b200_slot = (self, i)

    # ---- lookups -----------------------------------------------------------------------------------
    def index_of(self, p) -> int:
        slot = getattr(p, "_b200_slot", None)
        if slot is None or slot[0] is not self:
            raise KeyError("parameter does not belong to this arena")
        return slot[1]

    def grad_view(self, i: int) -> torch.Tensor:
        """A NEW view tensor over parameter i's gradient slice (fresh object => autograd may adopt it)."""
        o, n = self.offsets[i], self.params[i].numel()
        return self.g_flat[o:o + n].view(self.params[i].shape)

    def claim(self, i: int) -> bool:
        """First request for parameter i's gradient slice in the running backward pass? (called from backward
        Functions only; the set is cleared by an end-of-backward engine callback)."""
        if i in self._claimed:
            return False
        self._claimed.add(i)
        if not self._claim_reset_queued:
            self._claim_reset_queued = True
            try:
                torch.autograd.Variable._execution_engine.queue_callback(self._release_claims)
            except RuntimeError:        # not inside backward (direct call of a Function's backward in a test)
                self._claim_reset_queued = False
                self._claimed.clear()
        return True

    def _release_claims(self) -> None:
        self._claimed.clear()
        self._claim_reset_queued = False

    def grad_ptr(self, i: int) -> int:
        return self.g_flat.data_ptr() + 4 * self.offsets[i]

    def grad_in_arena(self, i: int) -> bool:
        g = self.params[i].grad
        return g is not None and g.data_ptr() == self.grad_ptr(i) and g.is_contiguous()

    def span(self, first: int, last: int):
        """Element range [start, end) covering parameters first..last inclusive (with their padding)."""
        start = self.offsets[first]
        end = self.offsets[last + 1] if last + 1 < len(self.params) else self.numel
        return start, end

    def ensure_moments(self) -> None:
        if self.m_flat is None:
            self.m_flat = torch.zeros_like(self.p_flat)
            self.v_flat = torch.zeros_like(self.p_flat)

    def moment_views(self, i: int):
        self.ensure_moments()
        o, n = self.offsets[i], self.params[i].numel()
        shape = self.params[i].shape
        return self.m_flat[o:o + n].view(shape), self.v_flat[o:o + n].view(shape)


def arena_of(params) -> Optional[FlatArena]:
    """The arena shared by all `params` if they cover it exactly, else None."""
    params = list(params)
    if not params:
        return None
    slot = getattr(params[0], "_b200_slot", None)
    if slot is None:
        return None
    arena = slot[0]
    if len(params) != len(arena.params):
        return None
    ids = {id(p) for p in arena.params}
    for p in params:
        s = getattr(p, "_b200_slot", None)
        if s is None or s[0] is not arena or id(p) not in ids:
            return None
    return arena
