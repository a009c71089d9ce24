"""f1 — the evaluate() path.

Drop-in for the `evaluate(model, test_loader, criterion, device)` function of the reference scripts
(multi-GPU-training-torch.py:136-153; multi-GPU-training-accelerate.py:60-75) and for the metric aggregation of
`run_training_loop` (:194-204):

    total_test_loss, n_correct, n_samples = b200ddp.evaluate(model, test_loader, criterion, device)
    train_loss, test_loss, n_correct, n_samples = b200ddp.all_reduce_metrics(...)      # ONE collective

What changes underneath: the forward pass runs in inference mode (no ReLU/dropout gate bytes, no pooling argmax codes
are written or kept), the 10-class head + cross-entropy + argmax + comparison with the labels are ONE kernel
(b200ddp_head_eval) that adds into a 3-float device buffer, and nothing synchronises with the host per batch — the
reference's loop calls `.item()` twice per batch (:148,:151). Return values keep the reference's shapes:
three 1-element float tensors on `device` (sum of per-sample losses, number of correct predictions, number of samples).
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.distributed as dist

from . import _lib
from .comm import get_comm


def _unwrap(model):
    return model.module if hasattr(model, "module") and isinstance(getattr(model, "module"), torch.nn.Module) else model


class EvalMetrics:
    """The device-side accumulator: metrics = [sum of losses, correct predictions, samples]."""

    def __init__(self, device) -> None:
        self.device = torch.device(device)
        self.metrics = torch.zeros(3, dtype=torch.float32, device=self.device)
        self._scratch = None

    def scratch(self, B: int) -> torch.Tensor:
        need = int(_lib.load().b200ddp_eval_scratch_bytes(B))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.zeros(need, dtype=torch.uint8, device=self.device)
        return self._scratch

    def add_batch(self, model, inputs: torch.Tensor, labels: torch.Tensor) -> None:
        """Forward `inputs` through `model` under no_grad and add the batch's loss sum / hits / size."""
        L = _lib.load()
        net = _unwrap(model)
        labels = labels.contiguous()
        if labels.dtype != torch.int64:
            raise TypeError("evaluate expects int64 class indices")
        B = int(labels.shape[0])
        sc = self.scratch(B)
        dev, st = _lib.dev_index(self.metrics), _lib.cur_stream(self.device)
        head = getattr(net, "classifier", None)
        head = head[-1] if isinstance(head, torch.nn.Sequential) and len(head) else None
        with torch.no_grad():
            if (hasattr(net, "forward_hidden") and isinstance(head, torch.nn.Linear) and head.out_features <= 16
                    and head.weight.is_cuda and head.weight.dtype == torch.float32):
                h = _lib.f32c(net.forward_hidden(inputs), "hidden activations")
                w = _lib.f32c(head.weight, "head weight")
                _lib.check(L.b200ddp_head_eval(h.data_ptr(), w.data_ptr(), _lib.ptr(head.bias), labels.data_ptr(), None, B,
                                               w.shape[0], w.shape[1], self.metrics.data_ptr(), sc.data_ptr(), sc.numel(),
                                               dev, st), "head_eval")
            else:
                z = _lib.f32c(model(inputs), "logits")
                _lib.check(L.b200ddp_logits_eval(z.data_ptr(), labels.data_ptr(), B, z.shape[1], self.metrics.data_ptr(),
                                                 sc.data_ptr(), sc.numel(), dev, st), "logits_eval")

    def result(self):
        m = self.metrics
        return m[0:1].clone(), m[1:2].clone(), m[2:3].clone()


def evaluate(model, test_loader, criterion=None, device=None):
    """Same signature and return value as the reference's evaluate() (torch.py:136-153): (total_test_loss, correct,
    total) as 1-element tensors on `device`. `criterion` is accepted for signature compatibility; the loss is the
    mean-reduced cross-entropy the reference constructs (nn.CrossEntropyLoss(), :248), summed over samples."""
    if device is None:
        device = next(_unwrap(model).parameters()).device
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.B200Error("b200ddp.evaluate needs a CUDA device (no CPU fallback)")
    model.eval()
    acc = EvalMetrics(device)
    for inputs, labels in test_loader:
        inputs = inputs.to(device, non_blocking=True)
        labels = labels.to(device, non_blocking=True)
        acc.add_batch(model, inputs, labels)
    return acc.result()


def all_reduce_metrics(*tensors: torch.Tensor, group=None) -> Sequence[torch.Tensor]:
    """Sum every given metric tensor over the ranks with ONE collective (the reference issues one all_reduce per
    scalar, torch.py:198-204): the tensors are packed into one small buffer, reduced, and written back in place."""
    if not tensors:
        return tensors
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return tensors
    flat = torch.cat([t.detach().reshape(-1).to(torch.float32) for t in tensors])
    if flat.is_cuda:
        comm = get_comm(flat.device, group)
        comm.allreduce_(flat, 1.0)
    else:
        dist.all_reduce(flat, group=group)
    off = 0
    with torch.no_grad():
        for t in tensors:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t).to(t.dtype))
            off += n
    return tensors
