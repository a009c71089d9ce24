"""f2 — input pipeline: GPU-side CIFAR transform + double-buffered host-to-device prefetch.

Reference: `load_datasets()` (data_and_toy_model.py:8-38) builds torchvision datasets whose transform —
Resize(224), RandomHorizontalFlip, ToTensor, Normalize — runs in PIL on two DataLoader workers
(multi-GPU-training-torch.py:88-97), and `train()` then copies every 77 MB fp32 batch to the GPU with a blocking
`.to(device)` (:110). In the real script that pipeline, not the model, is the bottleneck.

Here the batch crosses PCIe in the dataset's own format (uint8 [B,32,32,3], 393 KB for 128 images, 196x fewer
bytes) on a copy stream while the previous step computes, and `GpuTransform` (one CUDA kernel,
b200ddp_image_transform) produces the normalised fp32 NCHW batch, bit-identical to the torchvision result.

    loader = b200ddp.data.DeviceLoader(images_u8, labels, batch_size=128, sampler=train_sampler, device=device,
                                       transform=b200ddp.data.GpuTransform(224, train=True))
    for inputs, labels in loader:          # already on `device`; the loop's own `.to(device)` is then a no-op
        ...
"""
from __future__ import annotations

import ctypes
from typing import Iterable, Optional

import torch

from . import _lib

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)      # data_and_toy_model.py:18,26
CIFAR_STD = (0.2023, 0.1994, 0.2010)


class GpuTransform:
    """Resize(size) -> [RandomHorizontalFlip] -> ToTensor -> Normalize on a uint8 [B,H,W,3] CUDA batch."""

    def __init__(self, size: int = 224, train: bool = True, mean=CIFAR_MEAN, std=CIFAR_STD, p_flip: float = 0.5) -> None:
        self.size = int(size)
        self.train = bool(train)
        self.p_flip = float(p_flip)
        self._mean = (ctypes.c_float * 3)(*[float(m) for m in mean])
        self._std = (ctypes.c_float * 3)(*[float(s) for s in std])
        self._tables = {}

    def _table(self, n_in: int, device) -> torch.Tensor:
        key = (n_in, str(device))
        t = self._tables.get(key)
        if t is None:
            host = (ctypes.c_int * (5 * self.size))()
            _lib.check(_lib.load().b200ddp_resize_coeffs(n_in, self.size, host), "resize_coeffs")
            t = torch.tensor(list(host), dtype=torch.int32).to(device)
            self._tables[key] = t
        return t

    def draw_flips(self, batch: int) -> torch.Tensor:
        """One decision per image from torch's CPU generator, like RandomHorizontalFlip (`torch.rand(1) < p`)."""
        return (torch.rand(batch) < self.p_flip).to(torch.uint8)

    def __call__(self, images_u8: torch.Tensor, flips: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None):
        if images_u8.dtype != torch.uint8 or images_u8.dim() != 4 or images_u8.shape[-1] != 3:
            raise _lib.B200Error("GpuTransform expects a uint8 [B, H, W, 3] batch (the layout CIFAR-10 is stored in)")
        dev = _lib.dev_index(images_u8)
        x = images_u8 if images_u8.is_contiguous() else images_u8.contiguous()
        B, H, W, _ = x.shape
        if flips is None and self.train:
            flips = self.draw_flips(B).to(x.device, non_blocking=True)
        if flips is not None:
            flips = flips.to(device=x.device, dtype=torch.uint8).contiguous()
        if out is None:
            out = torch.empty((B, 3, self.size, self.size), dtype=torch.float32, device=x.device)
        th, tv = self._table(W, x.device), self._table(H, x.device)
        _lib.check(_lib.load().b200ddp_image_transform(x.data_ptr(), _lib.ptr(flips), out.data_ptr(), B, H, W, self.size,
                                                       self.size, th.data_ptr(), tv.data_ptr(), self._mean, self._std, dev,
                                                       _lib.cur_stream(x.device)), "image_transform")
        return out


class DeviceLoader:
    """Batches of a host-resident dataset, delivered ON the device, one batch ahead of the consumer.

    `images`: uint8 [N,H,W,3] (with `transform=GpuTransform(...)`) or any per-sample tensor; `labels`: [N].
    Index order comes from `sampler` (e.g. b200ddp.DistributedSampler, call its set_epoch as usual) or is sequential.
    Each batch is gathered into one of two pinned staging buffers, copied on a dedicated copy stream while the
    previous batch is being consumed, and (optionally) transformed on the consumer's stream.
    """

    def __init__(self, images: torch.Tensor, labels: torch.Tensor, batch_size: int, device, sampler: Optional[Iterable[int]] = None,
                 transform: Optional[GpuTransform] = None, drop_last: bool = False) -> None:
        if images.shape[0] != labels.shape[0]:
            raise ValueError("images and labels differ in length")
        self.images, self.labels = images, labels
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.B200Error("DeviceLoader needs a CUDA device")
        self.sampler = sampler
        self.transform = transform
        self.drop_last = drop_last
        self._copy = torch.cuda.Stream(self.device)
        shape = (self.batch_size,) + tuple(images.shape[1:])
        self._host_x = [torch.empty(shape, dtype=images.dtype).pin_memory() for _ in range(2)]
        self._host_y = [torch.empty((self.batch_size,), dtype=torch.int64).pin_memory() for _ in range(2)]
        self._host_f = [torch.empty((self.batch_size,), dtype=torch.uint8).pin_memory() for _ in range(2)]
        self._dev_x = [torch.empty(shape, dtype=images.dtype, device=self.device) for _ in range(2)]
        self._dev_y = [torch.empty((self.batch_size,), dtype=torch.int64, device=self.device) for _ in range(2)]
        self._dev_f = [torch.empty((self.batch_size,), dtype=torch.uint8, device=self.device) for _ in range(2)]
        self._ready = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        self._host_free = [torch.cuda.Event() for _ in range(2)]

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else self.images.shape[0]
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _batches(self):
        idx = list(self.sampler) if self.sampler is not None else list(range(self.images.shape[0]))
        for i in range(0, len(idx), self.batch_size):
            part = idx[i:i + self.batch_size]
            if len(part) < self.batch_size and self.drop_last:
                return
            yield torch.as_tensor(part, dtype=torch.int64)

    def _stage(self, slot: int, sel: torch.Tensor) -> int:
        n = int(sel.numel())
        self._host_free[slot].synchronize()          # the previous copy out of this pinned buffer has finished
        torch.index_select(self.images, 0, sel, out=self._host_x[slot][:n])
        self._host_y[slot][:n] = self.labels.index_select(0, sel)
        has_flip = self.transform is not None and self.transform.train
        if has_flip:
            self._host_f[slot][:n] = self.transform.draw_flips(n)
        with torch.cuda.stream(self._copy):
            self._copy.wait_event(self._consumed[slot])
            self._dev_x[slot][:n].copy_(self._host_x[slot][:n], non_blocking=True)
            self._dev_y[slot][:n].copy_(self._host_y[slot][:n], non_blocking=True)
            if has_flip:
                self._dev_f[slot][:n].copy_(self._host_f[slot][:n], non_blocking=True)
            self._host_free[slot].record(self._copy)
            self._ready[slot].record(self._copy)
        return n

    def __iter__(self):
        main = torch.cuda.current_stream(self.device)
        for s in range(2):
            self._consumed[s].record(main)
            self._host_free[s].record(self._copy)
        it = self._batches()
        pending = []
        first = next(it, None)
        if first is None:
            return
        pending.append((0, self._stage(0, first)))
        k = 1
        while pending:
            slot, n = pending.pop(0)
            nxt = next(it, None)
            if nxt is not None:
                pending.append((k % 2, self._stage(k % 2, nxt)))
                k += 1
            main = torch.cuda.current_stream(self.device)
            main.wait_event(self._ready[slot])
            x, y = self._dev_x[slot][:n], self._dev_y[slot][:n]
            if self.transform is not None:
                x = self.transform(x, self._dev_f[slot][:n] if self.transform.train else None)
                y = y.clone()                        # the labels leave the staging buffer too (the loss reads them later)
                self._consumed[slot].record(main)    # the staging buffers are free once transform and clone have read them
                yield x, y
            else:
                yield x, y
                self._consumed[slot].record(torch.cuda.current_stream(self.device))   # after the consumer's step
