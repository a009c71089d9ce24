"""`Accelerator` shim — the seven members reference multi-GPU-training-accelerate.py uses
(:19,115,122,129-131,53,96,106,108): `Accelerator()`, `.device`, `.prepare(model, optimizer, loader)`,
`.backward(loss)`, `.is_local_main_process`, `.wait_for_everyone()`, `.save_model(model, save_dir)`.

HF Accelerate is not installed in this image and its version is unpinned by the reference
(requirements.txt:3), so the behaviour below follows Accelerate's documented defaults and its parity
against the real package is UNPINNED:
  * process group from torchrun-style env (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*), single process
    when absent; one process per GPU, device = cuda:LOCAL_RANK;
  * `prepare` wraps the model in b200ddp's DistributedDataParallel, returns the optimizer unchanged,
    and re-creates the DataLoader over a sharded batch sampler: rank r takes batches r, r+ws, r+2ws ...
    of the original batch sampler (`split_batches=False`), short tails are completed by cycling from
    the start (`even_batches=True`), and each batch is moved to `.device` (the training loop of the
    reference does not call `.to(device)`, :43);
  * `save_model` writes the unwrapped module's weights on the main process as `model.safetensors`
    (torch.save `pytorch_model.bin` when safetensors is absent), unprefixed keys.
"""
from __future__ import annotations

import os
from typing import Any, Iterator, List

import torch
import torch.distributed as dist
from torch import nn
from torch.utils.data import BatchSampler, DataLoader

from .distributed import DistributedDataParallel


class BatchSamplerShard(BatchSampler):
    """Every `num_processes`-th batch of `batch_sampler`, starting at `process_index`."""

    def __init__(self, batch_sampler, num_processes: int = 1, process_index: int = 0, even_batches: bool = True) -> None:
        self.batch_sampler = batch_sampler
        self.num_processes = num_processes
        self.process_index = process_index
        self.even_batches = even_batches
        self.batch_size = getattr(batch_sampler, "batch_size", None)
        self.drop_last = getattr(batch_sampler, "drop_last", False)

    def __len__(self) -> int:
        n = len(self.batch_sampler)
        if n % self.num_processes == 0:
            return n // self.num_processes
        if self.even_batches or self.process_index < n % self.num_processes:
            return n // self.num_processes + 1
        return n // self.num_processes

    def __iter__(self) -> Iterator[List[int]]:
        ws, r, full = self.num_processes, self.process_index, self.batch_size
        initial: List[int] = []      # indices of the first `ws` batches: the pool the tail is completed from
        mine: List[int] = []
        batch: List[int] = []
        idx = -1
        for idx, batch in enumerate(self.batch_sampler):
            batch = list(batch)
            if not self.drop_last and idx < ws:
                initial += batch
            if idx % ws == r:
                mine = batch
            # a round is released only once its last batch turned out to be full
            if idx % ws == ws - 1 and (full is None or len(batch) == full):
                yield mine
                mine = []
        if self.drop_last or not initial:
            return
        if not self.even_batches:
            if mine:
                yield mine
            return
        if len(mine) == full:
            yield mine
        while len(initial) < ws * full:
            initial += initial
        if len(batch) == full:
            batch = []
            idx += 1
        cycle = 0
        while idx % ws != 0 or len(batch) > 0:
            end = cycle + full - len(batch)
            batch = batch + initial[cycle:end]
            if idx % ws == r:
                yield batch
            cycle = end
            batch = []
            idx += 1


class _DeviceLoader:
    """Iterates a DataLoader and moves every tensor of a batch to `device`."""

    def __init__(self, loader: DataLoader, device: torch.device) -> None:
        self.loader = loader
        self.device = device
        self.dataset = loader.dataset
        self.batch_sampler = loader.batch_sampler

    def __len__(self) -> int:
        return len(self.loader)

    def _move(self, obj: Any) -> Any:
        if torch.is_tensor(obj):
            return obj.to(self.device, non_blocking=True)
        if isinstance(obj, (list, tuple)):
            return type(obj)(self._move(o) for o in obj)
        if isinstance(obj, dict):
            return {k: self._move(v) for k, v in obj.items()}
        return obj

    def __iter__(self):
        for batch in self.loader:
            yield self._move(batch)


class Accelerator:
    def __init__(self, cpu: bool = False, **ddp_kwargs) -> None:
        self.num_processes = int(os.environ.get("WORLD_SIZE", "1"))
        self.process_index = int(os.environ.get("RANK", "0"))
        self.local_process_index = int(os.environ.get("LOCAL_RANK", str(self.process_index)))
        use_cuda = torch.cuda.is_available() and not cpu
        if self.num_processes > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            dist.init_process_group("nccl" if use_cuda else "gloo", rank=self.process_index,
                                    world_size=self.num_processes)
        if dist.is_available() and dist.is_initialized():
            self.num_processes = dist.get_world_size()
            self.process_index = dist.get_rank()
        if use_cuda:
            self.device = torch.device("cuda", self.local_process_index % max(torch.cuda.device_count(), 1))
            torch.cuda.set_device(self.device)
        else:
            self.device = torch.device("cpu")
        self._ddp_kwargs = ddp_kwargs

    # ---- properties the script reads ------------------------------------------------------------------
    @property
    def is_local_main_process(self) -> bool:
        return self.local_process_index == 0

    @property
    def is_main_process(self) -> bool:
        return self.process_index == 0

    # ---- prepare -----------------------------------------------------------------------------------------
    def prepare(self, *objs):
        out = tuple(self._prepare_one(o) for o in objs)
        return out if len(out) > 1 else out[0]

    def _prepare_one(self, obj):
        if isinstance(obj, DistributedDataParallel):
            return obj
        if isinstance(obj, nn.Module):
            obj = obj.to(self.device)
            if self.num_processes > 1 or self.device.type == "cuda":
                kw = dict(self._ddp_kwargs)
                if self.device.type == "cuda":
                    kw.setdefault("device_ids", [self.device.index])
                return DistributedDataParallel(obj, **kw)
            return obj
        if isinstance(obj, DataLoader):
            return self._prepare_loader(obj)
        return obj   # optimizers and anything else pass through unchanged

    def _prepare_loader(self, loader: DataLoader):
        if self.num_processes > 1:
            shard = BatchSamplerShard(loader.batch_sampler, self.num_processes, self.process_index)
            loader = DataLoader(loader.dataset, batch_sampler=shard, num_workers=loader.num_workers,
                                collate_fn=loader.collate_fn, pin_memory=loader.pin_memory,
                                timeout=loader.timeout, worker_init_fn=loader.worker_init_fn)
        return _DeviceLoader(loader, self.device)

    # ---- step helpers --------------------------------------------------------------------------------------
    def backward(self, loss: torch.Tensor, **kwargs) -> None:
        loss.backward(**kwargs)

    def wait_for_everyone(self) -> None:
        if self.num_processes > 1:
            dist.barrier()

    def unwrap_model(self, model: nn.Module) -> nn.Module:
        return model.module if isinstance(model, DistributedDataParallel) else model

    def save_model(self, model: nn.Module, save_directory: str, safe_serialization: bool = True) -> None:
        if not self.is_main_process:
            return
        os.makedirs(save_directory, exist_ok=True)
        state = {k: v.detach().to("cpu").contiguous().clone() for k, v in self.unwrap_model(model).state_dict().items()}
        if safe_serialization:
            try:
                from safetensors.torch import save_file
                save_file(state, os.path.join(save_directory, "model.safetensors"), metadata={"format": "pt"})
                return
            except ImportError:
                pass
        torch.save(state, os.path.join(save_directory, "pytorch_model.bin"))

    def print(self, *args, **kwargs) -> None:
        if self.is_local_main_process:
            print(*args, **kwargs)
