"""CPU restatement of the reference's training loop — TEST INFRASTRUCTURE / CPU BASELINE, NOT PRODUCT.

Only tests/, __graft_entry__.smoke() and bench.py (`cpu_baseline`, `--impl reference`) import this.

What it restates (reference = /root/reference, absent on the GPU box, so nothing here reads it):
  * `load_model`            data_and_toy_model.py:41-45   -> make_model(): torchvision AlexNet, 10-class head,
                            random init instead of the ImageNet download (no network; SURVEY 8(c) patch 4)
  * `load_datasets`         data_and_toy_model.py:8-38    -> synthetic_pool(): randn [N,3,224,224] + labels
                            (patch 3; CIFAR-10 cannot be downloaded)
  * `train`                 multi-GPU-training-torch.py:104-133 -> train_epoch(): same statement order
                            (H2D no-op on CPU, zero_grad, forward, criterion, backward, step, loss.item())
  * `evaluate`              multi-GPU-training-torch.py:136-153 -> evaluate_epoch(): same statement order
  * `basic_DDP_training_loop` :228-266                    -> ddp_worker(): setup (gloo), seeds, sampler,
                            stock DistributedDataParallel, CrossEntropyLoss, Adam(lr)
Everything numeric is stock torch / torchvision on the host cores: that IS the reference's CPU path.
oracle/make_golden.py checks, in the build container, that this restatement and the reference's own
functions produce identical loss traces.
"""
from __future__ import annotations

import os
import time
from typing import Callable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.optim as optim


def make_model(seed: int = 1234, hidden: int = 4096, dropout: float = 0.5, num_classes: int = 10) -> nn.Module:
    """Reference load_model() with random weights (SURVEY 8(c) patch 4)."""
    import torchvision
    torch.manual_seed(seed)
    m = torchvision.models.alexnet(weights=None, dropout=dropout)
    if hidden != 4096:
        m.classifier[1] = nn.Linear(256 * 6 * 6, hidden)
        m.classifier[4] = nn.Linear(hidden, hidden)
    m.classifier[6] = nn.Linear(hidden, num_classes)
    return m


def synthetic_pool(n: int, seed: int = 7, hw: int = 224):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 3, hw, hw, generator=g)
    y = torch.randint(0, 10, (n,), generator=g)
    return x, y


def train_epoch(model, loader, criterion: Callable, optimizer, device, max_steps: Optional[int] = None,
                losses: Optional[List[float]] = None):
    """Statement-for-statement restatement of reference train() (multi-GPU-training-torch.py:104-133)."""
    model.train()
    total_running_loss = torch.zeros(1, device=device)
    n_samples = torch.zeros(1, device=device)
    batch_idx = 0
    for inputs, labels in loader:
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        outputs = model(inputs)
        loss = criterion(outputs, labels)
        loss.backward()
        optimizer.step()
        batch_idx += 1
        batch_size = torch.tensor(inputs.shape[0], device=device)
        n_samples += batch_size
        v = loss.item()
        total_running_loss += v * batch_size
        if losses is not None:
            losses.append(v)
        if max_steps is not None and batch_idx >= max_steps:
            break
    return total_running_loss, n_samples


def evaluate_epoch(model, loader, criterion: Callable, device):
    """Statement-for-statement restatement of reference evaluate() (multi-GPU-training-torch.py:136-153)."""
    model.eval()
    correct = torch.zeros(1, device=device)
    total = torch.zeros(1, device=device)
    total_test_loss = torch.zeros(1, device=device)
    with torch.no_grad():
        for inputs, labels in loader:
            inputs, labels = inputs.to(device), labels.to(device)
            outputs = model(inputs)
            loss = criterion(outputs, labels)
            batch_size = torch.tensor(inputs.shape[0], device=device)
            total_test_loss += loss.item() * batch_size
            _, predicted = torch.max(outputs.data, 1)
            total += batch_size
            correct += (predicted == labels).sum().item()
    return total_test_loss, correct, total


def time_cpu_train(batch: int = 128, steps: int = 3, warmup: int = 1, threads: Optional[int] = None,
                   hidden: int = 4096, seed: int = 1234) -> dict:
    """samples/s of the reference's CPU path (ws = 1, all host threads): the bench's cpu_baseline."""
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = threads or cores
    torch.set_num_threads(threads)
    model = make_model(seed=seed, hidden=hidden)
    criterion = nn.CrossEntropyLoss()
    optimizer = optim.Adam(model.parameters(), lr=0.001)
    x, y = synthetic_pool(batch, seed=7)
    loader = [(x, y)]
    dev = torch.device("cpu")
    for _ in range(warmup):
        train_epoch(model, loader, criterion, optimizer, dev)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        train_epoch(model, loader, criterion, optimizer, dev)
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return {"samples_per_s": batch / med, "ms_per_step": med * 1e3, "cores": threads, "steps": steps,
            "warmup": warmup, "batch": batch, "all_ms": [t * 1e3 for t in times]}


def ddp_worker(rank: int, world: int, port: int, out_dir: str, n_samples: int, batch: int, lr: float,
               steps: int, model_seed: int, data_seed: int, dropout: float, threads: int = 2) -> None:
    """Restated glue of basic_DDP_training_loop (:228-266) on CPU/gloo with stock torch DDP."""
    from torch.nn.parallel import DistributedDataParallel as DDP
    from torch.utils.data import DataLoader, DistributedSampler, TensorDataset
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y = synthetic_pool(n_samples, seed=data_seed)
    ds = TensorDataset(x, y)
    sampler = DistributedSampler(ds, shuffle=True, num_replicas=world, rank=rank)
    sampler.set_epoch(0)
    loader = DataLoader(ds, batch_size=batch, sampler=sampler)
    model = make_model(seed=model_seed, dropout=dropout)
    ddp = DDP(model)
    criterion = nn.CrossEntropyLoss()
    optimizer = optim.Adam(ddp.parameters(), lr=lr)
    losses: List[float] = []
    train_epoch(ddp, loader, criterion, optimizer, torch.device("cpu"), max_steps=steps, losses=losses)
    torch.save({"losses": losses,
                "param_sums": {k: float(v.double().sum()) for k, v in ddp.state_dict().items()},
                "param_l2": {k: float(v.double().norm()) for k, v in ddp.state_dict().items()}},
               os.path.join(out_dir, f"trace_rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()
