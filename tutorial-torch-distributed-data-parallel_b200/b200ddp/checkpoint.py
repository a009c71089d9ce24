"""f3 — checkpoint write / resume.

Reference: every `checkpoint_epoch` epochs rank 0 runs `torch.save(model.state_dict(), ckpt_path)` inside the
training loop while every other rank waits at a barrier (multi-GPU-training-torch.py:217-223); the Accelerate
variant calls `accelerator.save_model(model, save_dir)` (multi-GPU-training-accelerate.py:104-108). The file holds
the 16 `module.*` tensors only, so a run cannot be resumed (README.md:51-52 points at these files).

`AsyncCheckpointer.save()` keeps the file format — `torch.load(path, map_location=...)` + `load_state_dict` works
exactly as before — but takes the write off the step's critical path: the parameter arena (one contiguous fp32
buffer) is copied device -> pinned host on a side stream, and a background thread serialises it once the copy has
landed; the training stream only records an event. With `optimizer=` / `sampler=` it also writes a side-car
`<path>.train` (Adam moments and step count, epoch, sampler seed/epoch, torch RNG state) from which
`load_checkpoint()` resumes the run — the reference has no such thing.
"""
from __future__ import annotations

import os
import threading
from collections import OrderedDict
from typing import Optional

import torch

from .arena import FlatArena, arena_of


def _arena_of_model(model) -> Optional[FlatArena]:
    a = getattr(model, "arena", None)
    if isinstance(a, FlatArena):
        return a
    return arena_of([p for p in model.parameters() if p.requires_grad])


class AsyncCheckpointer:
    def __init__(self, model, optimizer=None, sampler=None) -> None:
        self.model = model
        self.optimizer = optimizer
        self.sampler = sampler
        self.arena = _arena_of_model(model)
        self._thread: Optional[threading.Thread] = None
        self._error: Optional[BaseException] = None
        self._stream = None
        self._host = {}
        self._dev = {}
        if self.arena is not None and self.arena.device.type == "cuda":
            self._stream = torch.cuda.Stream(self.arena.device)

    # ---- snapshot -----------------------------------------------------------------------------------------------
    def _pinned(self, tag: str, like: torch.Tensor) -> torch.Tensor:
        t = self._host.get(tag)
        if t is None or t.numel() != like.numel():
            t = torch.empty(like.numel(), dtype=like.dtype).pin_memory()
            self._host[tag] = t
        return t

    def _snapshot(self, tensors: dict):
        """Two hops on the side stream: device -> device snapshot (fast; the only thing the training stream waits
        for, so that the next optimizer step cannot overwrite what is being saved), then snapshot -> pinned host at
        PCIe speed with nobody waiting. Returns (host views, done event). 180 GB of HBM make the extra copy cheap."""
        dev = self.arena.device
        main = torch.cuda.current_stream(dev)
        ready = torch.cuda.Event()
        ready.record(main)                       # everything enqueued so far (the optimizer step) precedes the snapshot
        out = {}
        with torch.cuda.stream(self._stream):
            self._stream.wait_event(ready)
            snaps = {}
            for tag, t in tensors.items():
                d = self._dev.get(tag)
                if d is None or d.numel() != t.numel():
                    d = self._dev[tag] = torch.empty_like(t)
                d.copy_(t, non_blocking=True)
                snaps[tag] = d
            snapped = torch.cuda.Event()
            snapped.record(self._stream)
            for tag, d in snaps.items():
                h = self._pinned(tag, d)
                h.copy_(d, non_blocking=True)
                out[tag] = h
            done = torch.cuda.Event()
            done.record(self._stream)
        main.wait_event(snapped)
        return out, done

    def _named_views(self, flat_host: torch.Tensor, prefix_names) -> "OrderedDict[str, torch.Tensor]":
        a = self.arena
        index = {id(p): i for i, p in enumerate(a.params)}
        sd = OrderedDict()
        for name, p in prefix_names:
            i = index.get(id(p))
            if i is None:
                sd[name] = p.detach().cpu()
            else:
                o = a.offsets[i]
                sd[name] = flat_host[o:o + p.numel()].view(p.shape)
        return sd

    def save(self, path: str, epoch: Optional[int] = None, extra: Optional[dict] = None, blocking: bool = False) -> None:
        """Write `model.state_dict()` to `path` (same keys / dtypes as the reference's torch.save) and, when an
        optimizer or sampler was given, the training state to `path + '.train'`. Returns as soon as the device-side
        snapshot has been enqueued; `wait()` joins the writer."""
        self.wait()
        model = self.model
        keys = list(model.state_dict().keys())
        named = dict(model.named_parameters())
        buffers = {k: v.detach().cpu().clone() for k, v in model.state_dict().items() if k not in named}
        train_state = None
        if self.arena is not None and self._stream is not None:
            tensors = {"p": self.arena.p_flat}
            if self.optimizer is not None and self.arena.m_flat is not None:
                tensors["m"], tensors["v"] = self.arena.m_flat, self.arena.v_flat
            host, done = self._snapshot(tensors)
        else:
            host, done = None, None
        if self.optimizer is not None or self.sampler is not None or epoch is not None or extra:
            train_state = {"epoch": epoch, "extra": extra or {}, "torch_rng": torch.get_rng_state()}
            if self.sampler is not None:
                train_state["sampler"] = {"epoch": getattr(self.sampler, "epoch", 0), "seed": getattr(self.sampler, "seed", 0)}
            if self.optimizer is not None:
                osd = self.optimizer.state_dict()
                steps = {k: (float(v["step"]) if "step" in v else 0.0) for k, v in osd["state"].items()}
                train_state["optimizer"] = {"param_groups": osd["param_groups"], "steps": steps,
                                            "state": None if host is not None and "m" in host else osd["state"]}

        def write():
            try:
                if done is not None:
                    done.synchronize()
                if host is not None:
                    views = self._named_views(host["p"], [(k, named[k]) for k in keys if k in named])
                    sd = OrderedDict((k, views[k] if k in views else buffers[k]) for k in keys)
                else:
                    sd = OrderedDict((k, v.detach().cpu()) for k, v in model.state_dict().items())
                tmp = path + ".tmp"
                torch.save(sd, tmp)
                os.replace(tmp, path)
                if train_state is not None:
                    if host is not None and "m" in host:
                        train_state["optimizer"]["m_flat"] = host["m"].clone()
                        train_state["optimizer"]["v_flat"] = host["v"].clone()
                    tmp = path + ".train.tmp"
                    torch.save(train_state, tmp)
                    os.replace(tmp, path + ".train")
            except BaseException as e:        # surfaced by wait()
                self._error = e

        if blocking or done is None:
            write()
            self._raise()
        else:
            self._thread = threading.Thread(target=write, name="b200ddp-checkpoint", daemon=True)
            self._thread.start()

    def wait(self) -> None:
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        self._raise()

    def _raise(self) -> None:
        if self._error is not None:
            e, self._error = self._error, None
            raise e


def load_checkpoint(path: str, model, optimizer=None, sampler=None, map_location="cpu") -> dict:
    """Load `path` into `model` (reference README.md:51-52 usage: torch.load(map_location) + load_state_dict) and, when
    `path + '.train'` exists, restore Adam's moments / step, the sampler's seed and epoch and torch's CPU RNG state.
    Returns {"epoch": ..., "extra": ...} ({} when there is no side-car)."""
    sd = torch.load(path, map_location=map_location)
    model.load_state_dict(sd)
    side = path + ".train"
    if not os.path.exists(side):
        return {}
    st = torch.load(side, map_location="cpu", weights_only=False)
    if sampler is not None and "sampler" in st:
        sampler.seed = st["sampler"]["seed"]
        sampler.set_epoch(st["sampler"]["epoch"])
    if "torch_rng" in st:
        torch.set_rng_state(st["torch_rng"])
    if optimizer is not None and "optimizer" in st:
        o = st["optimizer"]
        arena = _arena_of_model(model)
        if o.get("state") is not None:
            optimizer.load_state_dict({"state": o["state"], "param_groups": o["param_groups"]})
        elif arena is not None and "m_flat" in o:
            arena.ensure_moments()
            with torch.no_grad():
                arena.m_flat.copy_(o["m_flat"].to(arena.device))
                arena.v_flat.copy_(o["v_flat"].to(arena.device))
            for g, saved in zip(optimizer.param_groups, o["param_groups"]):
                for k, v in saved.items():
                    if k != "params":
                        g[k] = v
            idx = 0
            for g in optimizer.param_groups:
                for p in g["params"]:
                    stp = optimizer._init_state(p)
                    stp["step"] = torch.tensor(float(o["steps"].get(idx, 0.0)), dtype=torch.float32)
                    idx += 1
            first = None
            for p in arena.params:       # the flat kernel keeps ONE shared step tensor per arena
                s = optimizer.state[p]
                if first is None:
                    first = s["step"]
                else:
                    s["step"] = first
    return {"epoch": st.get("epoch"), "extra": st.get("extra", {})}
