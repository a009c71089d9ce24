#!/usr/bin/env python
"""bench.py — train samples/sec of the DDP training step of the toy model (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1] at N=1, configs[2] for N>1): torchvision-AlexNet-10 (hidden 4096,
57,044,810 fp32 parameters), per-rank batch 128 of synthetic [3,224,224] fp32 images, CrossEntropy,
Adam(lr 1e-3), dropout 0.5 on — one "step" = zero_grad, forward, loss, backward (bucketed gradient
all-reduce when N>1), optimizer step: reference multi-GPU-training-torch.py:118-126.

Own arm: one process per GPU (torchrun env for N>1), timing with CUDA events on the compute stream
bracketed by barrier + synchronize, max over ranks. `value` has the inputs resident in HBM (batches are
gathered from a 1024-sample device pool through DistributedSampler indices); `e2e` copies every step's
batch from pinned host memory and reads every step's loss back. The working set of a step (77 MB batch
+ 228 MB weights + 228 MB gradients + 456 MB Adam moments) is several times the 126 MB L2.

Reference arm (`--impl reference`): the reference's CPU path (oracle/reference_loop.py: stock torch +
torchvision on the host cores, the restatement checked against the reference's own train()).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

BATCH = 128
POOL = 1024
VIRTUAL_N = 50_000
P_PARAMS = 57_044_810


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="tf32", choices=["tf32", "fp32"])
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--bn", action="store_true", help="AlexNet-BN + convert_sync_batchnorm as the main job (configs[3])")
    ap.add_argument("--no-extras", action="store_true",
                    help="timed regions only: no stock-torch arm, no extra configs, no parity block, no CPU baseline")
    return ap.parse_args()


# ---- clocks ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed regions: NVML polled every 25 ms from a thread
    (a 200-step region lasts ~0.5 s), nvidia-smi -lms 20 as the fallback when pynvml cannot open the device."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index = index
        self.proc = None
        self.lines = []
        self.samples = []          # (sm_mhz, reasons_mask)
        self.max_mhz = None
        self.stop_flag = threading.Event()
        self.thread = None
        self.source = None
        # polling period: NVML queries take driver locks that kernel launches also need; 5 ms polling produced
        # occasional 10-20 ms launch stalls in a 50 ms timed region, 25 ms does not
        self.interval = float(os.environ.get("B200DDP_BENCH_CLOCK_MS", "25")) / 1e3

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def _poll(self, pynvml, h) -> None:
        while not self.stop_flag.is_set():
            try:
                mhz = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                try:
                    mask = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    mask = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((float(mhz), int(mask)))
            except Exception:
                pass
            time.sleep(self.interval)

    def start(self) -> None:
        if self.interval <= 0:
            self.source = None
            return
        try:
            pynvml, h = self._nvml_handle()
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, args=(pynvml, h), daemon=True)
            self.thread.start()
            self.source = f"nvml {self.interval * 1e3:.0f} ms"
            return
        except Exception:
            self.thread = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            self.source = "nvidia-smi -lms 20"
        except Exception:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None and not self.samples and self.source is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        sm, mx, reasons = [], [], set()
        if self.proc is None:
            self.stop_flag.set()
            self.thread.join(timeout=1)
            for mhz, mask in self.samples:
                sm.append(mhz)
                mx.append(self.max_mhz)
                for n, bit in self.REASONS:
                    if mask & bit:
                        reasons.add(n)
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            names = [n for n, _ in self.REASONS]
            order = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for ln in self.lines:
                f = [c.strip() for c in ln.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0]))
                    mx.append(float(f[1]))
                except ValueError:
                    continue
                for n, v in zip(order, f[3:7]):
                    if v.lower().startswith("active") and n in names:
                        reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples"], "source": self.source}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": max(mx), "samples": len(sm),
                "reasons": sorted(reasons), "source": self.source}


# ---- reference arm ----------------------------------------------------------------------------------------------
def _ref_ddp_worker(rank, world, port, batch, steps, warmup, hidden, threads, out_path):
    """The reference's own loop (oracle/reference_loop: stock DDP over gloo, the restated train()) on host cores."""
    import torch
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    from oracle import reference_loop as rl
    # this process was spawned from a torchrun worker: drop the launcher's rendezvous environment, in particular
    # TORCHELASTIC_USE_AGENT_STORE, which would make BOTH ranks TCP-store clients of a server nobody runs
    for k in list(os.environ):
        if k.startswith("TORCHELASTIC") or k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "GROUP_RANK",
                                                 "GROUP_WORLD_SIZE", "ROLE_RANK", "ROLE_WORLD_SIZE", "ROLE_NAME",
                                                 "MASTER_ADDR", "MASTER_PORT"):
            os.environ.pop(k, None)
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    model = DDP(rl.make_model(seed=1234, hidden=hidden))
    crit = torch.nn.CrossEntropyLoss()
    opt = torch.optim.Adam(model.parameters(), lr=0.001)
    x, y = rl.synthetic_pool(batch, seed=7 + rank)
    dev = torch.device("cpu")
    for _ in range(warmup):
        rl.train_epoch(model, [(x, y)], crit, opt, dev)
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        rl.train_epoch(model, [(x, y)], crit, opt, dev)
    dist.barrier()
    total = time.perf_counter() - t0
    if rank == 0:
        json.dump({"total_s": total}, open(out_path, "w"))
    dist.destroy_process_group()


def _free_port() -> int:
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_reference(args) -> None:
    """The reference's CPU path on the box's host cores. N = 1: one process, all threads. N > 1 (rank 0 only; the GPUs
    are not used by this arm): the reference's DDP loop over gloo with world size 2 — its own multi-process
    configuration (BASELINE.json configs[0]) — cores // 2 threads per rank; `config` states what really ran."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import reference_loop as rl
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ws = 1 if args.gpus <= 1 else 2
    threads = max(1, cores // ws)
    torch.set_num_threads(threads)
    batch = args.batch
    # size the sample from one cold step so that the whole run stays within a few minutes
    model = rl.make_model(seed=1234, hidden=args.hidden)
    crit = torch.nn.CrossEntropyLoss()
    opt = torch.optim.Adam(model.parameters(), lr=0.001)
    x, y = rl.synthetic_pool(min(batch, 32), seed=7)
    t0 = time.perf_counter()
    rl.train_epoch(model, [(x, y)], crit, opt, torch.device("cpu"))
    per_sample = (time.perf_counter() - t0) / x.shape[0]
    budget = 200.0
    est = per_sample * batch * (args.steps + args.warmup) * (2.0 if ws == 2 else 1.0)
    if est > budget and batch > 16:
        batch = max(16, int(batch * budget / est) // 16 * 16)
    if ws == 1:
        x, y = rl.synthetic_pool(batch, seed=7)
        for _ in range(args.warmup):
            rl.train_epoch(model, [(x, y)], crit, opt, torch.device("cpu"))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            rl.train_epoch(model, [(x, y)], crit, opt, torch.device("cpu"))
        total = time.perf_counter() - t0
    else:
        del model, opt
        import tempfile
        import torch.multiprocessing as mp
        out = os.path.join(tempfile.mkdtemp(), "ref.json")
        mp.spawn(_ref_ddp_worker, args=(ws, _free_port(), batch, args.steps, args.warmup, args.hidden, threads, out),
                 nprocs=ws, join=True)
        total = json.load(open(out))["total_s"]
    value = ws * batch * args.steps / total
    sample = (f"{args.steps} timed CPU training steps, {ws} process(es) x batch {batch} (the GPU arm's per-rank batch is "
              f"{args.batch}); stock torch {torch.__version__} + torchvision AlexNet-10 + Adam"
              f"{' + DistributedDataParallel over gloo' if ws > 1 else ''}, {threads} threads per process")
    cfg = workload_config(args, batch_override=batch)
    cfg["parallelism"] = f"cpu gloo ws={ws}" if ws > 1 else "cpu ws=1"
    cfg["global_batch"] = ws * batch
    cfg["reference_world"] = ws
    cfg["note"] = ("the reference arm runs on host cores only: at --gpus N > 1 it is the reference's CPU/gloo DDP loop with "
                   "world size 2, not N GPUs")
    line = {"impl": "reference", "metric": "train_samples_per_sec", "value": value, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads * ws, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, batch_override=None) -> dict:
    return {"workload": "AlexNet-10 DDP training step (BASELINE.json configs[1]; configs[2] when n_gpus>1): "
                        "zero_grad + forward + CrossEntropy + backward + bucketed grad all-reduce + Adam",
            "per_rank_batch": batch_override or args.batch, "global_batch": (batch_override or args.batch) * args.gpus,
            "input": "[3,224,224] fp32 synthetic", "hidden": args.hidden, "params": P_PARAMS if args.hidden == 4096 else None,
            "optimizer": "Adam lr=1e-3 (launched per gradient bucket inside backward, optimizer.step() is the join)",
            "dropout": 0.5, "parallelism": f"dp{args.gpus}",
            "allreduce": os.environ.get("B200DDP_ALLREDUCE", "own") if args.gpus > 1 else None,
            "l2": "working set per step (77 MB batch + 912 MB weights/grads/moments) >> 126 MB L2; inputs rotate "
                  "through a 1024-sample pool"}


# ---- per-kernel accounting (CUDA events around every C-ABI call, see b200ddp/_lib.py) ---------------------------------------
def _kernel_work(name: str, ints: tuple):
    """(label, algorithmic bytes, algorithmic flops) of one entry-point call, from its integer arguments."""
    try:
        if name == "b200ddp_adam_step_flat":
            return f"adam_step_flat n={ints[0]}", 28.0 * ints[0], 0.0
        if name == "b200ddp_conv2d_fwd":
            ldw, ko, nb, cin, h, w, cout, kh, kw, s, p = ints[:11]
            oh, ow = (h + 2 * p - kh) // s + 1, (w + 2 * p - kw) // s + 1
            fl = 2.0 * nb * oh * ow * cout * cin * kh * kw
            by = 4.0 * (nb * cin * h * w + nb * cout * oh * ow + cout * cin * kh * kw)
            return f"conv2d_fwd/dgrad Cin={cin} {h}x{w} -> Cout={cout} k{kh} s{s}", by, fl
        if name == "b200ddp_conv2d_wgrad":
            ko, nb, cin, h, w, cout, kh, kw, s, p = ints[:10]
            oh, ow = (h + 2 * p - kh) // s + 1, (w + 2 * p - kw) // s + 1
            fl = 2.0 * nb * oh * ow * cout * cin * kh * kw
            by = 4.0 * (nb * cin * h * w + nb * cout * oh * ow + cout * cin * kh * kw)
            return f"conv2d_wgrad Cin={cin} {h}x{w} Cout={cout} k{kh}", by, fl
        if name in ("b200ddp_linear_fwd", "b200ddp_linear_bwd_dx", "b200ddp_linear_bwd_dw"):
            bb, n, k = ints[:3]
            return f"{name[8:]} B={bb} N={n} K={k}", 4.0 * (n * k + bb * (n + k)), 2.0 * bb * n * k
        # C6: algorithmic bytes of DESIGN.md section 4 (one read + one write of the activation per pass, 1-byte codes)
        if name == "b200ddp_relu_maxpool_fwd":
            bb, c, h, w = ints[:4]
            oh, ow = (h - 3) // 2 + 1, (w - 3) // 2 + 1
            return "relu_maxpool_fwd", bb * c * (4.0 * h * w + 5.0 * oh * ow), 0.0
        if name in ("b200ddp_relu_maxpool_bwd_bias", "b200ddp_relu_maxpool_bwd"):
            bb, c, h, w = ints[:4]
            oh, ow = (h - 3) // 2 + 1, (w - 3) // 2 + 1
            return name[8:], bb * c * (4.0 * h * w + 5.0 * oh * ow), 0.0
        if name == "b200ddp_relu_bwd_bias":
            bb, c, hw = ints[:3]
            return "relu_bwd_bias", 12.0 * bb * c * hw, 0.0
        if name == "b200ddp_gate_bias_bwd":
            bb, n = ints[:2]
            return "gate_bias_bwd", 9.0 * bb * n, 0.0
        if name == "b200ddp_dropout_fwd":
            return "dropout_fwd", 9.0 * ints[0], 0.0
    except Exception:
        pass
    return name[8:], 0.0, 0.0


GROUPS = (("conv", ("conv2d_fwd", "conv2d_wgrad", "conv2d_weight_prep")),
          ("fc", ("linear_fwd", "linear_bwd", "gate_bias_bwd", "dropout")),
          ("pool_relu", ("relu_maxpool", "relu_bwd", "relu_fwd", "channel_sum")),
          ("adam", ("adam_step_flat",)),
          ("loss_head_clip", ("softmax_ce", "scale_by_device_scalar", "pack_scale_clip", "sqnorm", "clip_coef")))


def roofline_groups(kernels, step_ms, hbm_peak, tf32_peak):
    """Where the step stands, group by group: algorithmic work / summed kernel time, against the roofline
    that bounds the group (conv: dense TF32 tensor rate; everything else: HBM). Per-kernel CUDA-event timing of a
    serialised pass (side / optimizer streams off), so the shares add up to ~100 % of a step without overlap."""
    out = {}
    for gname, keys in GROUPS:
        ks = [k for k in kernels if any(t in k["kernel"] for t in keys)]
        ms = sum(k["ms_per_step"] for k in ks)
        if ms <= 0:
            continue
        fl = sum(k["alg_flops_per_call"] * k["calls_per_step"] for k in ks)
        by = sum(k["alg_bytes_per_call"] * k["calls_per_step"] for k in ks)
        g = {"ms_per_step": ms, "share_of_serial_step": None, "launches_per_step": sum(k["calls_per_step"] for k in ks)}
        if gname == "conv":
            ach = fl / ms / 1e9
            g.update({"bound": "tensor", "algorithmic_flops_per_step": fl, "achieved": ach, "unit": "TFLOP/s",
                      "peak": tf32_peak, "frac": ach / tf32_peak})
        else:
            ach = by / ms / 1e6
            g.update({"bound": "hbm", "algorithmic_bytes_per_step": by, "achieved": ach, "unit": "GB/s", "peak": hbm_peak,
                      "frac": ach / hbm_peak})
        out[gname] = g
    tot = sum(k["ms_per_step"] for k in kernels)
    for g in out.values():
        g["share_of_serial_step"] = g["ms_per_step"] / tot if tot > 0 else None
    out["_serial_sum_ms"] = tot
    out["_live_step_ms"] = step_ms
    return out


def summarize_kernels(calls, steps: int):
    agg = {}
    for name, ints, ms in calls:
        label, by, fl = _kernel_work(name, ints)
        a = agg.setdefault(label, {"kernel": label, "calls": 0, "ms": 0.0, "bytes": 0.0, "flops": 0.0})
        a["calls"] += 1
        a["ms"] += ms
        a["bytes"] += by
        a["flops"] += fl
    out = []
    for a in agg.values():
        ms_per_call = a["ms"] / max(a["calls"], 1)
        out.append({"kernel": a["kernel"], "calls_per_step": a["calls"] / steps, "ms_per_step": a["ms"] / steps,
                    "GBps": a["bytes"] / a["ms"] / 1e6 if a["ms"] > 0 and a["bytes"] else None,
                    "TFLOPs": a["flops"] / a["ms"] / 1e9 if a["ms"] > 0 and a["flops"] else None,
                    "alg_bytes_per_call": a["bytes"] / max(a["calls"], 1), "alg_flops_per_call": a["flops"] / max(a["calls"], 1),
                    "ms_per_call": ms_per_call})
    out.sort(key=lambda d: -d["ms_per_step"])
    return out


def ncu_traffic(label: str):
    """DRAM bytes per launch of this kernel from the committed `ncu --set full` capture (profiles/ncu_traffic.json),
    or None when that capture does not hold the kernel."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return table[label]["dram_bytes_per_launch"]
    except Exception:
        return None


def roofline_of_dominant(kernels, hbm_peak, tf32_peak, peak_src, step_ms):
    """Roofline object for the entry point with the largest share of the step (measured with CUDA events)."""
    if not kernels:
        return None
    k = kernels[0]
    ai = (k["alg_flops_per_call"] / k["alg_bytes_per_call"]) if k["alg_bytes_per_call"] else 0.0
    tensor_bound = ai > tf32_peak * 1e3 / hbm_peak        # arithmetic intensity above the machine balance point
    if tensor_bound and k["TFLOPs"]:
        return {"kernel": k["kernel"], "bound": "tensor", "achieved": k["TFLOPs"], "peak": tf32_peak, "unit": "TFLOP/s",
                "frac": k["TFLOPs"] / tf32_peak, "traffic": ncu_traffic(k["kernel"]),
                "algorithmic_bytes_per_launch": k["alg_bytes_per_call"], "algorithmic_flops_per_launch": k["alg_flops_per_call"],
                "avg_launch_ms": k["ms_per_call"], "share_of_step": k["ms_per_step"] / step_ms,
                "peak_source": "dense TF32 = bf16_tflops_sustained / 2 (MEASURED_PEAKS.json); operands are TF32"}
    return {"kernel": k["kernel"], "bound": "hbm", "achieved": k["GBps"], "peak": hbm_peak, "unit": "GB/s",
            "frac": (k["GBps"] or 0.0) / hbm_peak, "traffic": ncu_traffic(k["kernel"]), "algorithmic_bytes_per_launch": k["alg_bytes_per_call"],
            "avg_launch_ms": k["ms_per_call"], "share_of_step": k["ms_per_step"] / step_ms, "peak_source": peak_src}


# ---- own arm ------------------------------------------------------------------------------------------------------
class Job:
    """One training job of the workload: model + DDP wrapper + Adam + loss, a device-resident synthetic pool indexed
    through DistributedSampler, and the step the reference's train() performs (multi-GPU-training-torch.py:118-126)."""

    def __init__(self, args, dev, world, rank, local, hidden=None, bn=False, via_accelerate=False, batch=None):
        import torch
        import b200ddp
        from b200ddp.models import AlexNet
        self.torch, self.dev, self.world, self.rank = torch, dev, world, rank
        self.B = batch or args.batch
        torch.manual_seed(1234)
        model = AlexNet(hidden=hidden or args.hidden, dropout=0.5, mode=args.mode, bn=bn)
        if bn:
            model = b200ddp.SyncBatchNorm.convert_sync_batchnorm(model)      # README.md:79-81: convert, then wrap
        if via_accelerate:
            # reference multi-GPU-training-accelerate.py:113-131: optimizer built on the bare model, then prepare()
            from b200ddp.accelerate import Accelerator
            self.accelerator = Accelerator()
            self.opt = b200ddp.optim.Adam(model.parameters(), lr=0.001)
            self.ddp, self.opt = self.accelerator.prepare(model, self.opt)
        else:
            self.ddp = b200ddp.DistributedDataParallel(model.to(dev), device_ids=[local])
            self.opt = b200ddp.optim.Adam(self.ddp.parameters(), lr=0.001)
        self.model = self.ddp.module
        self.crit = b200ddp.CrossEntropyLoss()
        self.ddp.train()
        gcpu = torch.Generator().manual_seed(7)
        pool_y_cpu = torch.randint(0, 10, (POOL,), generator=gcpu)
        gdev = torch.Generator(device=dev).manual_seed(7)
        self.pool_x = torch.randn(POOL, 3, 224, 224, device=dev, generator=gdev)
        self.pool_y = pool_y_cpu.to(dev)

        class _Virtual:
            def __len__(self):
                return VIRTUAL_N
        self.sampler = b200ddp.DistributedSampler(_Virtual(), num_replicas=world, rank=rank, shuffle=True)
        self.idx_dev = None
        self.cursor = 0

    def prepare_indices(self, total_steps):
        torch = self.torch
        idx, epoch = [], 0
        while len(idx) < total_steps * self.B + self.B:
            self.sampler.set_epoch(epoch)
            idx += list(self.sampler)
            epoch += 1
        self.idx_dev = (torch.tensor(idx, dtype=torch.int64) % POOL).to(self.dev)
        self.cursor = 0

    def batch(self):
        i, B = self.cursor, self.B
        self.cursor += 1
        sel = self.idx_dev[i * B:(i + 1) * B]
        return self.pool_x.index_select(0, sel), self.pool_y.index_select(0, sel)

    def step(self, x, y):
        self.opt.zero_grad()
        loss = self.crit(self.ddp(x), y)
        loss.backward()
        self.opt.step()
        return loss

    def step_resident(self):
        x, y = self.batch()
        return self.step(x, y)

    def free(self):
        self.ddp.reducer.remove_hooks()
        for k in ("pool_x", "pool_y", "ddp", "opt", "model", "idx_dev"):
            setattr(self, k, None)
        import gc
        gc.collect()
        self.torch.cuda.empty_cache()


def timed_steps(job, steps, world, sync_all):
    """ms per step (max over ranks) of `steps` resident-input steps, CUDA events on the compute stream."""
    import torch
    import torch.distributed as dist
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = job.step_resident()
    e1.record()
    sync_all()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=job.dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) / steps, float(loss.detach())


def comm_report(job, world, steps, sync_all, step_ms):
    """R4 roofline: device time and NVLink bus bandwidth of every bucket collective, and the exposed communication
    (step time minus the time of the same step with the collectives skipped)."""
    import torch
    comm = job.ddp.comm
    comm.set_timing(True)
    for _ in range(3):
        job.step_resident()
    torch.cuda.synchronize()
    times = comm.bucket_times()
    comm.set_timing(False)
    buckets = []
    for ms, n in times:
        by = 4.0 * n
        buckets.append({"bytes": by, "ms": ms, "busbw_GBps": (by * 2 * (world - 1) / world / (ms * 1e-3) / 1e9) if ms > 0 else None})
    tot_b = sum(b["bytes"] for b in buckets)
    tot_ms = sum(b["ms"] for b in buckets if b["ms"] and b["ms"] > 0)
    red = job.ddp.reducer
    red.dry_run_collectives = True
    dry_ms, _ = timed_steps(job, max(10, min(steps, 30)), world, sync_all)
    red.dry_run_collectives = False
    # the dry run left the replicas with different parameters: make them identical again before anything else runs
    comm.broadcast_(job.ddp.arena.p_flat, 0)
    if job.ddp.arena.m_flat is not None:
        comm.broadcast_(job.ddp.arena.m_flat, 0)
        comm.broadcast_(job.ddp.arena.v_flat, 0)
    torch.cuda.synchronize()
    busbw = (tot_b * 2 * (world - 1) / world / (tot_ms * 1e-3) / 1e9) if tot_ms > 0 else None
    return {"algorithm": red.algo, "arena_mode": {0: "registered", 1: "symmetric window, peer mappings",
                                                 2: "symmetric window, NVSwitch multicast"}.get(comm.arena_mode),
            "buckets": buckets, "bytes_per_step": tot_b, "collective_ms_per_step": tot_ms,
            "busbw_GBps": busbw, "nvlink_peak_GBps": 900.0, "frac_of_nvlink": (busbw / 900.0) if busbw else None,
            "dry_step_ms": dry_ms, "exposed_comm_ms": step_ms - dry_ms,
            "note": "collective time is measured on the communication stream while backward runs beside it (it includes "
                    "waiting for the slowest rank at the kernel's first barrier); busbw = S*2(ws-1)/ws/t (nccl-tests)"}


def parity_block(job, world, rank):
    """UNTIMED multi-GPU parity check (the driver's scaling run is the only place more than one GPU is available):
      1. local gradients of one step under ddp.no_sync();  oracle mean = stock torch.distributed all_reduce / world;
         the product's `.grad` after a synchronised backward of the same batch must match it (rel-L2 <= 1e-6, all 16
         tensors) — for the plain reducer and for both clip-before-aggregation modes (README.md:92-95), whose oracle is
         the R3 restatement: clip the LOCAL gradients with torch's own clip_grad_value_ / clip_grad_norm_, then average;
      2. after all of that and the timed region, the parameter arenas of all ranks are bit-identical."""
    import torch
    import torch.distributed as dist
    ddp, model, crit, dev = job.ddp, job.model, job.crit, job.dev
    params = [p for p in model.parameters()]
    x, y = job.batch()
    out = {"ok": True, "checks": {}}
    was = job.opt.overlap_backward
    job.opt.overlap_backward = False            # gradients only: no optimizer step in here

    def backward(sync):
        for p in params:
            p.grad = None
        torch.manual_seed(4321 + rank)          # same dropout seeds in every pass
        if sync:
            crit(ddp(x), y).backward()
        else:
            with ddp.no_sync():
                crit(ddp(x), y).backward()
        torch.cuda.synchronize()
        return [p.grad.detach().clone() for p in params]

    local = backward(False)

    def rel_l2(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))

    def backward_err(a, b, mag):
        """|got - want| against the size of the TERMS that were summed (mean over ranks of |g_r|): the bound fp32
        summation in a different order can be held to even where the terms cancel (clamped gradients of opposite sign)."""
        return float((a.double() - b.double()).norm() / mag.double().norm().clamp_min(1e-30))

    def expect(clip_value=None, clip_norm=None):
        g = [t.clone() for t in local]
        holders = [torch.nn.Parameter(torch.empty_like(t)) for t in g]
        for h, t in zip(holders, g):
            h.grad = t
        if clip_value is not None:
            torch.nn.utils.clip_grad_value_(holders, clip_value)
        if clip_norm is not None:
            torch.nn.utils.clip_grad_norm_(holders, clip_norm)
        res, mags = [], []
        for h in holders:
            t = h.grad
            m = t.abs()
            dist.all_reduce(t)                  # stock c10d NCCL all-reduce as the oracle
            dist.all_reduce(m)
            res.append(t / world)
            mags.append(m / world)
        return res, mags

    gn = float(torch.sqrt(sum((t.double() ** 2).sum() for t in local)))
    gmed = 0.5 * gn / float(sum(t.numel() for t in local)) ** 0.5      # half the RMS gradient entry: clamps a good part
    cases = [("mean", {}, {}), ("clip_value_0.5", {"clip_grad_value": 0.5}, {"clip_value": 0.5}),
             ("clip_value_active", {"clip_grad_value": 4 * gmed}, {"clip_value": 4 * gmed}),
             ("clip_norm_1.0", {"clip_grad_norm": 1.0}, {"clip_norm": 1.0}),
             ("clip_norm_active", {"clip_grad_norm": 0.25 * gn}, {"clip_norm": 0.25 * gn})]
    for name, ours_kw, ref_kw in cases:
        ddp.set_clipping(**ours_kw)
        got = backward(True)
        want, mags = expect(**ref_kw)
        worst = max(rel_l2(a, b) for a, b in zip(got, want))
        worst_b = max(backward_err(a, b, m) for a, b, m in zip(got, want, mags))
        t = torch.tensor([worst, worst_b], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        worst, worst_b = float(t[0]), float(t[1])
        ok = worst_b <= 1e-6 and worst <= 1e-4
        out["checks"][name] = {"max_rel_l2": worst, "max_err_vs_summed_magnitudes": worst_b, "ok": ok}
        out["ok"] = out["ok"] and ok
    ddp.set_clipping()
    out["local_grad_norm"] = gn
    job.opt.overlap_backward = was
    for p in params:
        p.grad = None
    return out


def replicas_identical(job, world):
    import torch
    import torch.distributed as dist
    p = job.ddp.arena.p_flat
    words = p.view(torch.int32)
    digest = torch.stack([words.sum(dtype=torch.int64), (words.to(torch.int64) * 31 % 1000003).sum(dtype=torch.int64),
                          p.double().sum().view(torch.int64)])
    box = [torch.empty_like(digest) for _ in range(world)]
    dist.all_gather(box, digest)
    return all(torch.equal(box[0], t) for t in box)


def gpu_stock_arm(args, dev, world, rank, local, steps):
    """The bar BASELINE.md section 4 names: the reference's own stack — torchvision AlexNet-10, stock
    DistributedDataParallel (N > 1), nn.CrossEntropyLoss, torch.optim.Adam — with torch's default numerics (cuDNN TF32
    convolutions, cuBLAS fp32 Linear, NCCL, foreach Adam) on the SAME GPUs, same batch and timing protocol."""
    import torch
    import torch.distributed as dist
    from oracle import reference_loop as rl
    model = rl.make_model(seed=1234, hidden=args.hidden).to(dev)
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.Adam(net.parameters(), lr=0.001)
    crit = torch.nn.CrossEntropyLoss()
    net.train()
    g = torch.Generator(device=dev).manual_seed(7)
    pool_x = torch.randn(256, 3, 224, 224, device=dev, generator=g)
    pool_y = torch.randint(0, 10, (256,), device=dev, generator=g)
    B = args.batch

    def step(i):
        sel = torch.arange(i * 37, i * 37 + B, device=dev) % 256
        opt.zero_grad()
        loss = crit(net(pool_x.index_select(0, sel)), pool_y.index_select(0, sel))
        loss.backward()
        opt.step()
    for i in range(5):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t) / steps
    del net, model, opt, pool_x
    torch.cuda.empty_cache()
    return {"samples_per_s": world * B / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "torch": torch.__version__,
            "cudnn": torch.backends.cudnn.version(), "nccl": ".".join(map(str, torch.cuda.nccl.version())),
            "numerics": "torch defaults: cuDNN TF32 convolutions, fp32 cuBLAS Linear, foreach Adam",
            "what": "stock torchvision AlexNet-10 + torch DDP + nn.CrossEntropyLoss + torch.optim.Adam on the same GPU(s)"}


def e2e_u8(job, steps, world, sync_all):
    """End to end through the package's public input path (b200ddp.data): every step's batch starts as uint8 CIFAR-format
    images [B,32,32,3] + labels + flip flags in HOST memory, is staged into pinned buffers and copied to the device on
    the loader's copy stream, transformed on the GPU (Resize 224 / flip / ToTensor / Normalize, bit-exact with
    torchvision), trained on, and the step's loss is copied back to pinned host memory."""
    import torch
    import torch.distributed as dist
    import b200ddp
    B, dev = job.B, job.dev
    n_img = 8192
    g = torch.Generator().manual_seed(11)
    images = torch.randint(0, 256, (n_img, 32, 32, 3), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (n_img,), generator=g)
    order = (torch.arange((steps + 4) * B) * 7919 + job.rank * 104729) % n_img

    def loader_for(n):
        return b200ddp.data.DeviceLoader(images, labels, batch_size=B, device=dev, sampler=order[:n * B].tolist(),
                                         transform=b200ddp.data.GpuTransform(224, train=True))
    loss_host = torch.zeros(steps + 4, dtype=torch.float32).pin_memory()

    def run(n):
        k = 0
        for x, y in loader_for(n):
            loss = job.step(x, y)
            loss_host[k].copy_(loss.detach(), non_blocking=True)
            k += 1
        torch.cuda.current_stream(dev).synchronize()
    run(3)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    run(steps)
    e1.record()
    sync_all()
    wall = time.perf_counter() - t0
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t) / steps
    return {"value": world * B / (ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": B * 32 * 32 * 3 + B * 8 + B,
            "d2h_bytes_per_step": 4, "ms_per_step": ms, "wall_ms_per_step": 1e3 * wall / steps,
            "input": "uint8 CIFAR-format images [B,32,32,3] + int64 labels + flip bytes from pinned host memory; "
                     "b200ddp.data.DeviceLoader (copy stream, one batch ahead) + GpuTransform kernel on the device"}


def e2e_fp32(job, steps, world, sync_all):
    """Round-1 definition, kept for continuity: the already-transformed fp32 batch (77 MB) crosses PCIe every step."""
    import torch
    import torch.distributed as dist
    B, dev = job.B, job.dev
    gcpu = torch.Generator().manual_seed(5)
    ring = 4
    host_x = [torch.randn(B, 3, 224, 224, generator=gcpu).pin_memory() for _ in range(ring)]
    host_y = [torch.randint(0, 10, (B,), generator=gcpu).pin_memory() for _ in range(ring)]
    dev_x = [torch.empty(B, 3, 224, 224, device=dev) for _ in range(2)]
    dev_y = [torch.empty(B, dtype=torch.int64, device=dev) for _ in range(2)]
    loss_host = torch.zeros(steps + 2, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(dev)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    main = torch.cuda.current_stream(dev)

    def prefetch(j):
        slot = j % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])
            dev_x[slot].copy_(host_x[j % ring], non_blocking=True)
            dev_y[slot].copy_(host_y[j % ring], non_blocking=True)
            ready[slot].record(copy_stream)

    def run(n, base=0):
        for s in range(2):
            consumed[s].record(main)
        prefetch(0)
        for j in range(n):
            slot = j % 2
            if j + 1 < n:
                prefetch(j + 1)
            main.wait_event(ready[slot])
            loss = job.step(dev_x[slot], dev_y[slot])
            consumed[slot].record(main)
            loss_host[base + j].copy_(loss.detach(), non_blocking=True)
        main.synchronize()
    run(2)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(steps, base=2)
    e1.record()
    sync_all()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t) / steps
    return {"value": world * B / (ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": B * 3 * 224 * 224 * 4 + B * 8,
            "d2h_bytes_per_step": 4, "ms_per_step": ms}


def extra_config(args, dev, world, rank, local, sync_all, name, steps=12, **kw):
    """A few timed steps of another BASELINE.json config, appended to the default line because the driver runs plain
    `bench.py --gpus N`: configs[4] (scaled width through the Accelerate entry) and configs[3] (SyncBatchNorm)."""
    import torch
    try:
        job = Job(args, dev, world, rank, local, **kw)
        job.prepare_indices(3 * steps + 40)
        for _ in range(4):
            job.step_resident()
        ms, loss = timed_steps(job, steps, world, sync_all)
        res = {"samples_per_s": world * job.B / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "per_rank_batch": job.B,
               "params": int(sum(p.numel() for p in job.model.parameters())),
               "grad_bytes_per_step": int(job.ddp.arena.numel * 4), "final_loss": loss, "dtype": "tf32",
               "bucket_bytes": job.ddp.bucket_sizes()}
        if world > 1:
            c = comm_report(job, world, steps, sync_all, ms)
            res.update({"allreduce": c["algorithm"], "busbw_GBps": c["busbw_GBps"], "exposed_comm_ms": c["exposed_comm_ms"],
                        "dry_step_ms": c["dry_step_ms"]})
        job.free()
        return res
    except Exception as e:       # an extra config must never take the headline measurement down with it
        import traceback
        return {"failed": f"{type(e).__name__}: {e}", "trace": traceback.format_exc()[-600:]}


def run_b200(args) -> None:
    import gc

    import torch
    import torch.distributed as dist

    import b200ddp
    from b200ddp import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run --nproc-per-node N")
    assert torch.cuda.is_available(), "bench.py (own arm) needs a GPU: there is no CPU fallback"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    pinned_cores = None
    if world > 1 and os.environ.get("B200DDP_BENCH_PIN", "1") != "0" and hasattr(os, "sched_setaffinity"):
        # one process per GPU, each on its own slice of the host cores: eight launch threads that migrate over and
        # compete for the same cores show up as straggling ranks (every step waits for the slowest one)
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = max(1, len(cores) // world)
            pinned_cores = cores[local * per:(local + 1) * per]
            os.sched_setaffinity(0, pinned_cores)
        except Exception:
            pinned_cores = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    _lib.load()
    B = args.batch

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    job = Job(args, dev, world, rank, local, bn=args.bn)
    ddp, opt = job.ddp, job.opt
    job.prepare_indices(2 * (args.steps + args.warmup) + 200)
    # Reserve the activation pool up front: one block that the caching allocator then carves up, so that no cudaMalloc
    # (tens of ms for multi-hundred-MB segments) happens inside a timed region. 180 GB of HBM per GPU.
    b200ddp.reserve_memory_pool(float(os.environ.get("B200DDP_BENCH_RESERVE_GB", "24")), dev)

    # ---- warm-up -----------------------------------------------------------------------------------------------
    for _ in range(args.warmup):
        job.step_resident()
    sync_all()

    # ---- timed region: device-resident inputs ------------------------------------------------------------------------
    clocks = ClockSampler(local)
    clocks.start()                # every rank samples its own GPU; rank 0's goes into "clocks", the rest into "ranks"
    time.sleep(0.02)
    gc.collect()
    gc.disable()          # a collector pause inside a 0.1-0.5 s timed region is a multi-ms launch stall
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    launches0 = _lib.launch_count()
    sync_all()
    ev0.record()
    for k in range(args.steps):
        loss = job.step_resident()
        marks[k].record()
    ev1.record()
    sync_all()
    launches = _lib.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    final_loss = float(loss.detach())
    step_ms = [marks[k].elapsed_time(marks[k + 1]) for k in range(len(marks) - 1)]    # step-to-step period
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t)
    ms_step = ms_max / args.steps
    # host side: how long the Python thread needs to ENQUEUE a step (8 steps from an idle device: the launch queue never
    # fills, so this is pure host time). Close to ms_step = the step is launch-bound on this rank.
    sync_all()
    th0 = time.perf_counter()
    for _ in range(8):
        job.step_resident()
    host_ms = 1e3 * (time.perf_counter() - th0) / 8
    sync_all()

    # ---- timed regions: end to end -------------------------------------------------------------------------------------
    e2e = e2e_u8(job, args.steps, world, sync_all)
    e2e_f = e2e_fp32(job, min(args.steps, 50), world, sync_all)
    gc.enable()
    clock_info = clocks.stop()                              # covers the resident and the end-to-end regions
    rank_info = None
    if world > 1:
        mine = {"rank": rank, "resident_ms_per_step": ms / args.steps, "host_enqueue_ms_per_step": host_ms,
                "sm_mhz": clock_info.get("sm_mhz"), "sm_min_mhz": clock_info.get("sm_min_mhz"),
                "reasons": clock_info.get("reasons"), "cores": len(pinned_cores) if pinned_cores else None}
        box = [None] * world
        dist.all_gather_object(box, mine)
        rank_info = box
    identical = replicas_identical(job, world) if world > 1 else None

    # ---- communication roofline, multi-GPU parity (untimed) ---------------------------------------------------------------
    nccl, parity = None, None
    if world > 1:
        nccl = comm_report(job, world, args.steps, sync_all, ms_step)
        if not args.no_extras:
            parity = parity_block(job, world, rank)
            parity["replicas_identical_after_timed_region"] = bool(identical)
            parity["replicas_identical_at_end"] = bool(replicas_identical(job, world))
            parity["ok"] = bool(parity["ok"] and identical and parity["replicas_identical_at_end"])

    # ---- per-kernel CUDA-event timing of the same step, serialised (side / optimizer streams off: an event pair around
    #      a kernel that shares the SMs with another stream's kernel measures the interference, not the kernel) ----------
    prev_side = os.environ.get("B200DDP_WGRAD_SIDE_STREAM")
    os.environ["B200DDP_WGRAD_SIDE_STREAM"] = "0"
    prev_engine = os.environ.get("B200DDP_ENGINE")
    os.environ["B200DDP_ENGINE"] = "0"         # per-layer module path: the same kernels, one timed C-ABI call each
    was_overlap, opt.overlap_backward = opt.overlap_backward, False
    _lib.enable_timing()
    n_prof = min(args.steps, 10)
    for _ in range(n_prof):
        job.step_resident()
    calls = _lib.disable_timing()
    opt.overlap_backward = was_overlap
    if prev_engine is None:
        del os.environ["B200DDP_ENGINE"]
    else:
        os.environ["B200DDP_ENGINE"] = prev_engine
    if prev_side is None:
        del os.environ["B200DDP_WGRAD_SIDE_STREAM"]
    else:
        os.environ["B200DDP_WGRAD_SIDE_STREAM"] = prev_side
    kernels = summarize_kernels(calls, n_prof)
    n_arena = ddp.arena.numel
    job.free()

    # ---- the same step on stock torch, and the other BASELINE.json configs (untimed for the headline) --------------------
    stock, extras = None, None
    if not args.no_extras:
        try:
            stock = gpu_stock_arm(args, dev, world, rank, local, steps=min(args.steps, 30))
        except Exception as e:
            stock = {"failed": f"{type(e).__name__}: {e}"}
        extras = {
            "scaled_h8192_accelerate": extra_config(args, dev, world, rank, local, sync_all, "scaled", hidden=8192,
                                                    via_accelerate=True),
            "syncbn_alexnet_bn": extra_config(args, dev, world, rank, local, sync_all, "syncbn", bn=True),
        }
        if os.environ.get("B200DDP_BENCH_H16384", "0") == "1":
            extras["scaled_h16384_accelerate"] = extra_config(args, dev, world, rank, local, sync_all, "scaled16k",
                                                              hidden=16384, via_accelerate=True)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        value = world * B * args.steps / (ms_max * 1e-3)
        tf32_peak = float(peaks.get("bf16_tflops_sustained", 1400.0)) / 2.0     # dense TF32 = half the dense bf16 rate
        roof = roofline_of_dominant(kernels, hbm_peak, tf32_peak, peak_src, ms_step)
        adam_k = next((k for k in kernels if k["kernel"].startswith("adam_step_flat")), None)
        line = {
            "metric": "train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32" if args.mode == "tf32" else "fp32", "data": "synthetic",
            "config": workload_config(args),
            "clocks": clock_info,
            "step_ms_percentiles": ({"p10": sorted(step_ms)[len(step_ms) // 10], "p50": sorted(step_ms)[len(step_ms) // 2],
                                     "p90": sorted(step_ms)[(len(step_ms) * 9) // 10], "max": max(step_ms),
                                     "slowest_step_index": step_ms.index(max(step_ms))} if step_ms else None),
            "e2e": e2e,
            "e2e_fp32_h2d": e2e_f,
            "gpu_launches": int(launches),
            "roofline": roof,
            "roofline_groups": roofline_groups(kernels, ms_step, hbm_peak, tf32_peak),
            "kernels": kernels[:24],
            "final_loss": final_loss,
            "host_enqueue_ms_per_step": host_ms,
            "ranks": rank_info,
            "nccl": nccl,
            "parity_n": parity,
            "gpu_stock": stock,
            "vs_gpu_stock": (value / stock["samples_per_s"]) if stock and stock.get("samples_per_s") else None,
            "extra_configs": extras,
        }
        if adam_k is not None:
            line["roofline_adam"] = {"kernel": "adam_flat_kernel (O1: single-pass Adam over the flat arenas), timed alone in "
                                               "the serialised pass; in the live step it runs per bucket inside backward",
                                     "bound": "hbm", "achieved": adam_k["GBps"], "peak": hbm_peak, "unit": "GB/s",
                                     "frac": (adam_k["GBps"] or 0.0) / hbm_peak,
                                     "traffic": ncu_traffic(f"adam_step_flat n={n_arena}"),
                                     "algorithmic_bytes_per_launch": adam_k["alg_bytes_per_call"],
                                     "avg_launch_ms": adam_k["ms_per_call"], "peak_source": peak_src}
        if world == 1 and not args.no_cpu_baseline and not args.no_extras:
            try:
                from oracle import reference_loop as rl
                cb = rl.time_cpu_train(batch=B, steps=3, warmup=1, hidden=args.hidden)
                line["cpu_baseline"] = {"value": cb["samples_per_s"], "unit": "samples/s", "cores": cb["cores"],
                                        "kind": "port",
                                        "sample": f"3 timed + 1 warm-up CPU training steps of batch {B} "
                                                  f"(stock torch/torchvision AlexNet-10 + Adam, ws=1, {cb['cores']} threads); "
                                                  f"median {cb['ms_per_step']:.0f} ms/step"}
            except Exception as e:  # pragma: no cover
                line["cpu_baseline"] = {"value": None, "unit": "samples/s", "cores": None, "kind": "port",
                                        "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    ok = True
    if parity is not None:
        flag = torch.tensor([1 if parity["ok"] else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(int(flag))
    if world > 1:
        dist.barrier()
        from b200ddp import comm as _comm
        _comm.destroy_all()
        dist.destroy_process_group()
    if not ok:
        raise SystemExit("bench.py: multi-GPU parity check FAILED (see parity_n in the JSON line)")


def main() -> None:
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
