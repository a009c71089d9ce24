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
def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import reference_loop as rl
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    model = rl.make_model(seed=1234, hidden=args.hidden)
    crit = torch.nn.CrossEntropyLoss()
    opt = torch.optim.Adam(model.parameters(), lr=0.001)
    dev = torch.device("cpu")
    batch = args.batch
    x, y = rl.synthetic_pool(batch, seed=7)
    t0 = time.perf_counter()
    rl.train_epoch(model, [(x, y)], crit, opt, dev)          # first (cold) step, also sizes the sample
    first = time.perf_counter() - t0
    budget = 240.0
    if first * (args.steps + args.warmup) > budget and batch > 16:
        batch = max(16, int(batch * budget / (first * (args.steps + args.warmup))) // 16 * 16)
        x, y = x[:batch], y[:batch]
    for _ in range(max(args.warmup - 1, 0)):
        rl.train_epoch(model, [(x, y)], crit, opt, dev)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        rl.train_epoch(model, [(x, y)], crit, opt, dev)
        times.append(time.perf_counter() - t0)
    total = sum(times)
    value = batch * args.steps / total
    sample = (f"{args.steps} timed CPU training steps of batch {batch} (per-rank batch of the workload is {args.batch}); "
              f"stock torch {torch.__version__} + torchvision AlexNet-10 + Adam on {cores} host threads, ws=1")
    line = {"impl": "reference", "metric": "train_samples_per_sec", "value": value, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": workload_config(args, batch_override=batch),
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, batch_override=None) -> dict:
    return {"workload": "AlexNet-10 DDP training step (BASELINE.json configs[1]; configs[2] when n_gpus>1): "
                        "zero_grad + forward + CrossEntropy + backward + bucketed grad all-reduce + Adam",
            "per_rank_batch": batch_override or args.batch, "global_batch": (batch_override or args.batch) * args.gpus,
            "input": "[3,224,224] fp32 synthetic", "hidden": args.hidden, "params": P_PARAMS if args.hidden == 4096 else None,
            "optimizer": "Adam lr=1e-3", "dropout": 0.5, "parallelism": f"dp{args.gpus}",
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
    except Exception:
        pass
    return name[8:], 0.0, 0.0


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
def run_b200(args) -> None:
    import torch
    import torch.distributed as dist

    import b200ddp
    from b200ddp import _lib
    from b200ddp.models import AlexNet

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run --nproc-per-node N")
    assert torch.cuda.is_available(), "bench.py (own arm) needs a GPU: there is no CPU fallback"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    _lib.load()

    B = args.batch
    torch.manual_seed(1234)
    model = AlexNet(hidden=args.hidden, dropout=0.5, mode=args.mode).to(dev)
    ddp = b200ddp.DistributedDataParallel(model, device_ids=[local])
    opt = b200ddp.optim.Adam(ddp.parameters(), lr=0.001)
    crit = b200ddp.CrossEntropyLoss()
    ddp.train()

    # synthetic pool resident in HBM + sampler-driven batch indices (S1 stays on the path)
    gcpu = torch.Generator().manual_seed(7)
    pool_y_cpu = torch.randint(0, 10, (POOL,), generator=gcpu)
    gdev = torch.Generator(device=dev).manual_seed(7)
    pool_x = torch.randn(POOL, 3, 224, 224, device=dev, generator=gdev)
    pool_y = pool_y_cpu.to(dev)

    class _Virtual:
        def __len__(self):
            return VIRTUAL_N
    sampler = b200ddp.DistributedSampler(_Virtual(), num_replicas=world, rank=rank, shuffle=True)
    total_steps = args.steps + args.warmup
    idx = []
    epoch = 0
    while len(idx) < total_steps * B * 2 + B:
        sampler.set_epoch(epoch)
        idx += list(sampler)
        epoch += 1
    idx_dev = (torch.tensor(idx, dtype=torch.int64) % POOL).to(dev)

    def step_resident(i):
        sel = idx_dev[i * B:(i + 1) * B]
        x = pool_x.index_select(0, sel)
        y = pool_y.index_select(0, sel)
        opt.zero_grad()
        loss = crit(ddp(x), y)
        loss.backward()
        opt.step()
        return loss

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # Reserve the activation pool up front: one 24 GB block that the caching allocator then carves up, so that no
    # cudaMalloc (tens of ms for multi-hundred-MB segments, observed as single 30-65 ms steps among 2.34 ms ones)
    # happens inside a timed region. 180 GB of HBM per GPU: this is 13 % of it.
    b200ddp.reserve_memory_pool(float(os.environ.get("B200DDP_BENCH_RESERVE_GB", "24")), dev)

    # ---- warm-up -----------------------------------------------------------------------------------------------
    for i in range(args.warmup):
        step_resident(i)
    sync_all()

    # ---- timed region: device-resident inputs ------------------------------------------------------------------------
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        time.sleep(0.02)
    import gc
    gc.collect()
    gc.disable()          # a collector pause inside a 0.1-0.5 s timed region is a multi-ms launch stall
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    adam_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = _lib.launch_count()
    sync_all()
    ev0.record()
    for k in range(args.steps):
        i = args.warmup + k
        sel = idx_dev[i * B:(i + 1) * B]
        x = pool_x.index_select(0, sel)
        y = pool_y.index_select(0, sel)
        opt.zero_grad()
        loss = crit(ddp(x), y)
        loss.backward()
        adam_ev[k][0].record()
        opt.step()
        adam_ev[k][1].record()
    ev1.record()
    sync_all()
    launches = _lib.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    final_loss = float(loss.detach())
    adam_ms = [a.elapsed_time(b) for a, b in adam_ev]
    step_ms = [adam_ev[k][0].elapsed_time(adam_ev[k + 1][0]) for k in range(len(adam_ev) - 1)]    # step-to-step period
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t)

    # ---- timed region: end to end (pinned host batches -> device, loss read back every step) -----------------------------
    ring = 4
    host_x = [torch.randn(B, 3, 224, 224, generator=gcpu).pin_memory() for _ in range(ring)]
    host_y = [torch.randint(0, 10, (B,), generator=gcpu).pin_memory() for _ in range(ring)]
    dev_x = [torch.empty(B, 3, 224, 224, device=dev) for _ in range(2)]
    dev_y = [torch.empty(B, dtype=torch.int64, device=dev) for _ in range(2)]
    loss_host = torch.zeros(args.steps + 2, dtype=torch.float32).pin_memory()
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

    def e2e_steps(n, record_from=0):
        for s in range(2):
            consumed[s].record(main)
        prefetch(0)
        for j in range(n):
            slot = j % 2
            if j + 1 < n:
                prefetch(j + 1)
            main.wait_event(ready[slot])
            opt.zero_grad()
            loss = crit(ddp(dev_x[slot]), dev_y[slot])
            loss.backward()
            opt.step()
            consumed[slot].record(main)
            loss_host[record_from + j].copy_(loss.detach(), non_blocking=True)   # D2H of this step's loss
        main.synchronize()

    e2e_steps(2)                       # warm the path (pinned buffers, copy stream)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.perf_counter()
    e0.record()
    e2e_steps(args.steps, record_from=2)
    e1.record()
    sync_all()
    e2e_wall = time.perf_counter() - t_wall0
    e2e_ms = e0.elapsed_time(e1)
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms_max = float(te)
    h2d = B * 3 * 224 * 224 * 4 + B * 8
    d2h = 4

    gc.enable()
    # ---- per-kernel CUDA-event timing of the same step (device-resident inputs), for the roofline of the dominant kernel ----
    clock_info = clocks.stop() if rank == 0 else None      # covers both timed regions (resident and end to end)
    # (the side stream that overlaps weight-gradient kernels with dgrad is switched off for this pass: an event pair
    # around a kernel that shares the SMs with another stream's kernel measures the interference, not the kernel)
    prev_side = os.environ.get("B200DDP_WGRAD_SIDE_STREAM")
    os.environ["B200DDP_WGRAD_SIDE_STREAM"] = "0"
    _lib.enable_timing()
    n_prof = min(args.steps, 10)
    for k in range(n_prof):
        step_resident(k)
    calls = _lib.disable_timing()
    if prev_side is None:
        del os.environ["B200DDP_WGRAD_SIDE_STREAM"]
    else:
        os.environ["B200DDP_WGRAD_SIDE_STREAM"] = prev_side
    kernels = summarize_kernels(calls, n_prof)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        n_arena = ddp.arena.numel
        adam_t = statistics.mean(adam_ms) * 1e-3
        adam_bytes = 28.0 * n_arena
        achieved = adam_bytes / adam_t / 1e9
        value = world * B * args.steps / (ms_max * 1e-3)
        tf32_peak = float(peaks.get("bf16_tflops_sustained", 1400.0)) / 2.0     # dense TF32 = half the dense bf16 rate
        roof = roofline_of_dominant(kernels, hbm_peak, tf32_peak, peak_src, ms_max / args.steps)
        line = {
            "metric": "train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32" if args.mode == "tf32" else "fp32", "data": "synthetic",
            "config": workload_config(args),
            "clocks": clock_info,
            "step_ms_percentiles": ({"p10": sorted(step_ms)[len(step_ms) // 10], "p50": sorted(step_ms)[len(step_ms) // 2],
                                     "p90": sorted(step_ms)[(len(step_ms) * 9) // 10], "max": max(step_ms),
                                     "slowest_step_index": step_ms.index(max(step_ms))} if step_ms else None),
            "e2e": {"value": world * B * args.steps / (e2e_ms_max * 1e-3), "unit": "samples/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms_max / args.steps,
                    "wall_ms_per_step": 1e3 * e2e_wall / args.steps},
            "gpu_launches": int(launches),
            "roofline": roof,
            "roofline_adam": {"kernel": "adam_flat_kernel (O1: single-pass Adam over the flat arenas)", "bound": "hbm",
                              "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                              "traffic": ncu_traffic(f"adam_step_flat n={n_arena}"), "algorithmic_bytes_per_launch": adam_bytes,
                              "avg_launch_ms": adam_t * 1e3, "peak_source": peak_src,
                              "share_of_step": adam_t * 1e3 / (ms_max / args.steps)},
            "kernels": kernels[:24],
            "final_loss": final_loss,
            "nccl": None,
        }
        if world == 1 and not args.no_cpu_baseline:
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
    if world > 1:
        dist.barrier()
        from b200ddp import comm as _comm
        _comm.destroy_all()
        dist.destroy_process_group()


def main() -> None:
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
