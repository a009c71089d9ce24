#!/usr/bin/env python
"""Per-kernel timings at the toy model's batch-128 shapes, through the C ABI, CUDA events on the launching
stream, L2 flushed (256 MB write) before every timed launch. Prints one JSON object per kernel with the
algorithmic bytes / FLOPs of DESIGN.md section 4 and the fraction of the measured HBM peak.

    python scripts/kernel_bench.py [--only gemm] [--iters 10] [--prof]     (--prof: 2 un-timed launches each, for ncu)
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")):
    sys.path.insert(0, p)

import torch  # noqa: E402

from b200ddp import _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--only", default="")
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--prof", action="store_true")
ap.add_argument("--mode", type=int, default=1)
args = ap.parse_args()

DEV = torch.device("cuda", 0)
L = _lib.load()
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def st():
    return torch.cuda.current_stream(DEV).cuda_stream


def timeit(name, fn, bytes_alg, flops=0.0, note=""):
    if args.only and args.only not in name:
        return
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if args.prof:
        return
    ts = []
    for _ in range(args.iters):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    gbs = bytes_alg / ms / 1e6
    print(json.dumps({"kernel": name, "ms": round(ms, 4), "min_ms": round(ts[0], 4), "alg_bytes": int(bytes_alg),
                      "GBps": round(gbs, 1), "frac_hbm_peak": round(gbs / PEAK, 3),
                      "TFLOPs": round(flops / ms / 1e9, 2) if flops else None, "note": note}), flush=True)


def chk(rc, what):
    _lib.check(rc, what)


B = 128
mode = args.mode
# ---- Adam ----------------------------------------------------------------------------------------------------------------
n = 57_045_120
p, g, m, v = (torch.randn(n, device=DEV) * 0.01 for _ in range(4))
v.abs_()
timeit("adam_step_flat P=57.0M", lambda: chk(L.b200ddp_adam_step_flat(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n, 1e-3,
                                                                     0.9, 0.999, 1e-8, 0.0, 3, None, 0, st()), "adam"), 28.0 * n)
# ---- reducer kernels -----------------------------------------------------------------------------------------------------
scr = torch.zeros(int(L.b200ddp_sqnorm_scratch_bytes()), dtype=torch.uint8, device=DEV)
sq = torch.zeros(1, device=DEV)
timeit("sqnorm P=57.0M", lambda: chk(L.b200ddp_sqnorm(g.data_ptr(), n, sq.data_ptr(), 0, scr.data_ptr(), scr.numel(), 0, st()), "sqnorm"), 4.0 * n)
timeit("pack_scale_clip (clamp in place) P=57.0M", lambda: chk(L.b200ddp_pack_scale_clip(g.data_ptr(), g.data_ptr(), n, 1.0, 0.5, 0, 0, st()), "clip"), 8.0 * n)
del p, m, v

# ---- Linear layers -------------------------------------------------------------------------------------------------------
for (N, K, tag) in ((4096, 9216, "FC1"), (4096, 4096, "FC2")):
    x = torch.randn(B, K, device=DEV)
    w = torch.randn(N, K, device=DEV) / 64
    bias = torch.randn(N, device=DEV)
    y = torch.empty(B, N, device=DEV)
    gate = torch.empty(B, N, dtype=torch.uint8, device=DEV)
    dz = torch.randn(B, N, device=DEV)
    dx = torch.empty(B, K, device=DEV)
    dw = torch.empty(N, K, device=DEV)
    db = torch.empty(N, device=DEV)
    ws = torch.empty(int(L.b200ddp_linear_workspace_bytes(B, N, K)), dtype=torch.uint8, device=DEV)
    wbytes = 4.0 * N * K
    fl = 2.0 * B * N * K
    timeit(f"gemm linear_fwd {tag} mode{mode} (+bias+relu+dropout+gate, incl. split-K finalize)",
           lambda: chk(L.b200ddp_linear_fwd(x.data_ptr(), w.data_ptr(), bias.data_ptr(), y.data_ptr(), gate.data_ptr(), B, N, K, 1, 0.5,
                                            1234, mode, ws.data_ptr(), ws.numel(), 0, st()), "fwd"), wbytes + 4.0 * B * (N + K), fl)
    timeit(f"gemm linear_bwd_dx {tag} mode{mode}",
           lambda: chk(L.b200ddp_linear_bwd_dx(dz.data_ptr(), w.data_ptr(), dx.data_ptr(), B, N, K, mode, ws.data_ptr(), ws.numel(), 0, st()),
                       "dx"), wbytes + 4.0 * B * (N + K), fl)
    timeit(f"gemm linear_bwd_dw {tag} mode{mode}",
           lambda: chk(L.b200ddp_linear_bwd_dw(dz.data_ptr(), x.data_ptr(), dw.data_ptr(), db.data_ptr(), B, N, K, mode, 0, st()), "dw"),
           wbytes + 4.0 * B * (N + K), fl)
    timeit(f"gate_bias_bwd {tag}", lambda: chk(L.b200ddp_gate_bias_bwd(dz.data_ptr(), gate.data_ptr(), 0.5, y.data_ptr(), db.data_ptr(), B, N, 0,
                                                                      st()), "gate"), 9.0 * B * N)
    for pers, stg, epi in (("1", "3", "8"), ("1", "3", "16"), ("1", "2", "16"), ("1", "6", "16"), ("0", "6", "8")):
        os.environ["B200DDP_GEMM_PERSISTENT"], os.environ["B200DDP_GEMM_PERS_STAGES"] = pers, stg
        os.environ["B200DDP_GEMM_PERS_EPI"] = epi
        timeit(f"gemm linear_bwd_dw {tag} mode1 persistent={pers} stages={stg} epilogue warps={epi}", lambda: chk(
            L.b200ddp_linear_bwd_dw(dz.data_ptr(), x.data_ptr(), dw.data_ptr(), None, B, N, K, 1, 0, st()), "dw"),
            wbytes + 4.0 * B * (N + K), fl)
    for k in ("B200DDP_GEMM_PERSISTENT", "B200DDP_GEMM_PERS_STAGES", "B200DDP_GEMM_PERS_EPI"):
        os.environ.pop(k, None)
    # yardstick only (never on the product path): the library GEMMs stock torch would run for the same three products,
    # without the fused epilogues - cuBLAS fp32 (the reference's Linear numerics) and cuBLAS TF32
    for tf32 in (False, True):
        def lib(fn, tf32=tf32):
            def run():
                old = torch.backends.cuda.matmul.allow_tf32
                torch.backends.cuda.matmul.allow_tf32 = tf32
                fn()
                torch.backends.cuda.matmul.allow_tf32 = old
            return run
        kind = "tf32" if tf32 else "fp32"
        timeit(f"cublas {kind} matmul fwd {tag} (yardstick)", lib(lambda: torch.matmul(x, w.t(), out=y)), wbytes + 4.0 * B * (N + K), fl)
        timeit(f"cublas {kind} matmul dx {tag} (yardstick)", lib(lambda: torch.matmul(dz, w, out=dx)), wbytes + 4.0 * B * (N + K), fl)
        timeit(f"cublas {kind} matmul dw {tag} (yardstick)", lib(lambda: torch.matmul(dz.t(), x, out=dw)), wbytes + 4.0 * B * (N + K), fl)
    del x, w, dw, ws

# ---- head + loss ---------------------------------------------------------------------------------------------------------
x = torch.randn(B, 4096, device=DEV)
w = torch.randn(10, 4096, device=DEV) / 64
bias = torch.randn(10, device=DEV)
z = torch.empty(B, 10, device=DEV)
lab = torch.randint(0, 10, (B,), device=DEV)
loss = torch.zeros((), device=DEV)
dzh = torch.empty(B, 10, device=DEV)
timeit("head linear_fwd N=10", lambda: chk(L.b200ddp_linear_fwd(x.data_ptr(), w.data_ptr(), bias.data_ptr(), z.data_ptr(), None, B, 10, 4096, 0,
                                                               0.0, 0, 0, None, 0, 0, st()), "head"), 4.0 * (B * 4096 + 10 * 4096 + B * 10))
timeit("softmax_ce B=128", lambda: chk(L.b200ddp_softmax_ce(z.data_ptr(), lab.data_ptr(), loss.data_ptr(), dzh.data_ptr(), B, 10, None, 0, st()),
                                       "ce"), 4.0 * B * 20)

# ---- pool / relu ---------------------------------------------------------------------------------------------------------
for (C, H, tag) in ((64, 55, "pool1"), (192, 27, "pool2"), (256, 13, "pool3")):
    xin = torch.randn(B, C, H, H, device=DEV)
    OH = (H - 3) // 2 + 1
    yo = torch.empty(B, C, OH, OH, device=DEV)
    idx = torch.empty(B, C, OH, OH, dtype=torch.uint8, device=DEV)
    dyo = torch.randn_like(yo)
    dxi = torch.empty_like(xin)
    timeit(f"relu_maxpool_fwd {tag}", lambda: chk(L.b200ddp_relu_maxpool_fwd(xin.data_ptr(), yo.data_ptr(), idx.data_ptr(), B, C, H, H, 1, 0, st()),
                                                   "pf"), 4.0 * xin.numel() + 5.0 * yo.numel())
    timeit(f"relu_maxpool_bwd {tag}", lambda: chk(L.b200ddp_relu_maxpool_bwd(dyo.data_ptr(), idx.data_ptr(), dxi.data_ptr(), B, C, H, H, 0, st()),
                                                   "pb"), 4.0 * xin.numel() + 5.0 * yo.numel())
    dbo = torch.empty(C, device=DEV)
    scr = torch.zeros(int(L.b200ddp_bias_grad_scratch_bytes(B, C)), dtype=torch.uint8, device=DEV)
    timeit(f"relu_maxpool_bwd_bias {tag}", lambda: chk(L.b200ddp_relu_maxpool_bwd_bias(dyo.data_ptr(), idx.data_ptr(), dxi.data_ptr(), dbo.data_ptr(),
                                                                                        B, C, H, H, scr.data_ptr(), scr.numel(), 0, st()),
                                                        "pbb"), 4.0 * xin.numel() + 5.0 * yo.numel())
for (C, tag) in ((384, "conv3"), (256, "conv4")):
    yr = torch.randn(B, C, 13, 13, device=DEV)
    gr = torch.randn_like(yr)
    dxr = torch.empty_like(yr)
    dbr = torch.empty(C, device=DEV)
    scr = torch.zeros(int(L.b200ddp_bias_grad_scratch_bytes(B, C)), dtype=torch.uint8, device=DEV)
    timeit(f"relu_bwd_bias {tag}", lambda: chk(L.b200ddp_relu_bwd_bias(gr.data_ptr(), yr.data_ptr(), dxr.data_ptr(), dbr.data_ptr(), B, C, 169,
                                                                       scr.data_ptr(), scr.numel(), 0, st()), "rbb"), 12.0 * yr.numel())
    timeit(f"relu_bwd {tag}", lambda: chk(L.b200ddp_relu_bwd(gr.data_ptr(), yr.data_ptr(), dxr.data_ptr(), yr.numel(), 0, st()), "rb"),
           12.0 * yr.numel())
xin = torch.randn(B, 384, 13, 13, device=DEV)
timeit("relu_fwd conv3", lambda: chk(L.b200ddp_relu_fwd(xin.data_ptr(), xin.numel(), 0, st()), "relu"), 8.0 * xin.numel())

# ---- SyncBN (AlexNet-BN layer 1 shape) -------------------------------------------------------------------------------------
C, HW = 64, 55 * 55
xb = torch.randn(B, C, 55, 55, device=DEV)
scr = torch.zeros(int(L.b200ddp_bn_scratch_bytes(B, C)), dtype=torch.uint8, device=DEV)
stats = torch.empty(2 * C + 1, device=DEV)
timeit("bn_stats [128,64,55,55]", lambda: chk(L.b200ddp_bn_stats(xb.data_ptr(), B, C, HW, 1e-5, stats.data_ptr(), stats.data_ptr() + 4 * C,
                                                                scr.data_ptr(), scr.numel(), 0, st()), "bns"), 4.0 * xb.numel())
yb = torch.empty_like(xb)
wbn = torch.ones(C, device=DEV)
timeit("bn_elemt [128,64,55,55]", lambda: chk(L.b200ddp_bn_elemt(xb.data_ptr(), stats.data_ptr(), stats.data_ptr() + 4 * C, wbn.data_ptr(),
                                                                wbn.data_ptr(), yb.data_ptr(), B, C, HW, 1, 0, st()), "bne"), 8.0 * xb.numel())

# ---- convolutions (implicit GEMM, tcgen05 TF32) ---------------------------------------------------------------------------
from b200ddp import nn as bnn  # noqa: E402

for (Cin, H, Cout, K, s_, p_, tag) in ((3, 224, 64, 11, 4, 2, "conv1"), (64, 27, 192, 5, 1, 2, "conv2"), (192, 13, 384, 3, 1, 1, "conv3"),
                                       (384, 13, 256, 3, 1, 1, "conv4"), (256, 13, 256, 3, 1, 1, "conv5")):
    xin = torch.randn(B, Cin, H, H, device=DEV)
    conv = bnn.Conv2d(Cin, Cout, K, stride=s_, padding=p_).to(DEV)
    OH = (H + 2 * p_ - K) // s_ + 1
    Kk = Cin * K * K
    fl = 2.0 * B * OH * OH * Cout * Kk
    w = conv.weight.detach()
    ko = bnn._k_order(Cin)
    wmat, ldw = bnn.conv_weight_matrix(w, 2 if ko else 0)
    bias = conv.bias.detach()
    yb_ = 4.0 * (xin.numel() + B * Cout * OH * OH + w.numel())
    timeit(f"conv2d_fwd {tag} (k_order {ko})", lambda: bnn.conv2d_forward_raw(xin, wmat, ldw, bias, Cout, K, K, s_, p_, False, ko), yb_, fl)
    if bnn._use_slab(Cin, K, K, s_, Cout, H, p_):
        wm5, ld5 = bnn.conv_weight_matrix(w, 5)
        timeit(f"conv2d_fwd {tag} (k_order 2, slab kernel)", lambda: bnn.conv2d_forward_raw(xin, wm5, ld5, bias, Cout, K, K, s_, p_, False, 2), yb_, fl)
    dyc = torch.randn(B, Cout, OH, OH, device=DEV)
    dwc = torch.empty_like(w)
    ktab = bnn._conv_tables(DEV, Cin, H, H, K, K) if ko == 0 else None
    nb = int(L.b200ddp_conv2d_wgrad_workspace_bytes(B, Cin, H, H, Cout, K, K, s_, p_))
    wsc = torch.empty(max(nb, 256), dtype=torch.uint8, device=DEV)
    timeit(f"conv2d_wgrad {tag} (k_order {ko}, incl. split finalize)",
           lambda: chk(L.b200ddp_conv2d_wgrad(xin.data_ptr(), dyc.data_ptr(), dwc.data_ptr(), _lib.ptr(ktab), ko, B, Cin, H, H,
                                              Cout, K, K, s_, p_, wsc.data_ptr(), wsc.numel(), 0, st()), "wgrad"), yb_, fl)
    if bnn._use_slab(Cin, K, K, s_, Cout, H, p_):
        timeit(f"conv2d_wgrad {tag} (k_order 2, slab kernel, incl. finalize)",
               lambda: chk(L.b200ddp_conv2d_wgrad(xin.data_ptr(), dyc.data_ptr(), dwc.data_ptr(), None, 2, B, Cin, H, H,
                                                  Cout, K, K, s_, p_, wsc.data_ptr(), wsc.numel(), 0, st()), "wgrad"), yb_, fl)
    if s_ == 1:
        kod = bnn._k_order(Cout)
        wt, ldt = bnn.conv_weight_matrix(w, 3 if kod else 1)
        timeit(f"conv2d_dgrad {tag} (k_order {kod})", lambda: bnn.conv2d_forward_raw(dyc, wt, ldt, None, Cin, K, K, 1, K - 1 - p_, False, kod), yb_, fl)
    del xin, dyc
