#!/usr/bin/env python
"""clock64 timeline of CTA (0,0) of conv2d_fwd (developer tool; needs a library built with
B200DDP_NVCC_DEFINES=-DB200_CONV_TRACE). Per k-block: producer thread 0 of group 0 stamps
0 loads issued-from, 6 empty-wait start, 1 slot acquired, 7 stores done, 2 arrived; the MMA thread stamps
3 wait start, 4 full seen, 5 after commit."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from b200ddp import _lib, nn as bnn  # noqa: E402

DEV = torch.device("cuda", 0)
_lib.load()
L = ctypes.CDLL(_lib.lib_path())
L.b200ddp_debug_conv_trace.argtypes = [ctypes.c_void_p]
B = 128
CASES = ((64, 27, 192, 5, 1, 2, "conv2 fwd", False), (64, 27, 192, 5, 1, 2, "conv2 dgrad", True),
         (384, 13, 256, 3, 1, 1, "conv4 fwd", False))
for (Cin, H, Cout, K, s_, p_, tag, dgrad) in CASES:
    conv = bnn.Conv2d(Cin, Cout, K, stride=s_, padding=p_).to(DEV)
    w = conv.weight.detach()
    OH = (H + 2 * p_ - K) // s_ + 1
    if not dgrad:
        x = torch.randn(B, Cin, H, H, device=DEV)
        wm, ld = bnn.conv_weight_matrix(w, 2)
        fn = lambda: bnn.conv2d_forward_raw(x, wm, ld, conv.bias.detach(), Cout, K, K, s_, p_, False, 1)
        nkb = Cin * K * K // 32
    else:
        x = torch.randn(B, Cout, OH, OH, device=DEV)
        wm, ld = bnn.conv_weight_matrix(w, 3)
        fn = lambda: bnn.conv2d_forward_raw(x, wm, ld, None, Cin, K, K, 1, K - 1 - p_, False, 1)
        nkb = Cout * K * K // 32
    fn(); torch.cuda.synchronize()
    buf = torch.zeros(nkb * 8, dtype=torch.int64, device=DEV)
    L.b200ddp_debug_conv_trace(ctypes.c_void_p(buf.data_ptr()))
    fn(); torch.cuda.synchronize()
    L.b200ddp_debug_conv_trace(None)
    t = buf.view(nkb, 8).cpu()
    t0 = int(t[0, 3])
    print(f"== {tag}: {nkb} k-blocks; total MMA loop {int(t[nkb-1,5]) - t0} cycles = {(int(t[nkb-1,5]) - t0)/nkb:.0f} per k-block")
    lo, hi = nkb // 3, min(nkb // 3 + 12, nkb)
    print(" kb | prod: loads@  emptywait@ acquired@ stored@ arrived@ | mma: wait@ full@ committed@ | wait issue period  full-arrive")
    for kb in range(lo, hi):
        r = [int(v) - t0 if int(v) else None for v in t[kb]]
        per = int(t[kb + 1, 3]) - int(t[kb, 3]) if kb + 1 < nkb else 0
        fa = (r[4] - r[2]) if r[2] is not None else None
        print(f"{kb:4d} | {r[0]} {r[6]} {r[1]} {r[7]} {r[2]} | {r[3]} {r[4]} {r[5]} | {r[4]-r[3]} {r[5]-r[4]} {per} {fa}")
    w_ = (t[:, 4] - t[:, 3]).float(); i_ = (t[:, 5] - t[:, 4]).float()
    print(f" mean wait {w_.mean():.0f}  mean issue+commit {i_.mean():.0f}")
    g0 = t[0::3]
    print(f" group 0 means: loads-issue->emptywait {(g0[:,6]-g0[:,0]).float().mean():.0f}  empty wait {(g0[:,1]-g0[:,6]).float().mean():.0f} "
          f" tma+stores {(g0[:,7]-g0[:,1]).float().mean():.0f}  fence+arrive {(g0[:,2]-g0[:,7]).float().mean():.0f} "
          f" arrive->full seen by MMA {(g0[:,4]-g0[:,2]).float().mean():.0f}")
