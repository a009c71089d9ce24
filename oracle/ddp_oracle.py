"""CPU oracle for the DDP training step — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; nothing under tutorial-torch-distributed-data-parallel_b200/ does.

The reference (annalena-k/tutorial-torch-distributed-data-parallel) contains no arithmetic of its own:
every number on its hot path is produced by third-party code it calls, un-vendored and unpinned in
requirements.txt:1-3 — torch (installed: 2.11.0+cu128), torchvision (0.26.0+cu128) and accelerate
(absent). This file restates, in plain numpy (float64 accumulation unless noted), the published
algorithms of exactly the upstream routines the reference's call sites reach; each function cites the
reference line that triggers it and the upstream file that defines it.

Parity pin: oracle/make_golden.py runs the real upstream routines (and the reference's own train()
loop, imported from /root/reference) in the build container and stores their outputs under
tests/golden/; tests/test_oracle_golden.py checks every function below against those vectors.
The Accelerate path has no upstream to run here: its parity is UNPINNED.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np

# ----------------------------------------------------------------------------------------------------
# S1  DistributedSampler — reference multi-GPU-training-torch.py:80-83,175-177
#     upstream torch/utils/data/distributed.py:107-141
# ----------------------------------------------------------------------------------------------------


def sampler_indices(perm: Sequence[int], world: int, rank: int, drop_last: bool = False) -> np.ndarray:
    """Indices rank `rank` sees, given the epoch's global permutation `perm` of range(N).

    The permutation comes from torch.randperm(N, Generator(seed + epoch)) — a third-party RNG stream
    (mt19937 + torch's shuffle) that is taken as input, not re-implemented."""
    order = list(perm)
    n = len(order)
    if drop_last and n % world != 0:
        num = math.ceil((n - world) / world)
    else:
        num = math.ceil(n / world)
    total = num * world
    if not drop_last:
        pad = total - n
        if pad > 0:
            order += (order * math.ceil(pad / n))[:pad]
    else:
        order = order[:total]
    return np.asarray(order[rank:total:world], dtype=np.int64)


# ----------------------------------------------------------------------------------------------------
# R1  bucket assignment — reference :245 (DDP ctor); upstream reducer.hpp:592-598,
#     torch/nn/parallel/distributed.py:1219-1244
# ----------------------------------------------------------------------------------------------------


def bucket_assignment(nbytes: Sequence[int], caps: Sequence[int]) -> List[List[int]]:
    buckets: List[List[int]] = []
    cur: List[int] = []
    size = 0
    cap_i = 0
    for i, b in enumerate(nbytes):
        cur.append(i)
        size += b
        if size >= caps[cap_i]:
            buckets.append(cur)
            cur, size = [], 0
            cap_i = min(cap_i + 1, len(caps) - 1)
    if cur:
        buckets.append(cur)
    return buckets


# ----------------------------------------------------------------------------------------------------
# R2/R4  gradient averaging — reference :125 (loss.backward() under DDP)
#        upstream default_comm_hooks.hpp:36-40 (pre-divide) + allreduce(SUM)
# ----------------------------------------------------------------------------------------------------


def ddp_average(grads_per_rank: Sequence[np.ndarray]) -> np.ndarray:
    ws = len(grads_per_rank)
    acc = np.zeros_like(grads_per_rank[0], dtype=np.float64)
    for g in grads_per_rank:
        acc += g.astype(np.float64) * (1.0 / ws)
    return acc.astype(np.float32)


# ----------------------------------------------------------------------------------------------------
# R3  clip before aggregation — reference README.md:92-95
#     upstream torch/nn/utils/clip_grad.py:165-169 (norm), :291-299 (value)
# ----------------------------------------------------------------------------------------------------


def clip_grad_value(g: np.ndarray, c: float) -> np.ndarray:
    out = np.clip(g, -c, c).astype(np.float32)
    out[np.isnan(g)] = np.nan
    return out


def clip_grad_norm(grads: Sequence[np.ndarray], max_norm: float) -> Tuple[List[np.ndarray], float]:
    total = math.sqrt(sum(float(np.sum(g.astype(np.float64) ** 2)) for g in grads))
    coef = max_norm / (total + 1e-6)
    coef = min(coef, 1.0) if coef == coef else coef
    return [(g.astype(np.float64) * coef).astype(np.float32) for g in grads], total


# ----------------------------------------------------------------------------------------------------
# O1  Adam — reference :249 (optim.Adam(lr=0.001)), :126 (step)
#     upstream torch/optim/adam.py:347-551 / :554-800
# ----------------------------------------------------------------------------------------------------


def adam_step(p, g, m, v, t: int, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0):
    """One step in float32 arithmetic (as upstream); returns new (p, m, v)."""
    f = np.float32
    p, g, m, v = (np.asarray(a, dtype=f) for a in (p, g, m, v))
    if weight_decay:
        g = g + f(weight_decay) * p
    m = m + (g - m) * f(1.0 - beta1)
    v = v * f(beta2) + f(1.0 - beta2) * g * g
    bc1 = 1.0 - beta1 ** t
    bc2_sqrt = math.sqrt(1.0 - beta2 ** t)
    denom = np.sqrt(v) / f(bc2_sqrt) + f(eps)
    p = p + f(-(lr / bc1)) * (m / denom)
    return p.astype(f), m.astype(f), v.astype(f)


# ----------------------------------------------------------------------------------------------------
# F1-F3  Linear (+ReLU+Dropout) and cross-entropy — reference :121-122 via torchvision alexnet.py:38-44
# ----------------------------------------------------------------------------------------------------

_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10_raw(c0, c1, c2, c3, k0, k1) -> np.ndarray:
    """Philox4x32-10 (Salmon et al., SC'11) on arrays of 32-bit words; returns [..., 4] uint32."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_PHILOX_M0) * c0
        p1 = np.uint64(_PHILOX_M1) * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & mask
        n1 = p1 & mask
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ k1) & mask
        n3 = p0 & mask
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(_PHILOX_W0)) & mask
        k1 = (k1 + np.uint64(_PHILOX_W1)) & mask
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def philox4x32_10(ctr: np.ndarray, seed: int) -> np.ndarray:
    """Counter words (lo, hi, 0, 0), key (seed_lo, seed_hi) — the layout csrc/common.cuh uses."""
    ctr = np.asarray(ctr, dtype=np.uint64)
    z = np.zeros_like(ctr)
    return philox4x32_10_raw(ctr & np.uint64(0xFFFFFFFF), ctr >> np.uint64(32), z, z, seed & 0xFFFFFFFF,
                             (seed >> 32) & 0xFFFFFFFF)


def dropout_keep_mask(n: int, p: float, seed: int) -> np.ndarray:
    """keep[i] for i in range(n): element i uses counter i//4, word i%4; keep iff word >= p * 2^32."""
    idx = np.arange(n, dtype=np.uint64)
    words = philox4x32_10(idx >> np.uint64(2), seed)
    w = words[np.arange(n), (idx & np.uint64(3)).astype(np.int64)]
    thr = min(int(p * 4294967296.0), 0xFFFFFFFF)
    return w >= np.uint32(thr)


def linear_fwd(x, w, b=None, relu=False, keep: Optional[np.ndarray] = None, p: float = 0.0):
    """y = dropout(relu(x w^T + b)); returns (y float32, gate uint8 [0 off, 1 on, 2 on*1/(1-p)])."""
    z = x.astype(np.float64) @ w.astype(np.float64).T
    if b is not None:
        z = z + b.astype(np.float64)
    gate = np.ones(z.shape, dtype=np.uint8)
    if relu:
        gate[~(z > 0)] = 0
        z = np.where(z > 0, z, 0.0)
    if keep is not None:
        scale = 1.0 / (1.0 - p)
        z = np.where(keep, z * scale, 0.0)
        gate = np.where(keep, np.where(gate > 0, 2, 0), 0).astype(np.uint8)
    return z.astype(np.float32), gate


def linear_bwd(dy, gate, x, w, p: float = 0.0):
    """Returns (dx, dw, db) for y = dropout(relu(x w^T + b)) given the saved gate."""
    scale = 1.0 / (1.0 - p) if p > 0 else 1.0
    g = gate.astype(np.float64)
    factor = np.where(g == 0, 0.0, np.where(g == 2, scale, 1.0))
    dz = dy.astype(np.float64) * factor
    dx = dz @ w.astype(np.float64)
    dw = dz.T @ x.astype(np.float64)
    db = dz.sum(axis=0)
    return dx.astype(np.float32), dw.astype(np.float32), db.astype(np.float32)


def softmax_ce(logits, labels):
    """Mean CE over the batch and its gradient — nn.CrossEntropyLoss() (reference :248)."""
    z = logits.astype(np.float64)
    z = z - z.max(axis=1, keepdims=True)
    lse = np.log(np.exp(z).sum(axis=1, keepdims=True))
    logp = z - lse
    B = z.shape[0]
    loss = -logp[np.arange(B), labels].mean()
    d = np.exp(logp)
    d[np.arange(B), labels] -= 1.0
    return np.float32(loss), (d / B).astype(np.float32)


# ----------------------------------------------------------------------------------------------------
# C6  ReLU + MaxPool2d(3, 2) — torchvision alexnet.py:23-24,26-27,33-34
# ----------------------------------------------------------------------------------------------------


def relu_maxpool_fwd(x, relu=True):
    B, C, H, W = x.shape
    OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
    win = np.empty((B, C, OH, OW, 9), dtype=x.dtype)
    for kh in range(3):
        for kw in range(3):
            win[..., kh * 3 + kw] = x[:, :, kh:kh + 2 * OH - 1:2, kw:kw + 2 * OW - 1:2]
    idx = win.argmax(axis=-1).astype(np.uint8)           # first maximum, like torch
    y = win.max(axis=-1)
    if relu:
        closed = ~(y > 0)
        y = np.where(closed, 0, y)
        idx = np.where(closed, 255, idx).astype(np.uint8)
    return y.astype(np.float32), idx


def relu_maxpool_bwd(dy, idx, in_shape):
    B, C, H, W = in_shape
    OH, OW = dy.shape[2], dy.shape[3]
    dx = np.zeros(in_shape, dtype=np.float64)
    for kh in range(3):
        for kw in range(3):
            sel = (idx == kh * 3 + kw)
            dx[:, :, kh:kh + 2 * OH - 1:2, kw:kw + 2 * OW - 1:2] += np.where(sel, dy, 0.0)
    return dx.astype(np.float32)


# ----------------------------------------------------------------------------------------------------
# B1  SyncBatchNorm — reference README.md:79-81; upstream torch/nn/modules/_functions.py:7-209.
#     Oracle = plain BatchNorm over the concatenated global batch (the method upstream's own test
#     uses, torch/testing/_internal/distributed/distributed_test.py:5583-5640).
# ----------------------------------------------------------------------------------------------------


def bn_local_stats(x, eps=1e-5):
    """Per-rank (mean, invstd) over (B, H, W): aten::batch_norm_stats."""
    xd = x.astype(np.float64)
    mean = xd.mean(axis=(0, 2, 3))
    var = xd.var(axis=(0, 2, 3))
    return mean, 1.0 / np.sqrt(var + eps)


def bn_merge_stats(means, invstds, counts, eps=1e-5):
    """aten::batch_norm_gather_stats_with_counts: returns (mean, invstd, unbiased var)."""
    n = float(sum(counts))
    mean = sum(m * c for m, c in zip(means, counts)) / n
    m2 = sum(((1.0 / i ** 2 - eps) + (m - mean) ** 2) * c for m, i, c in zip(means, invstds, counts))
    var = m2 / n
    return mean, 1.0 / np.sqrt(var + eps), m2 / (n - 1.0)


def syncbn_forward(xs: Sequence[np.ndarray], weight, bias, running_mean, running_var, eps=1e-5, momentum=0.1,
                   relu=False):
    """All ranks' inputs -> (outputs per rank, new running_mean, new running_var, mean, invstd)."""
    allx = np.concatenate([x.astype(np.float64) for x in xs], axis=0)
    mean = allx.mean(axis=(0, 2, 3))
    var = allx.var(axis=(0, 2, 3))
    n = allx.shape[0] * allx.shape[2] * allx.shape[3]
    invstd = 1.0 / np.sqrt(var + eps)
    sh = (1, -1, 1, 1)
    ys = []
    for x in xs:
        y = (x.astype(np.float64) - mean.reshape(sh)) * (invstd * weight.astype(np.float64)).reshape(sh) + \
            bias.astype(np.float64).reshape(sh)
        if relu:
            y = np.where(y > 0, y, 0.0)
        ys.append(y.astype(np.float32))
    rm = (1 - momentum) * running_mean + momentum * mean
    rv = (1 - momentum) * running_var + momentum * var * n / (n - 1)
    return ys, rm.astype(np.float32), rv.astype(np.float32), mean, invstd


def syncbn_backward(dys: Sequence[np.ndarray], xs: Sequence[np.ndarray], weight, mean, invstd,
                    ys_relu: Optional[Sequence[np.ndarray]] = None):
    """Returns (dx per rank, per-rank grad_weight, per-rank grad_bias) — _functions.py:144-209."""
    sh = (1, -1, 1, 1)
    dys = [d.astype(np.float64) for d in dys]
    if ys_relu is not None:
        dys = [np.where(y > 0, d, 0.0) for d, y in zip(dys, ys_relu)]
    xmu = [x.astype(np.float64) - mean.reshape(sh) for x in xs]
    sum_dy_r = [d.sum(axis=(0, 2, 3)) for d in dys]
    sum_dy_xmu_r = [(d * xm).sum(axis=(0, 2, 3)) for d, xm in zip(dys, xmu)]
    n = sum(x.shape[0] * x.shape[2] * x.shape[3] for x in xs)
    sum_dy, sum_dy_xmu = sum(sum_dy_r), sum(sum_dy_xmu_r)
    dxs = []
    for d, xm in zip(dys, xmu):
        dx = (d - (sum_dy / n).reshape(sh) - xm * (invstd ** 2 * sum_dy_xmu / n).reshape(sh)) * \
             (invstd * weight.astype(np.float64)).reshape(sh)
        dxs.append(dx.astype(np.float32))
    gws = [(s * invstd).astype(np.float32) for s in sum_dy_xmu_r]
    gbs = [s.astype(np.float32) for s in sum_dy_r]
    return dxs, gws, gbs


# ---- f1: evaluate() -------------------------------------------------------------------------------------------------
def evaluate_batches(logits_per_batch: Sequence[np.ndarray], labels_per_batch: Sequence[np.ndarray]):
    """Reference evaluate() (multi-GPU-training-torch.py:136-153) given each batch's logits:
    total_test_loss += mean CE * batch_size; predicted = argmax (first maximum, torch.max); correct += (pred == y).
    Returns (total_test_loss, correct, total) as Python floats."""
    total_loss, correct, total = 0.0, 0.0, 0.0
    for z, y in zip(logits_per_batch, labels_per_batch):
        loss, _ = softmax_ce(z, y)
        bs = z.shape[0]
        total_loss += float(loss) * bs
        pred = np.argmax(z, axis=1)              # numpy argmax = first maximal index, like torch.max
        correct += float((pred == y).sum())
        total += bs
    return total_loss, correct, total
