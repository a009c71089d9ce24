"""Layers of the toy model as autograd Functions over the C ABI (F1-F3, C6, B1).

Module names and parameter layouts follow torch / torchvision so state dicts interchange:
`Linear` keeps `weight [out,in]` and `bias [out]` (torchvision alexnet.py:39,42 and
data_and_toy_model.py:44), `SyncBatchNorm` keeps `weight, bias, running_mean, running_var,
num_batches_tracked` (torch/nn/modules/batchnorm.py:745-842) and
`SyncBatchNorm.convert_sync_batchnorm` mirrors batchnorm.py:844-905 (README.md:79-81 of the
reference prescribes it before wrapping in DDP).

There is no CPU path: these Functions raise for non-CUDA tensors.
"""
from __future__ import annotations

from typing import Optional

import os

import torch
import torch.distributed as dist
from torch import nn
from torch.nn.modules.batchnorm import _BatchNorm

from . import _lib
from .comm import get_comm

MODE_FP32 = 0   # strict fp32 SIMT kernels (parity against the reference's fp32 sgemm)
MODE_TF32 = 1   # tcgen05 TF32 tensor cores, fp32 accumulate (throughput)

_seed_counter = [0]


def _next_seed() -> int:
    """Dropout seed drawn from torch's CPU generator: follows torch.manual_seed, differs per call."""
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


def _mode_id(mode) -> int:
    if mode in (MODE_FP32, "fp32", "strict"):
        return MODE_FP32
    if mode in (MODE_TF32, "tf32"):
        return MODE_TF32
    raise ValueError(f"unknown compute mode {mode!r} (use 'fp32' or 'tf32')")


# ---------------------------------------------------------------------------------------------------------
# Linear (+ bias + ReLU + dropout)
# ---------------------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu, p, seed, mode):
        L = _lib.load()
        dev = _lib.dev_index(x)
        x2 = _lib.f32c(x.reshape(-1, x.shape[-1]), "input")
        w = _lib.f32c(weight, "weight")
        B, K = x2.shape
        N = w.shape[0]
        if w.shape[1] != K:
            raise ValueError(f"Linear: input features {K} != weight in_features {w.shape[1]}")
        y = torch.empty((B, N), dtype=torch.float32, device=x.device)
        # no gate bytes when nothing will be differentiated (evaluate(), reference :136-153: forward under no_grad)
        need_gate = (bool(relu) or p > 0.0) and any(ctx.needs_input_grad[:3])
        gate = torch.empty((B, N), dtype=torch.uint8, device=x.device) if need_gate else None
        ws_bytes = int(L.b200ddp_linear_workspace_bytes(B, N, K))
        ws = _lib.workspace(x.device, ws_bytes, "linear")
        st = _lib.cur_stream(x.device)
        _lib.check(L.b200ddp_linear_fwd(x2.data_ptr(), w.data_ptr(), _lib.ptr(bias), y.data_ptr(), _lib.ptr(gate),
                                        B, N, K, int(bool(relu)), float(p), int(seed), mode, ws.data_ptr(),
                                        ws.numel(), dev, st), "linear_fwd")
        ctx.save_for_backward(x2, w, gate)
        ctx.weight_param = weight
        ctx.has_bias = bias is not None
        ctx.bias_param = bias
        ctx.p = float(p)
        ctx.mode = mode
        ctx.in_shape = x.shape
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        x2, w, gate = ctx.saved_tensors
        dev = _lib.dev_index(x2)
        st = _lib.cur_stream(x2.device)
        B, K = x2.shape
        N = w.shape[0]
        dy2 = _lib.f32c(dy.reshape(B, N), "grad_output")
        need_db = ctx.has_bias and ctx.needs_input_grad[2]
        db = _grad_target(ctx.bias_param, (N,), x2.device)[0] if need_db else None
        if gate is not None:
            dz = torch.empty_like(dy2)
            _lib.check(L.b200ddp_gate_bias_bwd(dy2.data_ptr(), gate.data_ptr(), ctx.p, dz.data_ptr(), _lib.ptr(db),
                                               B, N, dev, st), "gate_bias_bwd")
            db_done = True
        else:
            dz = dy2
            db_done = False
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((B, K), dtype=torch.float32, device=x2.device)
            ws_bytes = int(L.b200ddp_linear_workspace_bytes(B, N, K))
            ws = _lib.workspace(x2.device, ws_bytes, "linear")
            _lib.check(L.b200ddp_linear_bwd_dx(dz.data_ptr(), w.data_ptr(), dx.data_ptr(), B, N, K, ctx.mode,
                                               ws.data_ptr(), ws.numel(), dev, st), "linear_bwd_dx")
            dx = dx.view(ctx.in_shape)
        dw = None
        if ctx.needs_input_grad[1]:
            dw = _grad_target(ctx.weight_param, (N, K), x2.device)[0]
            _lib.check(L.b200ddp_linear_bwd_dw(dz.data_ptr(), x2.data_ptr(), dw.data_ptr(),
                                               None if db_done else _lib.ptr(db), B, N, K, ctx.mode, dev, st),
                       "linear_bwd_dw")
        elif need_db and not db_done:
            _lib.check(L.b200ddp_gate_bias_bwd(dz.data_ptr(), None, 0.0, None, db.data_ptr(), B, N, dev, st),
                       "bias_bwd")
        return dx, dw, db, None, None, None, None


def _grad_target(param, shape, device):
    """Where a wgrad kernel should write, as (tensor, adopted). `adopted` = the tensor is the parameter's own
    arena slice, which autograd will take over as `.grad`: only when no accumulation is pending (`.grad is None`)
    AND this is the parameter's first use in the current backward pass (a Parameter used twice in one graph gets
    the arena view once and a fresh buffer afterwards, so that autograd sums two different tensors instead of
    counting one aliased tensor twice). Otherwise a fresh buffer that autograd adds up."""
    slot = getattr(param, "_b200_slot", None) if param is not None else None
    if slot is not None and param.grad is None and slot[0].claim(slot[1]):
        return slot[0].grad_view(slot[1]), True
    return torch.empty(shape, dtype=torch.float32, device=device), False


class Linear(nn.Linear):
    """nn.Linear with the following ReLU and Dropout of the AlexNet classifier folded into its epilogue."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, relu: bool = False,
                 dropout: float = 0.0, mode="tf32", device=None, dtype=None) -> None:
        super().__init__(in_features, out_features, bias=bias, device=device, dtype=dtype)
        self.relu = bool(relu)
        self.dropout = float(dropout)
        self.mode = _mode_id(mode)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        p = self.dropout if self.training else 0.0
        seed = _next_seed() if p > 0.0 else 0
        return _LinearFn.apply(x, self.weight, self.bias, self.relu, p, seed, self.mode)

    def extra_repr(self) -> str:
        return super().extra_repr() + f", relu={self.relu}, dropout={self.dropout}, mode={self.mode}"


# ---------------------------------------------------------------------------------------------------------
# Conv2d (+ bias + ReLU) as an implicit GEMM on tcgen05 (csrc/conv.cu)
# ---------------------------------------------------------------------------------------------------------
_CONV_TABLES: dict = {}


def _conv_tables(device, Cin: int, H: int, W: int, KH: int, KW: int):
    """int32 device table of pairs indexed by k = (ci, kh, kw): {input offset, window-position bits
    1 << kh | 1 << (16 + kw)} (the kernels AND them with a per-pixel mask of in-image rows / columns).
    Padded to a multiple of 384 pairs with {0, 1 << 31} = never valid (host logic; built once per shape)."""
    key = (str(device), Cin, H, W, KH, KW)
    t = _CONV_TABLES.get(key)
    if t is None:
        K = Cin * KH * KW
        kp = (K + 383) // 384 * 384            # covers 128-row and 3 x 128-row tile grids
        k = torch.arange(kp, dtype=torch.int64)
        ci, rem = k // (KH * KW), k % (KH * KW)
        kh, kw = rem // KW, rem % KW
        koff = (ci * H * W + kh * W + kw).to(torch.int32)
        if KH > 15 or KW > 15:
            raise ValueError("b200ddp Conv2d (table path): kernel sizes above 15 are not supported")
        kpos = ((1 << kh) | (1 << (16 + kw))).to(torch.int32)
        koff[K:] = 0
        kpos[K:] = -(1 << 31)
        t = torch.stack([koff, kpos], dim=1).contiguous().to(device)
        _CONV_TABLES[key] = t
    return t


def _k_order(cin: int) -> int:
    """1 = reduction index ordered (ci//32, kh, kw, ci%32) (needs Cin % 32 == 0), 0 = torch's (ci, kh, kw) + table."""
    return 1 if cin % 32 == 0 else 0


def _use_slab(Cin: int, KH: int, KW: int, stride: int, Cout: int, W: int, pad: int) -> bool:
    """k_order 2: the first layer's own kernel (csrc/conv1.cu) - image slab staged in shared memory by TMA."""
    if os.environ.get("B200DDP_CONV1_SLAB", "1") == "0":
        return False
    return (Cin, KH, KW, stride, Cout, pad) == (3, 11, 11, 4, 64, 2) and W % 4 == 0 and 11 <= W <= 240


def _slab_ld(Cin: int, KH: int) -> int:
    return (Cin * ((KH * 12 + 7) // 8 * 8) + 31) // 32 * 32


def conv2d_forward_raw(x, wmat, ldw, bias, Cout, KH, KW, stride, pad, relu, k_order):
    """y = conv(x, wmat) through b200ddp_conv2d_fwd; x [N,Cin,H,W] contiguous fp32 CUDA; wmat in `k_order`."""
    L = _lib.load()
    N, Cin, H, W = x.shape
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    ktab = _conv_tables(x.device, Cin, H, W, KH, KW) if k_order == 0 else None
    y = torch.empty((N, Cout, OH, OW), dtype=torch.float32, device=x.device)
    _lib.check(L.b200ddp_conv2d_fwd(x.data_ptr(), wmat.data_ptr(), int(ldw), _lib.ptr(bias), y.data_ptr(),
                                    _lib.ptr(ktab), k_order, N, Cin, H, W, Cout, KH, KW, stride, pad,
                                    int(bool(relu)), _lib.dev_index(x), _lib.cur_stream(x.device)), "conv2d_fwd")
    return y


def conv_weight_matrix(w, mode: int):
    """Weight matrix for the implicit GEMM (b200ddp_conv2d_weight_prep): mode 0 padded copy / 2 channel-block-major
    forward weights [Cout][K]; 1 / 3 flipped, transposed dgrad weights [Cin][Cout*KH*KW]."""
    L = _lib.load()
    Cout, Cin, KH, KW = w.shape
    if mode == 5:
        rows, ld = Cout, _slab_ld(Cin, KH)
    elif mode in (0, 2):
        rows, ld = Cout, ((Cin * KH * KW + 31) // 32 * 32 if mode == 0 else Cin * KH * KW)
    else:
        rows, ld = Cin, Cout * KH * KW
    out = torch.empty((rows, ld), dtype=torch.float32, device=w.device)
    _lib.check(L.b200ddp_conv2d_weight_prep(w.data_ptr(), out.data_ptr(), Cout, Cin, KH, KW, ld, mode,
                                            _lib.dev_index(w), _lib.cur_stream(w.device)), "weight_prep")
    return out, ld


# ---- side stream for weight gradients -------------------------------------------------------------------------------
# dgrad and wgrad of a convolution depend only on the output gradient. dgrad is on the critical path of the backward
# pass; at these sizes its grid is 169 or 338 tiles on 148 SMs, i.e. the last wave leaves most SMs idle. wgrad (one wave
# of ~140 CTAs) is launched on a second stream so that the block scheduler fills those idle SMs with it. The main
# stream re-joins before anything reads the weight gradients: the bucket all-reduce (reducer.py), optimizer.step and an
# end-of-backward engine callback (so `.grad` is safe on the current stream when backward() returns).
_SIDE = {}


class _SideWork:
    def __init__(self, device) -> None:
        self.stream = torch.cuda.Stream(device)
        self.pending = []
        self.callback_queued = False


def _side(device) -> "_SideWork":
    key = (device.type, device.index)
    w = _SIDE.get(key)
    if w is None:
        w = _SIDE[key] = _SideWork(device)
    return w


def join_side_work(device=None) -> None:
    """Make the current stream wait for every weight-gradient kernel still running on the side stream."""
    for key, w in _SIDE.items():
        if device is not None and (device.type, device.index) != key:
            continue
        if w.pending:
            cur = torch.cuda.current_stream(w.stream.device)
            for ev in w.pending:
                cur.wait_event(ev)
            w.pending.clear()
        w.callback_queued = False


def _overlap_wgrad() -> bool:
    return os.environ.get("B200DDP_WGRAD_SIDE_STREAM", "1") != "0"


def _use_s2d(Cin: int, KH: int, KW: int, stride: int) -> bool:
    """Space-to-depth form of a strided convolution over few channels (the toy model's first layer)."""
    if os.environ.get("B200DDP_CONV1_S2D", "0") != "1":
        return False
    return 2 <= stride <= 8 and Cin * stride * stride <= 64 and KH == KW and KH > stride


_S2D_BUFS = {}


def _s2d_geometry(H: int, W: int, KH: int, KW: int, stride: int, pad: int):
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    KHp, KWp = (KH + stride - 1) // stride, (KW + stride - 1) // stride
    return OH, OW, KHp, KWp, OH + KHp - 1, OW + KWp - 1


def _conv_forward_s2d(x, w, bias, stride, pad, relu, prepared):
    """y = conv(x, w) through b200ddp_space_to_depth + the channel-block-major stride-1 kernel."""
    L = _lib.load()
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    OH, OW, KHp, KWp, Hs, Ws = _s2d_geometry(H, W, KH, KW, stride, pad)
    key = (str(x.device), N, Cin, H, W, stride, pad, KH, KW)
    xs = _S2D_BUFS.get(key)
    if xs is None:       # zeroed once: the kernel never writes the border / the channels beyond Cin*S*S
        xs = torch.zeros((N, 64, Hs, Ws), dtype=torch.float32, device=x.device)
        _S2D_BUFS.clear()
        _S2D_BUFS[key] = xs
    dev, st = _lib.dev_index(x), _lib.cur_stream(x.device)
    _lib.check(L.b200ddp_space_to_depth(x.data_ptr(), xs.data_ptr(), N, Cin, H, W, stride, pad, 64, Hs, Ws, dev, st),
               "space_to_depth")
    if prepared is not None and "fwd_s2d" in prepared:
        wmat, ldw = prepared["fwd_s2d"]
    else:
        ldw = 64 * KHp * KWp
        wmat = torch.empty((Cout, ldw), dtype=torch.float32, device=w.device)
        _lib.check(L.b200ddp_conv2d_weight_prep(w.data_ptr(), wmat.data_ptr(), Cout, Cin, KH, KW, ldw, 4 | (stride << 8),
                                                dev, st), "weight_prep(s2d)")
    y = conv2d_forward_raw(xs, wmat, ldw, bias, Cout, KHp, KWp, 1, 0, relu, 1)
    assert y.shape[-2:] == (OH, OW)
    return y


class _Conv2dFn(torch.autograd.Function):
    """conv (+bias) [+ReLU in the epilogue] [+ReLU + 3x3/2 max pool]. Backward: the pool / ReLU backward kernel also
    produces the bias gradient (sum of the conv output gradient over batch and plane), then wgrad and dgrad."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, relu, pool, prepared, mode):
        L = _lib.load()
        x = _lib.f32c(x, "input")
        w = _lib.f32c(weight, "weight")
        Cout, Cin, KH, KW = w.shape
        if x.shape[1] != Cin:
            raise ValueError(f"Conv2d: input has {x.shape[1]} channels, weight expects {Cin}")
        K = Cin * KH * KW
        ko = _k_order(Cin)
        ctx.dgrad_weights = None
        ctx.mode = mode
        if mode == MODE_FP32:
            # strict mode: direct fp32 convolution on torch's own weight layout (no GEMM copies, no TF32 rounding)
            N, _, H, W = x.shape
            OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
            y = torch.empty((N, Cout, OH, OW), dtype=torch.float32, device=x.device)
            _lib.check(L.b200ddp_conv2d_fwd_fp32(x.data_ptr(), w.data_ptr(), _lib.ptr(bias), y.data_ptr(), N, Cin, H, W,
                                                 Cout, KH, KW, stride, pad, int(bool(relu and not pool)),
                                                 _lib.dev_index(x), _lib.cur_stream(x.device)), "conv2d_fwd_fp32")
        elif _use_s2d(Cin, KH, KW, stride):
            # few input channels + stride S (conv1): space-to-depth turns it into a stride-1 ceil(K/S)^2 convolution
            # over 64 channels whose gather reads consecutive pixels; the backward (wgrad) keeps the original x
            y = _conv_forward_s2d(x, w, bias, stride, pad, relu and not pool, prepared)
        elif prepared is not None:      # GEMM copies made by prepare_conv_weights() for this forward/backward pair
            wmat, ldw = prepared["fwd"]
            ctx.dgrad_weights = prepared.get("dgrad")
            y = conv2d_forward_raw(x, wmat, ldw, bias, Cout, KH, KW, stride, pad, relu and not pool, prepared.get("ko", ko))
        else:
            if _use_slab(Cin, KH, KW, stride, Cout, x.shape[3], pad):
                wmat, ldw = conv_weight_matrix(w, 5)
                ko = 2
            elif ko == 1:
                wmat, ldw = conv_weight_matrix(w, 2)
            elif K % 4 == 0 and w.data_ptr() % 16 == 0:
                wmat, ldw = w, K
            else:   # TMA needs 16-byte aligned rows: zero-padded copy
                wmat, ldw = conv_weight_matrix(w, 0)
            y = conv2d_forward_raw(x, wmat, ldw, bias, Cout, KH, KW, stride, pad, relu and not pool, ko)
        ctx.stride, ctx.pad, ctx.relu, ctx.pool = stride, pad, relu, pool
        ctx.weight_param, ctx.bias_param = weight, bias
        if pool:
            B, C, H, W = y.shape
            if H < 3 or W < 3:
                raise ValueError("Conv2d(pool=True): the convolution output is smaller than the 3x3 pooling window")
            OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
            out = torch.empty((B, C, OH, OW), dtype=torch.float32, device=x.device)
            train = any(ctx.needs_input_grad[:3])          # inference: no argmax codes written or kept
            idx = torch.empty((B, C, OH, OW), dtype=torch.uint8, device=x.device) if train else None
            _lib.check(L.b200ddp_relu_maxpool_fwd(y.data_ptr(), out.data_ptr(), _lib.ptr(idx), B, C, H, W, int(relu),
                                                  _lib.dev_index(x), _lib.cur_stream(x.device)), "relu_maxpool_fwd")
            ctx.save_for_backward(x, w, idx)
            return out
        ctx.save_for_backward(x, w, y if relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        x, w, aux = ctx.saved_tensors
        dev, st = _lib.dev_index(x), _lib.cur_stream(x.device)
        N, Cin, H, W = x.shape
        Cout, _, KH, KW = w.shape
        stride, pad = ctx.stride, ctx.pad
        OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
        dy = _lib.f32c(dy, "grad_output")
        want_db = ctx.bias_param is not None and ctx.needs_input_grad[2]
        db = _grad_target(ctx.bias_param, (Cout,), x.device)[0] if want_db else None
        have_db = False
        if want_db and (ctx.pool or ctx.relu):
            nb = int(L.b200ddp_bias_grad_scratch_bytes(N, Cout))
            scratch = _lib.workspace(x.device, nb, "bias_grad")
        if ctx.pool:
            dz = torch.empty((N, Cout, OH, OW), dtype=torch.float32, device=x.device)
            if want_db:
                _lib.check(L.b200ddp_relu_maxpool_bwd_bias(dy.data_ptr(), aux.data_ptr(), dz.data_ptr(), db.data_ptr(), N,
                                                           Cout, OH, OW, scratch.data_ptr(), scratch.numel(), dev, st),
                           "relu_maxpool_bwd_bias")
                have_db = True
            else:
                _lib.check(L.b200ddp_relu_maxpool_bwd(dy.data_ptr(), aux.data_ptr(), dz.data_ptr(), N, Cout, OH, OW, dev,
                                                      st), "relu_maxpool_bwd")
        elif ctx.relu:
            dz = torch.empty_like(dy)
            if want_db:
                _lib.check(L.b200ddp_relu_bwd_bias(dy.data_ptr(), aux.data_ptr(), dz.data_ptr(), db.data_ptr(), N, Cout,
                                                   OH * OW, scratch.data_ptr(), scratch.numel(), dev, st), "relu_bwd_bias")
                have_db = True
            else:
                _lib.check(L.b200ddp_relu_bwd(dy.data_ptr(), aux.data_ptr(), dz.data_ptr(), dy.numel(), dev, st), "relu_bwd")
        else:
            dz = dy
        if want_db and not have_db:
            nb = int(L.b200ddp_channel_sum_scratch_bytes(N, Cout))
            scratch = _lib.workspace(x.device, nb, "chsum", zero=True)
            _lib.check(L.b200ddp_channel_sum(dz.data_ptr(), db.data_ptr(), N, Cout, OH * OW, scratch.data_ptr(),
                                             scratch.numel(), dev, st), "channel_sum")
        # dgrad first: it is on the critical path and should own the SMs' first wave; wgrad then fills its tail
        if ctx.mode == MODE_FP32:
            dx = dw = None
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x)
                _lib.check(L.b200ddp_conv2d_dgrad_fp32(dz.data_ptr(), w.data_ptr(), dx.data_ptr(), N, Cin, H, W, Cout, KH,
                                                       KW, stride, pad, dev, st), "conv2d_dgrad_fp32")
            if ctx.needs_input_grad[1]:
                dw = _grad_target(ctx.weight_param, tuple(w.shape), x.device)[0]
                _lib.check(L.b200ddp_conv2d_wgrad_fp32(x.data_ptr(), dz.data_ptr(), dw.data_ptr(), N, Cin, H, W, Cout, KH,
                                                       KW, stride, pad, dev, st), "conv2d_wgrad_fp32")
            return dx, dw, db, None, None, None, None, None, None
        dx = None
        if ctx.needs_input_grad[0]:
            if stride != 1:
                raise NotImplementedError("b200ddp Conv2d: input gradients are implemented for stride 1 only "
                                          "(the toy model's strided conv is its first layer); mode='fp32' has no such limit")
            ko = _k_order(Cout)          # dgrad reduces over (co, kh', kw')
            if ctx.dgrad_weights is not None:
                wt, ldt = ctx.dgrad_weights
            else:
                wt, ldt = conv_weight_matrix(w, 3 if ko == 1 else 1)
            dx = conv2d_forward_raw(dz, wt, ldt, None, Cin, KH, KW, 1, KH - 1 - pad, False, ko)
        dw = None
        if ctx.needs_input_grad[1]:
            dw, adopted = _grad_target(ctx.weight_param, tuple(w.shape), x.device)
            ko = 2 if _use_slab(Cin, KH, KW, stride, Cout, W, pad) else _k_order(Cin)
            ktab = _conv_tables(x.device, Cin, H, W, KH, KW) if ko == 0 else None
            nb = int(L.b200ddp_conv2d_wgrad_workspace_bytes(N, Cin, H, W, Cout, KH, KW, stride, pad))
            wst = st
            side = None
            # The side stream is used only when (a) there is a dgrad to overlap with and (b) the kernel writes the
            # parameter's own arena slice: a fresh buffer would be summed into `.grad` by autograd on the MAIN stream
            # (gradient accumulation, no_sync) with nothing ordering that add after the side-stream kernel.
            if ctx.needs_input_grad[0] and adopted and _overlap_wgrad():
                side = _side(x.device)
                main = torch.cuda.current_stream(x.device)
                ready = torch.cuda.Event()
                ready.record(main)
                side.stream.wait_event(ready)
                for t in (x, dz):
                    t.record_stream(side.stream)
                wst = side.stream.cuda_stream
                # split-K partials: the two streams never share a workspace (a main-stream wgrad, e.g. the first
                # layer's, could otherwise overwrite partials a side-stream finalize has not read yet); the side
                # buffer is allocated under the side stream so the caching allocator orders its reuse there
                ws = _lib.workspace(x.device, max(nb, 256), "conv_wgrad_side", stream=side.stream)
            else:
                ws = _lib.workspace(x.device, max(nb, 256), "conv_wgrad")
            _lib.check(L.b200ddp_conv2d_wgrad(x.data_ptr(), dz.data_ptr(), dw.data_ptr(), _lib.ptr(ktab), ko, N, Cin, H,
                                              W, Cout, KH, KW, stride, pad, ws.data_ptr(), ws.numel(), dev, wst),
                       "conv2d_wgrad")
            if side is not None:
                done = torch.cuda.Event()
                done.record(side.stream)
                side.pending.append(done)
                if not side.callback_queued:
                    side.callback_queued = True
                    torch.autograd.Variable._execution_engine.queue_callback(lambda d=x.device: join_side_work(d))
        return dx, dw, db, None, None, None, None, None, None


class Conv2d(nn.Conv2d):
    """nn.Conv2d (square stride / padding, no dilation / groups) with the following ReLU (`relu=True`) or the
    following ReLU + MaxPool2d(3, 2) (`pool=True`, alexnet.py:22-24) optionally folded in."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True, relu=False, pool=False,
                 mode="tf32", device=None, dtype=None) -> None:
        super().__init__(in_channels, out_channels, kernel_size, stride=stride, padding=padding, bias=bias,
                         device=device, dtype=dtype)
        if self.stride[0] != self.stride[1] or self.padding[0] != self.padding[1]:
            raise ValueError("b200ddp Conv2d needs equal stride / padding in both dimensions")
        self.mode = _mode_id(mode)                # 'tf32': tcgen05 implicit GEMM (conv.cu); 'fp32': strict SIMT (conv_fp32.cu)
        self.pool = bool(pool)
        self.relu = bool(relu) or self.pool       # the pooled form always includes the ReLU in front of the pool
        self._prepared = None                     # set by prepare_conv_weights(), consumed by the next forward
        self._prep_bufs = {}

    def _prep_jobs(self, need_dgrad: bool, in_width=None):
        """(jobs, prepared) for prepare_conv_weights: GEMM copies of the weights for forward and, optionally, dgrad."""
        w = self.weight.data
        Cout, Cin, KH, KW = w.shape
        K = Cin * KH * KW
        jobs, prepared = [], {}

        def buf(tag, rows, ld):
            t = self._prep_bufs.get(tag)
            if t is None or t.shape != (rows, ld) or t.device != w.device:
                t = torch.empty((rows, ld), dtype=torch.float32, device=w.device)
                self._prep_bufs[tag] = t
            return t
        if _use_s2d(Cin, KH, KW, self.stride[0]):
            S = self.stride[0]
            ld = 64 * ((KH + S - 1) // S) * ((KW + S - 1) // S)
            out = buf("fwd_s2d", Cout, ld)
            jobs.append((w, out, 4 | (S << 8), ld))
            prepared["fwd_s2d"] = (out, ld)
        elif in_width is not None and _use_slab(Cin, KH, KW, self.stride[0], Cout, in_width, self.padding[0]):
            ld = _slab_ld(Cin, KH)
            out = buf("fwd", Cout, ld)
            jobs.append((w, out, 5, ld))
            prepared["fwd"], prepared["ko"] = (out, ld), 2
        elif _k_order(Cin) == 1:
            out = buf("fwd", Cout, K)
            jobs.append((w, out, 2, K))
            prepared["fwd"] = (out, K)
        elif K % 4 == 0 and w.data_ptr() % 16 == 0:
            prepared["fwd"] = (w, K)
        else:
            ld = (K + 31) // 32 * 32
            out = buf("fwd", Cout, ld)
            jobs.append((w, out, 0, ld))
            prepared["fwd"] = (out, ld)
        if need_dgrad and self.stride[0] == 1:
            ld = Cout * KH * KW
            out = buf("dgrad", Cin, ld)
            jobs.append((w, out, 3 if _k_order(Cout) == 1 else 1, ld))
            prepared["dgrad"] = (out, ld)
        return jobs, prepared

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        prepared, self._prepared = self._prepared, None
        return _Conv2dFn.apply(x, self.weight, self.bias, self.stride[0], self.padding[0], self.relu, self.pool, prepared,
                               self.mode)


def prepare_conv_weights(convs, first_needs_dgrad: bool = False, in_width=None) -> None:
    """One launch (b200ddp_conv2d_weight_prep_multi) that writes the GEMM copies of the weights of every Conv2d in
    `convs` (in network order) for the forward pass that follows and its backward: forward copies for all layers,
    dgrad copies for all but the first (whose input needs no gradient unless `first_needs_dgrad`)."""
    L = _lib.load()
    convs = list(convs)
    if not convs:
        return
    jobs = []
    for i, conv in enumerate(convs):
        if (not conv.weight.is_cuda or conv.weight.dtype != torch.float32 or not conv.weight.is_contiguous()
                or conv.mode == MODE_FP32):
            conv._prepared = None
            continue
        j, prepared = conv._prep_jobs(need_dgrad=torch.is_grad_enabled() and (i > 0 or first_needs_dgrad),
                                      in_width=in_width if i == 0 else None)
        jobs += j
        conv._prepared = prepared
    if not jobs:
        return
    dev = convs[0].weight.device
    for k in range(0, len(jobs), 16):
        part = jobs[k:k + 16]
        arr = (_lib.WeightPrepJob * len(part))()
        for a, (w, out, mode, ld) in zip(arr, part):
            Cout, Cin, KH, KW = w.shape
            a.w, a.out, a.Cout, a.Cin, a.KH, a.KW, a.ld_out, a.mode = w.data_ptr(), out.data_ptr(), Cout, Cin, KH, KW, ld, mode
        _lib.check(L.b200ddp_conv2d_weight_prep_multi(arr, len(part), dev.index or 0, _lib.cur_stream(dev)),
                   "conv2d_weight_prep_multi")


# ---------------------------------------------------------------------------------------------------------
# softmax cross-entropy
# ---------------------------------------------------------------------------------------------------------
class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        L = _lib.load()
        dev = _lib.dev_index(logits)
        z = _lib.f32c(logits, "logits")
        if z.dim() != 2:
            raise ValueError("CrossEntropyLoss expects [batch, classes] logits")
        if labels.dtype != torch.int64:
            raise TypeError("CrossEntropyLoss expects int64 class indices")
        y = labels.contiguous()
        B, C = z.shape
        _check_label_flag()                 # an out-of-range label seen by an EARLIER call surfaces here
        loss = torch.empty((), dtype=torch.float32, device=z.device)
        dlogits = torch.empty_like(z) if ctx.needs_input_grad[0] else None      # evaluate(): loss only
        _lib.check(L.b200ddp_softmax_ce(z.data_ptr(), y.data_ptr(), loss.data_ptr(), _lib.ptr(dlogits), B, C,
                                        _label_flag().data_ptr(), dev, _lib.cur_stream(z.device)), "softmax_ce")
        if dlogits is not None:
            ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dlogits,) = ctx.saved_tensors
        L = _lib.load()
        _check_label_flag()
        g = gout.to(torch.float32).reshape(1).contiguous()
        out = dlogits.clone()               # the saved tensor stays intact (retain_graph / a second backward)
        _lib.check(L.b200ddp_scale_by_device_scalar(out.data_ptr(), out.numel(), g.data_ptr(), 0,
                                                    _lib.dev_index(out), _lib.cur_stream(out.device)), "ce_bwd")
        return out, None


_LABEL_FLAG = []


def _label_flag() -> torch.Tensor:
    """One int32 in pinned (device-mapped) host memory that softmax_ce raises for a label outside [0, C): the host
    reads it without synchronising, so the report arrives asynchronously - like stock torch's device assert - at the
    next loss / backward call that runs after the kernel has executed."""
    if not _LABEL_FLAG:
        _LABEL_FLAG.append(torch.zeros(1, dtype=torch.int32).pin_memory())
    return _LABEL_FLAG[0]


def _check_label_flag() -> None:
    if _LABEL_FLAG and int(_LABEL_FLAG[0][0]) != 0:
        _LABEL_FLAG[0][0] = 0
        raise _lib.B200Error("CrossEntropyLoss: a class index outside [0, n_classes) other than ignore_index=-100 "
                             "was passed in an earlier call (that row was given zero loss and zero gradient)")


class CrossEntropyLoss(nn.Module):
    """Mean softmax cross-entropy over the batch (nn.CrossEntropyLoss() defaults incl. ignore_index=-100, reference :248)."""

    def forward(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return _CrossEntropyFn.apply(logits, labels)


# ---------------------------------------------------------------------------------------------------------
# ReLU, ReLU+MaxPool(3,2), Dropout
# ---------------------------------------------------------------------------------------------------------
class _ReluMaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, relu):
        L = _lib.load()
        x = _lib.f32c(x, "input")
        B, C, H, W = x.shape
        OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
        y = torch.empty((B, C, OH, OW), dtype=torch.float32, device=x.device)
        idx = torch.empty((B, C, OH, OW), dtype=torch.uint8, device=x.device) if ctx.needs_input_grad[0] else None
        _lib.check(L.b200ddp_relu_maxpool_fwd(x.data_ptr(), y.data_ptr(), _lib.ptr(idx), B, C, H, W, int(relu),
                                              _lib.dev_index(x), _lib.cur_stream(x.device)), "relu_maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.in_shape = (B, C, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        (idx,) = ctx.saved_tensors
        B, C, H, W = ctx.in_shape
        dy = _lib.f32c(dy, "grad_output")
        dx = torch.empty(ctx.in_shape, dtype=torch.float32, device=dy.device)
        _lib.check(L.b200ddp_relu_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), B, C, H, W,
                                              _lib.dev_index(dy), _lib.cur_stream(dy.device)), "relu_maxpool_bwd")
        return dx, None


class ReLUMaxPool2d(nn.Module):
    """nn.ReLU followed by nn.MaxPool2d(kernel_size=3, stride=2) in one pass (alexnet.py:23-24)."""

    def __init__(self, relu: bool = True) -> None:
        super().__init__()
        self.relu = relu

    def forward(self, x):
        return _ReluMaxPoolFn.apply(x, self.relu)


class _ReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        L = _lib.load()
        if not x.is_contiguous():
            x = x.contiguous()
        _lib.f32c(x, "input")
        _lib.check(L.b200ddp_relu_fwd(x.data_ptr(), x.numel(), _lib.dev_index(x), _lib.cur_stream(x.device)),
                   "relu_fwd")
        ctx.mark_dirty(x)
        ctx.save_for_backward(x)
        return x

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        (y,) = ctx.saved_tensors
        dy = _lib.f32c(dy, "grad_output")
        dx = torch.empty_like(dy)
        _lib.check(L.b200ddp_relu_bwd(dy.data_ptr(), y.data_ptr(), dx.data_ptr(), dy.numel(), _lib.dev_index(dy),
                                      _lib.cur_stream(dy.device)), "relu_bwd")
        return dx


class ReLU(nn.Module):
    """In-place ReLU (nn.ReLU(inplace=True), alexnet.py:29,31)."""

    def forward(self, x):
        return _ReluFn.apply(x)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        L = _lib.load()
        x = _lib.f32c(x, "input")
        y = torch.empty_like(x)
        gate = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
        _lib.check(L.b200ddp_dropout_fwd(x.data_ptr(), y.data_ptr(), gate.data_ptr(), x.numel(), float(p), int(seed),
                                         _lib.dev_index(x), _lib.cur_stream(x.device)), "dropout_fwd")
        ctx.save_for_backward(gate)
        ctx.p = float(p)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        (gate,) = ctx.saved_tensors
        dy = _lib.f32c(dy, "grad_output")
        dx = torch.empty_like(dy)
        n_last = dy.shape[-1]
        rows = dy.numel() // n_last
        _lib.check(L.b200ddp_gate_bias_bwd(dy.data_ptr(), gate.data_ptr(), ctx.p, dx.data_ptr(), None, rows, n_last,
                                           _lib.dev_index(dy), _lib.cur_stream(dy.device)), "dropout_bwd")
        return dx, None, None


class Dropout(nn.Module):
    """nn.Dropout(p) with a Philox mask that does not depend on launch geometry (alexnet.py:38,41)."""

    def __init__(self, p: float = 0.5) -> None:
        super().__init__()
        if not 0.0 <= p < 1.0:
            raise ValueError("dropout probability has to be in [0, 1)")
        self.p = float(p)

    def forward(self, x):
        if not self.training or self.p == 0.0:
            return x
        return _DropoutFn.apply(x, self.p, _next_seed())


# ---------------------------------------------------------------------------------------------------------
# SyncBatchNorm
# ---------------------------------------------------------------------------------------------------------
class _SyncBatchNormFn(torch.autograd.Function):
    """Restates torch/nn/modules/_functions.py:7-209 on top of the b200ddp_bn_* kernels."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, momentum, comm, world, fuse_relu):
        L = _lib.load()
        x = _lib.f32c(x, "input")
        dev, st = _lib.dev_index(x), _lib.cur_stream(x.device)
        B, C = x.shape[0], x.shape[1]
        HW = x.numel() // (B * C)
        if B * HW == 1 and world < 2:
            raise ValueError("Expected more than 1 value per channel when training")
        scratch_bytes = int(L.b200ddp_bn_scratch_bytes(B, C))
        scratch = _lib.workspace(x.device, scratch_bytes, "bn", zero=True)
        stats = torch.empty(2 * C + 1, dtype=torch.float32, device=x.device)   # mean | invstd | count
        _lib.check(L.b200ddp_bn_stats(x.data_ptr(), B, C, HW, float(eps), stats.data_ptr(),
                                      stats.data_ptr() + 4 * C, scratch.data_ptr(), scratch.numel(), dev, st),
                   "bn_stats")
        stats[2 * C].fill_(float(B * HW))
        if world > 1:
            gathered = torch.empty(world * (2 * C + 1), dtype=torch.float32, device=x.device)
            if not comm.small_collective(stats, gathered, sum_mode=False):      # one kernel over peer memory, else NCCL
                comm.allgather(stats, gathered)
        else:
            gathered = stats
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        invstd = torch.empty(C, dtype=torch.float32, device=x.device)
        _lib.check(L.b200ddp_bn_gather_stats(gathered.data_ptr(), world, C, float(eps), float(momentum),
                                             _lib.ptr(running_mean), _lib.ptr(running_var), mean.data_ptr(),
                                             invstd.data_ptr(), dev, st), "bn_gather_stats")
        y = torch.empty_like(x)
        _lib.check(L.b200ddp_bn_elemt(x.data_ptr(), mean.data_ptr(), invstd.data_ptr(), _lib.ptr(weight),
                                      _lib.ptr(bias), y.data_ptr(), B, C, HW, int(fuse_relu), dev, st), "bn_elemt")
        ctx.save_for_backward(x, weight, mean, invstd, y if fuse_relu else None)
        ctx.comm, ctx.world, ctx.count_local = comm, world, B * HW
        ctx.weight_param, ctx.bias_param = weight, bias
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.load()
        x, weight, mean, invstd, y_relu = ctx.saved_tensors
        dy = _lib.f32c(dy, "grad_output")
        dev, st = _lib.dev_index(x), _lib.cur_stream(x.device)
        B, C = x.shape[0], x.shape[1]
        HW = x.numel() // (B * C)
        scratch_bytes = int(L.b200ddp_bn_scratch_bytes(B, C))
        scratch = _lib.workspace(x.device, scratch_bytes, "bn", zero=True)
        sums = torch.empty(2 * C, dtype=torch.float32, device=x.device)   # sum_dy | sum_dy_xmu
        gw = _grad_target(ctx.weight_param, (C,), x.device)[0] if weight is not None and ctx.needs_input_grad[1] else None
        gb = _grad_target(ctx.bias_param, (C,), x.device)[0] if ctx.bias_param is not None and ctx.needs_input_grad[2] else None
        _lib.check(L.b200ddp_bn_bwd_reduce(dy.data_ptr(), x.data_ptr(), _lib.ptr(y_relu), mean.data_ptr(),
                                           invstd.data_ptr(), B, C, HW, sums.data_ptr(), sums.data_ptr() + 4 * C,
                                           _lib.ptr(gw), _lib.ptr(gb), scratch.data_ptr(), scratch.numel(), dev, st),
                   "bn_bwd_reduce")
        total = ctx.count_local
        if ctx.world > 1:
            summed = torch.empty_like(sums)
            if ctx.comm.small_collective(sums, summed, sum_mode=True):
                sums = summed
            else:
                ctx.comm.allreduce_(sums, 1.0)
            total = ctx.count_local * ctx.world   # equal local batch sizes (DistributedSampler pads)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _lib.check(L.b200ddp_bn_bwd_elemt(dy.data_ptr(), x.data_ptr(), _lib.ptr(y_relu), mean.data_ptr(),
                                              invstd.data_ptr(), _lib.ptr(weight), sums.data_ptr(),
                                              sums.data_ptr() + 4 * C, 1.0 / float(total), dx.data_ptr(), B, C, HW,
                                              dev, st), "bn_bwd_elemt")
        return dx, gw, gb, None, None, None, None, None, None, None


class SyncBatchNorm(_BatchNorm):
    """Batch normalisation whose statistics span all ranks (drop-in for torch.nn.SyncBatchNorm).

    `fuse_relu=True` additionally folds the ReLU that follows the layer in AlexNet-BN into the
    normalisation kernel (and its mask into the backward kernels)."""

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: Optional[float] = 0.1, affine: bool = True,
                 track_running_stats: bool = True, process_group=None, fuse_relu: bool = False, device=None,
                 dtype=None) -> None:
        super().__init__(num_features, eps, momentum, affine, track_running_stats, device=device, dtype=dtype)
        self.process_group = process_group
        self.fuse_relu = fuse_relu

    def _check_input_dim(self, input) -> None:
        if input.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {input.dim()}D input)")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        self._check_input_dim(x)
        if not x.is_cuda:
            raise ValueError("SyncBatchNorm expected input tensor to be on GPU")   # as upstream batchnorm.py:800-809
        if self.momentum is None:
            eaf = 0.0
        else:
            eaf = self.momentum
        if self.training and self.track_running_stats:
            self.num_batches_tracked.add_(1)
            if self.momentum is None:
                eaf = 1.0 / float(self.num_batches_tracked.item())
        use_batch_stats = self.training or (self.running_mean is None and self.running_var is None)
        if not use_batch_stats:
            # eval: running statistics, no communication (batchnorm.py:791-828)
            L = _lib.load()
            xc = _lib.f32c(x, "input")
            B, C = xc.shape[0], xc.shape[1]
            HW = xc.numel() // (B * C)
            invstd = torch.rsqrt(self.running_var + self.eps)
            y = torch.empty_like(xc)
            _lib.check(L.b200ddp_bn_elemt(xc.data_ptr(), self.running_mean.data_ptr(), invstd.data_ptr(),
                                          _lib.ptr(self.weight), _lib.ptr(self.bias), y.data_ptr(), B, C, HW,
                                          int(self.fuse_relu), _lib.dev_index(xc), _lib.cur_stream(xc.device)),
                       "bn_elemt")
            return y
        world = 1
        comm = None
        if dist.is_available() and dist.is_initialized():
            world = dist.get_world_size(self.process_group)
            if world > 1:
                comm = get_comm(x.device, self.process_group)
        rm = self.running_mean if self.track_running_stats else None
        rv = self.running_var if self.track_running_stats else None
        return _SyncBatchNormFn.apply(x, self.weight, self.bias, rm, rv, self.eps, eaf, comm, world, self.fuse_relu)

    @classmethod
    def convert_sync_batchnorm(cls, module: nn.Module, process_group=None) -> nn.Module:
        """Recursively replace every BatchNorm*D child by SyncBatchNorm, sharing parameters and buffers
        (torch batchnorm.py:844-905), then fold each `SyncBatchNorm -> ReLU` pair of a Sequential into the
        normalisation kernels (models.fuse_bn_relu) so that no eager activation remains between them."""
        from .models import fuse_bn_relu
        return fuse_bn_relu(cls._convert(module, process_group))

    @classmethod
    def _convert(cls, module: nn.Module, process_group=None) -> nn.Module:
        out = module
        if isinstance(module, _BatchNorm) and not isinstance(module, SyncBatchNorm):
            out = cls(module.num_features, module.eps, module.momentum, module.affine, module.track_running_stats,
                      process_group)
            if module.affine:
                with torch.no_grad():
                    out.weight = module.weight
                    out.bias = module.bias
            out.running_mean = module.running_mean
            out.running_var = module.running_var
            out.num_batches_tracked = module.num_batches_tracked
            out.training = module.training
            if hasattr(module, "qconfig"):
                out.qconfig = module.qconfig
        for name, child in module.named_children():
            out.add_module(name, cls._convert(child, process_group))
        del module
        return out
