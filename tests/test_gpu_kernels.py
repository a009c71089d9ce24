"""Parity of every CUDA kernel, called THROUGH THE C ABI (ctypes on libb200ddp.so), against the CPU
oracle, the upstream golden vectors, and size-independent properties at full AlexNet sizes.

Tolerances (stated per the precision contract in DESIGN.md):
  integer / byte results (gates, pool indices, dropout masks, bucket plans) ... bit exact
  fp32 elementwise / reductions ................................................ rtol 2e-6 .. 1e-5
  strict-fp32 GEMM (mode 0) .................................................... 2e-5 of the output scale
  TF32 tensor-core GEMM (mode 1), fp32 accumulate .............................. 3e-3 of the output scale
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from b200ddp import _lib
from oracle import ddp_oracle as O

pytestmark = pytest.mark.gpu

DEV = torch.device("cuda", 0)


def L():
    return _lib.load()


def st():
    return torch.cuda.current_stream(DEV).cuda_stream


def call(name, *args):
    _lib.check(getattr(L(), name)(*args), name)


def dev(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(DEV)


def rel_err(got, want):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


def test_library_is_native_and_counts_launches():
    assert _lib.is_loaded() or L() is not None
    sm, maj, mnr = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    call("b200ddp_device_info", 0, ctypes.byref(sm), ctypes.byref(maj), ctypes.byref(mnr))
    assert maj.value == 10, f"these kernels are built for sm_100a only, device is sm_{maj.value}{mnr.value}"
    before = _lib.launch_count()
    x = torch.ones(1024, device=DEV)
    call("b200ddp_relu_fwd", x.data_ptr(), x.numel(), 0, st())
    assert _lib.launch_count() == before + 1


# ---- O1 Adam ---------------------------------------------------------------------------------------------------
def test_adam_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "adam.npz"))
    for wd in (0.0, 0.01):
        p, m, v = dev(g["p0"]), torch.zeros(1003, device=DEV), torch.zeros(1003, device=DEV)
        for t in (1, 2, 3):
            gr = dev(g[f"g{t}"])
            call("b200ddp_adam_step_flat", p.data_ptr(), gr.data_ptr(), m.data_ptr(), v.data_ptr(), 1003, 1e-3, 0.9,
                 0.999, 1e-8, wd, t, None, 0, st())
            tag = f"wd{wd}_t{t}"
            np.testing.assert_allclose(p.cpu().numpy(), g[f"p_{tag}"], rtol=2e-6, atol=2e-7)
            np.testing.assert_allclose(m.cpu().numpy(), g[f"m_{tag}"], rtol=2e-6, atol=1e-7)
            np.testing.assert_allclose(v.cpu().numpy(), g[f"v_{tag}"], rtol=2e-6, atol=1e-9)


def test_adam_full_size_against_torch_and_edge_sizes():
    n = 57_044_810 + 6                      # the arena of the toy model is a little larger than P
    torch.manual_seed(0)
    p = torch.randn(n, device=DEV)
    gr = torch.randn(n, device=DEV) * 0.01
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=1e-3)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for t in (1, 2):
        ref.grad = gr.clone()
        opt.step()
        call("b200ddp_adam_step_flat", p.data_ptr(), gr.data_ptr(), m.data_ptr(), v.data_ptr(), n, 1e-3, 0.9, 0.999,
             1e-8, 0.0, t, None, 0, st())
    assert torch.allclose(p, ref.detach(), rtol=2e-6, atol=1e-7)
    assert torch.allclose(m, opt.state[ref]["exp_avg"], rtol=2e-6, atol=1e-9)
    # zero gradient, zero state: parameters must not move (arena padding stays put)
    z = torch.zeros(64, device=DEV)
    q = torch.zeros(64, device=DEV)
    call("b200ddp_adam_step_flat", q.data_ptr(), z.data_ptr(), z.clone().data_ptr(), z.clone().data_ptr(), 64, 1e-3,
         0.9, 0.999, 1e-8, 0.0, 1, None, 0, st())
    assert float(q.abs().sum()) == 0.0
    # empty input is a no-op; bad arguments are reported, not executed
    call("b200ddp_adam_step_flat", None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, None, 0, st())
    rc = L().b200ddp_adam_step_flat(p.data_ptr() + 4, gr.data_ptr(), m.data_ptr(), v.data_ptr(), 8, 1e-3, 0.9, 0.999,
                                    1e-8, 0.0, 1, None, 0, st())
    assert rc == 1 and "aligned" in _lib.last_error()
    # device-resident gradient scale (used for fused averaging)
    p2, m2, v2 = torch.ones(1000, device=DEV), torch.zeros(1000, device=DEV), torch.zeros(1000, device=DEV)
    g2 = torch.full((1000,), 4.0, device=DEV)
    sc = torch.tensor([0.25], device=DEV)
    call("b200ddp_adam_step_flat", p2.data_ptr(), g2.data_ptr(), m2.data_ptr(), v2.data_ptr(), 1000, 1e-3, 0.9, 0.999,
         1e-8, 0.0, 1, sc.data_ptr(), 0, st())
    assert torch.allclose(m2, torch.full_like(m2, 0.1))


# ---- R2 / R3 ---------------------------------------------------------------------------------------------------
def test_pack_scale_clip_against_oracle():
    rng = np.random.default_rng(0)
    for n in (1, 3, 4, 1001, 1 << 20):
        g = (rng.standard_normal(n) * 2).astype(np.float32)
        if n > 10:
            g[5] = np.nan
        src, dst = dev(g), torch.empty(n, device=DEV)
        call("b200ddp_pack_scale_clip", dst.data_ptr(), src.data_ptr(), n, 0.125, 0.5, 0, 0, st())
        want = O.clip_grad_value(g, 0.5) * np.float32(0.125)
        np.testing.assert_array_equal(dst.cpu().numpy(), want)
        call("b200ddp_pack_scale_clip", dst.data_ptr(), src.data_ptr(), n, 1.0, 0.0, 0, 0, st())
        np.testing.assert_array_equal(dst.cpu().numpy(), g)                  # pure copy
        call("b200ddp_pack_scale_clip", src.data_ptr(), src.data_ptr(), n, 1.0, 0.5, 1, 0, st())   # in place, NaN -> 0
        assert torch.isfinite(src).all() and float(src.abs().max()) <= 0.5
    # unaligned source (view at an odd offset)
    base = torch.randn(1030, device=DEV)
    out = torch.empty(1024, device=DEV)
    call("b200ddp_pack_scale_clip", out.data_ptr(), base[3:].data_ptr(), 1024, 2.0, 0.0, 0, 0, st())
    assert torch.equal(out, base[3:1027] * 2.0)


def test_norm_clip_chain_against_upstream_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "clip.npz"))
    flat = np.concatenate([g[f"g{i}"].reshape(-1) for i in range(4)])
    scratch = torch.zeros(int(L().b200ddp_sqnorm_scratch_bytes()), dtype=torch.uint8, device=DEV)
    for mn in (1.0, 100.0):
        buf = torch.zeros(((flat.size + 63) // 64) * 64, device=DEV)
        buf[: flat.size] = dev(flat)
        sq, coef, norm = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
        half = 64
        call("b200ddp_sqnorm", buf.data_ptr(), half, sq.data_ptr(), 0, scratch.data_ptr(), scratch.numel(), 0, st())
        call("b200ddp_sqnorm", buf[half:].data_ptr(), buf.numel() - half, sq.data_ptr(), 1, scratch.data_ptr(),
             scratch.numel(), 0, st())
        call("b200ddp_clip_coef", sq.data_ptr(), mn, 0, coef.data_ptr(), norm.data_ptr(), 0, st())
        before = buf.clone()
        call("b200ddp_scale_by_device_scalar", buf.data_ptr(), buf.numel(), coef.data_ptr(), 0, 0, st())
        assert abs(float(norm) - float(g[f"norm_total_{mn}"])) < 2e-6 * float(norm)
        want = np.concatenate([g[f"norm_{mn}_g{i}"].reshape(-1) for i in range(4)])
        np.testing.assert_allclose(buf[: flat.size].cpu().numpy(), want, rtol=2e-6, atol=1e-7)
        if mn == 100.0:
            assert torch.equal(buf, before)                   # coef == 1: untouched, bit for bit


def test_sqnorm_large_and_nonfinite_policies():
    n = 57_044_864
    x = torch.randn(n, device=DEV)
    scratch = torch.zeros(int(L().b200ddp_sqnorm_scratch_bytes()), dtype=torch.uint8, device=DEV)
    sq = torch.zeros(1, device=DEV)
    for _ in range(2):          # twice: the ticket must reset itself
        call("b200ddp_sqnorm", x.data_ptr(), n, sq.data_ptr(), 0, scratch.data_ptr(), scratch.numel(), 0, st())
        want = float((x.double() ** 2).sum())
        assert abs(float(sq) - want) < 2e-6 * want
    coef, norm = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    bad = torch.tensor([float("nan")], device=DEV)
    call("b200ddp_clip_coef", bad.data_ptr(), 1.0, 0, coef.data_ptr(), norm.data_ptr(), 0, st())
    assert torch.isnan(coef).all()                            # stock: NaN propagates
    call("b200ddp_clip_coef", bad.data_ptr(), 1.0, 1, coef.data_ptr(), norm.data_ptr(), 0, st())
    assert float(coef) == 0.0
    y = torch.tensor([1.0, float("nan"), float("inf"), -2.0] * 4, device=DEV)
    call("b200ddp_scale_by_device_scalar", y.data_ptr(), y.numel(), coef.data_ptr(), 1, 0, st())
    assert float(y.abs().sum()) == 0.0                        # exact zeros, not NaN * 0


# ---- F1 / F2 / F3 linear ------------------------------------------------------------------------------------
TOL = {0: 2e-5, 1: 3e-3}


def _linear_all(x, w, b, dy, relu, p, seed, mode):
    B, K = x.shape
    N = w.shape[0]
    xd, wd, bd, dyd = dev(x), dev(w), (dev(b) if b is not None else None), dev(dy)
    y = torch.empty(B, N, device=DEV)
    gate = torch.empty(B, N, dtype=torch.uint8, device=DEV) if (relu or p > 0) else None
    ws = torch.empty(int(L().b200ddp_linear_workspace_bytes(B, N, K)), dtype=torch.uint8, device=DEV)
    call("b200ddp_linear_fwd", xd.data_ptr(), wd.data_ptr(), _lib.ptr(bd), y.data_ptr(), _lib.ptr(gate), B, N, K,
         int(relu), p, seed, mode, ws.data_ptr(), ws.numel(), 0, st())
    dz = torch.empty_like(dyd)
    db = torch.empty(N, device=DEV)
    if gate is not None:
        call("b200ddp_gate_bias_bwd", dyd.data_ptr(), gate.data_ptr(), p, dz.data_ptr(), db.data_ptr(), B, N, 0, st())
    else:
        dz = dyd
    dx = torch.empty(B, K, device=DEV)
    dw = torch.empty(N, K, device=DEV)
    call("b200ddp_linear_bwd_dx", dz.data_ptr(), wd.data_ptr(), dx.data_ptr(), B, N, K, mode, ws.data_ptr(), ws.numel(),
         0, st())
    call("b200ddp_linear_bwd_dw", dz.data_ptr(), xd.data_ptr(), dw.data_ptr(), None if gate is not None else db.data_ptr(),
         B, N, K, mode, 0, st())
    torch.cuda.synchronize()
    return (y.cpu().numpy(), None if gate is None else gate.cpu().numpy(), dx.cpu().numpy(), dw.cpu().numpy(),
            db.cpu().numpy())


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "tf32"])
def test_linear_golden(golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    y, gate, dx, dw, db = _linear_all(g["lin_x"], g["lin_w"], g["lin_b"], g["lin_dy"], True, 0.0, 0, mode)
    assert rel_err(y, g["lin_y"]) < TOL[mode]
    assert rel_err(dx, g["lin_dx"]) < TOL[mode] * 4
    assert rel_err(dw, g["lin_dw"]) < TOL[mode] * 4
    assert rel_err(db, g["lin_db"]) < 1e-5


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "tf32"])
@pytest.mark.parametrize("B,N,K,relu,p", [(128, 256, 512, True, 0.5), (106, 200, 300, True, 0.0), (5, 64, 36, False, 0.0),
                                          (128, 4096, 1024, True, 0.5), (257, 132, 4100, False, 0.25)])
def test_linear_against_oracle_with_tails(mode, B, N, K, relu, p):
    rng = np.random.default_rng(B * 7 + N)
    x = rng.standard_normal((B, K)).astype(np.float32)
    w = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    dy = rng.standard_normal((B, N)).astype(np.float32)
    seed = 0x1234567 + B
    y, gate, dx, dw, db = _linear_all(x, w, b, dy, relu, p, seed, mode)
    keep = O.dropout_keep_mask(B * N, p, seed).reshape(B, N) if p > 0 else None
    y_ref, gate_ref = O.linear_fwd(x, w, b, relu=relu, keep=keep, p=p)
    z = x.astype(np.float64) @ w.astype(np.float64).T + b
    scale = np.abs(z).max()
    sure = np.abs(z) > 50 * TOL[mode] * scale          # away from the ReLU kink the gate is unambiguous
    if gate is not None:
        assert (gate[sure] == gate_ref[sure]).all()
        if p > 0:                                       # the dropout decision itself is integer: bit exact everywhere
            assert ((gate == 0) | (gate == 2) | ~keep).all() and ((gate != 2) | keep).all()
    assert np.abs(y - y_ref)[sure].max() < TOL[mode] * scale * (2 if p > 0 else 1)
    g_use = gate if gate is not None else np.ones((B, N), dtype=np.uint8)
    dx_ref, dw_ref, db_ref = O.linear_bwd(dy, g_use, x, w, p)
    assert rel_err(dx, dx_ref) < TOL[mode] * 4
    assert rel_err(dw, dw_ref) < TOL[mode] * 4
    assert rel_err(db, db_ref) < 1e-5


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "tf32"])
def test_linear_fc1_full_size_against_fp64(mode):
    B, N, K = 128, 4096, 9216
    torch.manual_seed(1)
    x = torch.randn(B, K, device=DEV)
    w = torch.randn(N, K, device=DEV) / 96.0
    b = torch.randn(N, device=DEV)
    dz = torch.randn(B, N, device=DEV)
    y = torch.empty(B, N, device=DEV)
    ws = torch.empty(int(L().b200ddp_linear_workspace_bytes(B, N, K)), dtype=torch.uint8, device=DEV)
    call("b200ddp_linear_fwd", x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, B, N, K, 0, 0.0, 0, mode,
         ws.data_ptr(), ws.numel(), 0, st())
    dx = torch.empty(B, K, device=DEV)
    dw = torch.empty(N, K, device=DEV)
    db = torch.empty(N, device=DEV)
    call("b200ddp_linear_bwd_dx", dz.data_ptr(), w.data_ptr(), dx.data_ptr(), B, N, K, mode, ws.data_ptr(), ws.numel(),
         0, st())
    call("b200ddp_linear_bwd_dw", dz.data_ptr(), x.data_ptr(), dw.data_ptr(), db.data_ptr(), B, N, K, mode, 0, st())
    y64 = x.double() @ w.double().T + b.double()
    assert float((y.double() - y64).abs().max() / y64.abs().max()) < TOL[mode]
    dx64 = dz.double() @ w.double()
    assert float((dx.double() - dx64).abs().max() / dx64.abs().max()) < TOL[mode] * 2
    dw64 = dz.double().T @ x.double()
    assert float((dw.double() - dw64).abs().max() / dw64.abs().max()) < TOL[mode] * 2
    assert torch.allclose(db.double(), dz.double().sum(0), rtol=1e-5, atol=1e-5)
    # linearity (size-independent property): f(2x) == 2 f(x) exactly for power-of-two scaling without bias
    y1, y2 = torch.empty(B, N, device=DEV), torch.empty(B, N, device=DEV)
    x2 = x * 2
    call("b200ddp_linear_fwd", x.data_ptr(), w.data_ptr(), None, y1.data_ptr(), None, B, N, K, 0, 0.0, 0, mode,
         ws.data_ptr(), ws.numel(), 0, st())
    call("b200ddp_linear_fwd", x2.data_ptr(), w.data_ptr(), None, y2.data_ptr(), None, B, N, K, 0, 0.0, 0, mode,
         ws.data_ptr(), ws.numel(), 0, st())
    assert torch.equal(y2, y1 * 2)


@pytest.mark.parametrize("stages,epi", [(6, 8), (3, 8), (3, 16), (6, 16)])
@pytest.mark.parametrize("B,N,K", [(128, 4096, 9216), (128, 4096, 4096), (100, 2100, 2500), (13, 4096, 4096)])
def test_linear_dw_persistent_kernel_is_bit_identical_to_the_tile_kernel(monkeypatch, stages, epi, B, N, K):
    """Same k-blocks in the same order into the same accumulator layout: the persistent CTAs (two TMEM accumulators,
    ring running across tile boundaries) must reproduce the one-tile-per-CTA kernel bit for bit, tails included."""
    torch.manual_seed(B + N)
    x = torch.randn(B, K, device=DEV)
    dz = torch.randn(B, N, device=DEV)
    out = {}
    for tag, env in (("tile", "0"), ("persistent", "1")):
        monkeypatch.setenv("B200DDP_GEMM_PERSISTENT", env)
        monkeypatch.setenv("B200DDP_GEMM_PERS_STAGES", str(stages))
        monkeypatch.setenv("B200DDP_GEMM_PERS_EPI", str(epi))
        dw = torch.full((N, K), float("nan"), device=DEV)
        for _ in range(2):                                   # twice: barrier phases and accumulators start clean again
            call("b200ddp_linear_bwd_dw", dz.data_ptr(), x.data_ptr(), dw.data_ptr(), None, B, N, K, 1, 0, st())
        torch.cuda.synchronize()
        out[tag] = dw
    assert torch.equal(out["tile"], out["persistent"])
    dw64 = dz.double().T @ x.double()
    assert float((out["persistent"].double() - dw64).abs().max() / dw64.abs().max()) < TOL[1] * 2


def test_head_small_n_against_oracle():
    rng = np.random.default_rng(3)
    for B, K in ((128, 4096), (106, 4096), (7, 33)):
        x = rng.standard_normal((B, K)).astype(np.float32)
        w = (rng.standard_normal((10, K)) / 64).astype(np.float32)
        b = rng.standard_normal(10).astype(np.float32)
        dy = rng.standard_normal((B, 10)).astype(np.float32)
        y, gate, dx, dw, db = _linear_all(x, w, b, dy, False, 0.0, 0, 0)
        y_ref, _ = O.linear_fwd(x, w, b)
        dx_ref, dw_ref, db_ref = O.linear_bwd(dy, np.ones((B, 10), np.uint8), x, w)
        assert rel_err(y, y_ref) < 1e-5 and rel_err(dx, dx_ref) < 1e-5
        assert rel_err(dw, dw_ref) < 1e-5 and rel_err(db, db_ref) < 1e-5


# ---- loss -----------------------------------------------------------------------------------------------------
def test_softmax_ce_golden_and_sizes(golden_dir):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    cases = [(g["ce_z"], g["ce_labels"])]
    rng = np.random.default_rng(5)
    for B in (1, 106, 128, 1024, 3000):
        cases.append(((rng.standard_normal((B, 10)) * 5).astype(np.float32), rng.integers(0, 10, B)))
    for z, lab in cases:
        zd, ld = dev(z), torch.as_tensor(lab, dtype=torch.int64).to(DEV)
        loss, dz = torch.zeros((), device=DEV), torch.empty_like(zd)
        call("b200ddp_softmax_ce", zd.data_ptr(), ld.data_ptr(), loss.data_ptr(), dz.data_ptr(), z.shape[0], 10, None, 0,
             st())
        l_ref, d_ref = O.softmax_ce(z, np.asarray(lab))
        assert abs(float(loss) - float(l_ref)) < 2e-6 * max(1.0, abs(float(l_ref)))
        np.testing.assert_allclose(dz.cpu().numpy(), d_ref, rtol=1e-5, atol=1e-8)
    assert abs(float(loss) - float(torch.nn.functional.cross_entropy(zd, ld))) < 1e-5
    flag = torch.zeros(1, dtype=torch.int32, device=DEV)
    bad = torch.tensor([3, 11], dtype=torch.int64, device=DEV)
    call("b200ddp_softmax_ce", zd.data_ptr(), bad.data_ptr(), loss.data_ptr(), None, 2, 10, flag.data_ptr(), 0, st())
    assert int(flag) == 1


# ---- pool / relu / dropout --------------------------------------------------------------------------------------
def test_relu_maxpool_golden_and_alexnet_shapes(golden_dir):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    shapes = [None, (4, 64, 55, 55), (4, 192, 27, 27), (3, 256, 13, 13), (1, 1, 3, 3)]
    rng = np.random.default_rng(1)
    for shp in shapes:
        x = g["pool_x"] if shp is None else rng.standard_normal(shp).astype(np.float32)
        B, C, H, W = x.shape
        OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
        xd = dev(x)
        y = torch.empty(B, C, OH, OW, device=DEV)
        idx = torch.empty(B, C, OH, OW, dtype=torch.uint8, device=DEV)
        call("b200ddp_relu_maxpool_fwd", xd.data_ptr(), y.data_ptr(), idx.data_ptr(), B, C, H, W, 1, 0, st())
        y_ref, idx_ref = O.relu_maxpool_fwd(x)
        np.testing.assert_array_equal(y.cpu().numpy(), y_ref)            # bit exact
        np.testing.assert_array_equal(idx.cpu().numpy(), idx_ref)
        dy = g["pool_dy"] if shp is None else rng.standard_normal((B, C, OH, OW)).astype(np.float32)
        dx = torch.empty(B, C, H, W, device=DEV)
        call("b200ddp_relu_maxpool_bwd", dev(dy).data_ptr(), idx.data_ptr(), dx.data_ptr(), B, C, H, W, 0, st())
        np.testing.assert_allclose(dx.cpu().numpy(), O.relu_maxpool_bwd(dy, idx_ref, x.shape), rtol=0, atol=1e-6)
        if shp is None:
            np.testing.assert_array_equal(y.cpu().numpy(), g["pool_y"])
            np.testing.assert_allclose(dx.cpu().numpy(), g["pool_dx"], rtol=0, atol=1e-6)
    # against torch's own kernels at the first pool's full batch-128 size
    xt = torch.randn(128, 64, 55, 55, device=DEV, requires_grad=True)
    yt = torch.nn.functional.max_pool2d(torch.relu(xt), 3, 2)
    dyt = torch.randn_like(yt)
    yt.backward(dyt)
    y = torch.empty_like(yt)
    idx = torch.empty(yt.shape, dtype=torch.uint8, device=DEV)
    call("b200ddp_relu_maxpool_fwd", xt.data_ptr(), y.data_ptr(), idx.data_ptr(), 128, 64, 55, 55, 1, 0, st())
    dx = torch.empty_like(xt)
    call("b200ddp_relu_maxpool_bwd", dyt.data_ptr(), idx.data_ptr(), dx.data_ptr(), 128, 64, 55, 55, 0, st())
    assert torch.equal(y, yt.detach()) and torch.allclose(dx, xt.grad, rtol=0, atol=1e-6)


@pytest.mark.parametrize("shape", [(128, 64, 55, 55), (128, 192, 27, 27), (128, 256, 13, 13), (3, 5, 9, 7)])
def test_pool_and_relu_backward_with_fused_bias_gradient(shape):
    """relu_maxpool_bwd_bias / relu_bwd_bias: dx bit-identical to the plain backward kernels, db = dx.sum((0,2,3))
    (fp64 fold), twice in a row on the same scratch (the ticket must be left at zero)."""
    B, C, H, W = shape
    torch.manual_seed(B * C + H)
    x = torch.randn(B, C, H, W, device=DEV)
    OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
    y = torch.empty(B, C, OH, OW, device=DEV)
    idx = torch.empty(B, C, OH, OW, dtype=torch.uint8, device=DEV)
    call("b200ddp_relu_maxpool_fwd", x.data_ptr(), y.data_ptr(), idx.data_ptr(), B, C, H, W, 1, 0, st())
    nb = int(L().b200ddp_bias_grad_scratch_bytes(B, C))
    scratch = torch.zeros(nb, dtype=torch.uint8, device=DEV)
    for rep in range(2):
        dy = torch.randn(B, C, OH, OW, device=DEV)
        ref = torch.empty_like(x)
        call("b200ddp_relu_maxpool_bwd", dy.data_ptr(), idx.data_ptr(), ref.data_ptr(), B, C, H, W, 0, st())
        dx = torch.empty_like(x)
        db = torch.empty(C, device=DEV)
        call("b200ddp_relu_maxpool_bwd_bias", dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), db.data_ptr(), B, C, H, W,
             scratch.data_ptr(), scratch.numel(), 0, st())
        assert torch.equal(dx, ref)
        want = ref.double().sum((0, 2, 3))
        assert float((db.double() - want).abs().max()) <= 1e-5 * float(want.abs().max()) + 1e-6
        # ReLU-only variant on the input-sized tensor
        g = torch.randn_like(x)
        yr = torch.relu(x)
        dx2 = torch.empty_like(x)
        db2 = torch.empty(C, device=DEV)
        call("b200ddp_relu_bwd_bias", g.data_ptr(), yr.data_ptr(), dx2.data_ptr(), db2.data_ptr(), B, C, H * W,
             scratch.data_ptr(), scratch.numel(), 0, st())
        want2 = g * (x > 0)
        assert torch.equal(dx2, want2)
        w2 = want2.double().sum((0, 2, 3))
        assert float((db2.double() - w2).abs().max()) <= 1e-5 * float(w2.abs().max()) + 1e-6
    # against autograd through relu -> max_pool2d
    xt = x.clone().requires_grad_()
    yt = torch.nn.functional.max_pool2d(torch.relu(xt), 3, 2)
    yt.backward(dy)
    assert torch.allclose(dx, xt.grad, rtol=0, atol=1e-6)


def test_relu_and_dropout_kernels():
    x = torch.randn(1_000_003, device=DEV)
    y = x.clone()
    call("b200ddp_relu_fwd", y.data_ptr(), y.numel(), 0, st())
    assert torch.equal(y, torch.relu(x))
    dy = torch.randn_like(x)
    dx = torch.empty_like(x)
    call("b200ddp_relu_bwd", dy.data_ptr(), y.data_ptr(), dx.data_ptr(), x.numel(), 0, st())
    assert torch.equal(dx, dy * (x > 0))
    n, p, seed = 128 * 9216, 0.5, 0xABCDEF0123
    xs = torch.randn(n, device=DEV)
    out = torch.empty_like(xs)
    gate = torch.empty(n, dtype=torch.uint8, device=DEV)
    call("b200ddp_dropout_fwd", xs.data_ptr(), out.data_ptr(), gate.data_ptr(), n, p, seed, 0, st())
    keep = O.dropout_keep_mask(n, p, seed)
    np.testing.assert_array_equal(gate.cpu().numpy() == 2, keep)        # Philox mask bit exact vs the oracle
    assert torch.equal(out, torch.where(torch.as_tensor(keep).to(DEV), xs * 2.0, torch.zeros_like(xs)))
    dxs = torch.empty_like(xs)
    dys = torch.randn_like(xs)
    call("b200ddp_gate_bias_bwd", dys.data_ptr(), gate.data_ptr(), p, dxs.data_ptr(), None, 128, 9216, 0, st())
    assert torch.equal(dxs, torch.where(torch.as_tensor(keep).to(DEV), dys * 2.0, torch.zeros_like(xs)))


# ---- B1 SyncBatchNorm -----------------------------------------------------------------------------------------
def _bn_two_ranks(xs, w, b, dys, relu):
    """Both 'ranks' on one GPU: the collectives are replaced by concatenation / addition of their payloads."""
    C = xs[0].shape[1]
    world = len(xs)
    stats = []
    xd = [dev(x) for x in xs]
    for x in xd:
        B, HW = x.shape[0], x.shape[2] * x.shape[3]
        scratch = torch.zeros(int(L().b200ddp_bn_scratch_bytes(B, C)), dtype=torch.uint8, device=DEV)
        s = torch.empty(2 * C + 1, device=DEV)
        call("b200ddp_bn_stats", x.data_ptr(), B, C, HW, 1e-5, s.data_ptr(), s.data_ptr() + 4 * C, scratch.data_ptr(),
             scratch.numel(), 0, st())
        s[2 * C] = B * HW
        stats.append(s)
    gathered = torch.cat(stats)
    mean, invstd = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    call("b200ddp_bn_gather_stats", gathered.data_ptr(), world, C, 1e-5, 0.1, rm.data_ptr(), rv.data_ptr(),
         mean.data_ptr(), invstd.data_ptr(), 0, st())
    wd, bd = dev(w), dev(b)
    ys, sums, gws, gbs = [], [], [], []
    for x, dy in zip(xd, dys):
        B, HW = x.shape[0], x.shape[2] * x.shape[3]
        y = torch.empty_like(x)
        call("b200ddp_bn_elemt", x.data_ptr(), mean.data_ptr(), invstd.data_ptr(), wd.data_ptr(), bd.data_ptr(),
             y.data_ptr(), B, C, HW, int(relu), 0, st())
        ys.append(y)
        scratch = torch.zeros(int(L().b200ddp_bn_scratch_bytes(B, C)), dtype=torch.uint8, device=DEV)
        s, gw, gb = torch.empty(2 * C, device=DEV), torch.empty(C, device=DEV), torch.empty(C, device=DEV)
        call("b200ddp_bn_bwd_reduce", dev(dy).data_ptr(), x.data_ptr(), y.data_ptr() if relu else None, mean.data_ptr(),
             invstd.data_ptr(), B, C, HW, s.data_ptr(), s.data_ptr() + 4 * C, gw.data_ptr(), gb.data_ptr(),
             scratch.data_ptr(), scratch.numel(), 0, st())
        sums.append(s); gws.append(gw); gbs.append(gb)
    tot = sum(sums)
    count = sum(x.shape[0] * x.shape[2] * x.shape[3] for x in xs)
    dxs = []
    for x, dy, y in zip(xd, dys, ys):
        B, HW = x.shape[0], x.shape[2] * x.shape[3]
        dx = torch.empty_like(x)
        call("b200ddp_bn_bwd_elemt", dev(dy).data_ptr(), x.data_ptr(), y.data_ptr() if relu else None, mean.data_ptr(),
             invstd.data_ptr(), wd.data_ptr(), tot.data_ptr(), tot.data_ptr() + 4 * C, 1.0 / count, dx.data_ptr(), B, C,
             HW, 0, st())
        dxs.append(dx)
    return ys, rm, rv, dxs, sum(gws), sum(gbs)


def test_syncbn_golden_global_batch(golden_dir):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    ys, rm, rv, dxs, gw, gb = _bn_two_ranks([g["bn_x0"], g["bn_x1"]], g["bn_w"], g["bn_b"],
                                            [g["bn_dy"][:3], g["bn_dy"][3:]], relu=True)
    np.testing.assert_allclose(torch.cat(ys).cpu().numpy(), g["bn_y"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(rm.cpu().numpy(), g["bn_rm"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(rv.cpu().numpy(), g["bn_rv"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(torch.cat(dxs).cpu().numpy(), g["bn_dx"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(gw.cpu().numpy(), g["bn_gw"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(gb.cpu().numpy(), g["bn_gb"], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("C,HW", [(64, 55), (192, 27), (384, 13)])
def test_syncbn_alexnet_shapes_against_torch_batchnorm(C, HW):
    torch.manual_seed(C)
    xs = [(torch.randn(16, C, HW, HW) * 3 + 2).numpy() for _ in range(2)]
    w, b = torch.randn(C).numpy(), torch.randn(C).numpy()
    dys = [torch.randn(16, C, HW, HW).numpy() for _ in range(2)]
    ys, rm, rv, dxs, gw, gb = _bn_two_ranks(xs, w, b, dys, relu=False)
    xa = torch.cat([torch.as_tensor(x) for x in xs]).to(DEV).requires_grad_()
    wt, bt = dev(w).requires_grad_(), dev(b).requires_grad_()
    rmt, rvt = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    ya = torch.nn.functional.batch_norm(xa, rmt, rvt, wt, bt, training=True, momentum=0.1, eps=1e-5)
    ya.backward(torch.cat([torch.as_tensor(d) for d in dys]).to(DEV))
    assert torch.allclose(torch.cat(ys), ya.detach(), rtol=1e-4, atol=1e-4)
    assert torch.allclose(rm, rmt, rtol=1e-5, atol=1e-6) and torch.allclose(rv, rvt, rtol=1e-5, atol=1e-6)
    assert torch.allclose(torch.cat(dxs), xa.grad, rtol=1e-3, atol=1e-5)
    assert torch.allclose(gw, wt.grad, rtol=1e-4, atol=1e-2) and torch.allclose(gb, bt.grad, rtol=1e-4, atol=1e-2)


# ---- R4 communicator on a single rank -----------------------------------------------------------------------------
def test_native_comm_single_rank_roundtrip():
    from b200ddp import _build
    path = _build.nccl_library_path()
    call("b200ddp_nccl_load", path.encode() if path else None)
    ver = ctypes.c_int()
    call("b200ddp_nccl_version", ctypes.byref(ver))
    assert ver.value >= 21800
    uid = ctypes.create_string_buffer(128)
    call("b200ddp_comm_unique_id", uid, 128)
    h = ctypes.c_void_p()
    call("b200ddp_comm_create", ctypes.byref(h), uid, 0, 1, 0)
    try:
        x = torch.arange(1 << 20, dtype=torch.float32, device=DEV)
        want = x * 0.5
        call("b200ddp_bucket_reduce_launch", h, x.data_ptr(), x.numel(), 0.5, st())   # PreMulSum rides the collective
        call("b200ddp_bucket_reduce_join", h, st())
        torch.cuda.synchronize()
        assert torch.equal(x, want)
        call("b200ddp_comm_allreduce", h, x.data_ptr(), x.numel(), 1.0, st())
        assert torch.equal(x, want)
        y = torch.empty_like(x)
        call("b200ddp_comm_allgather", h, x.data_ptr(), y.data_ptr(), x.numel(), st())
        call("b200ddp_comm_broadcast", h, y.data_ptr(), y.numel() * 4, 0, st())
        torch.cuda.synchronize()
        assert torch.equal(y, want)
    finally:
        call("b200ddp_comm_destroy", h)
