"""C1-C5: the implicit-GEMM tcgen05 convolution kernels through the C ABI against float64 torch convolutions
(the published definition of nn.Conv2d), at the toy model's five layer shapes, with ragged pixel counts,
bias + ReLU epilogues, dgrad (flipped-weight forward) and split-K wgrad.

Tolerance: TF32 operands (fp32 truncated to 10 mantissa bits), fp32 accumulation -> 3e-3 of the output scale,
the same bound as the TF32 Linear kernels and the numeric class the reference itself uses for convolutions
(cuDNN TF32)."""
import pytest
import torch

import b200ddp
from b200ddp import _lib
from b200ddp import nn as bnn

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
TOL = 3e-3

#           Cin  H   Cout K  s  p
LAYERS = [(3, 224, 64, 11, 4, 2), (64, 27, 192, 5, 1, 2), (192, 13, 384, 3, 1, 1), (384, 13, 256, 3, 1, 1),
          (256, 13, 256, 3, 1, 1)]


def _err(got, want):
    return float((got.double() - want).abs().max() / want.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("layer", range(5))
@pytest.mark.parametrize("batch", [4, 5])
def test_conv_forward_matches_float64(layer, batch):
    Cin, H, Cout, K, s, p = LAYERS[layer]
    torch.manual_seed(layer * 10 + batch)
    x = torch.randn(batch, Cin, H, H, device=DEV)
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p, relu=(layer in (2, 3))).to(DEV)
    y = conv(x)
    ref = torch.nn.functional.conv2d(x.double(), conv.weight.double(), conv.bias.double(), stride=s, padding=p)
    if conv.relu:
        ref = torch.relu(ref)
    assert y.shape == ref.shape
    assert _err(y, ref) < TOL


@pytest.mark.parametrize("layer", range(5))
def test_conv_backward_matches_float64(layer):
    Cin, H, Cout, K, s, p = LAYERS[layer]
    batch = 6
    torch.manual_seed(100 + layer)
    x = torch.randn(batch, Cin, H, H, device=DEV, requires_grad=(layer > 0))
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p).to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    x64 = x.detach().double().requires_grad_(layer > 0)
    w64 = conv.weight.detach().double().requires_grad_()
    b64 = conv.bias.detach().double().requires_grad_()
    torch.nn.functional.conv2d(x64, w64, b64, stride=s, padding=p).backward(dy.double())
    assert _err(conv.weight.grad, w64.grad) < TOL
    assert _err(conv.bias.grad, b64.grad) < 1e-5
    if layer > 0:
        assert _err(x.grad, x64.grad) < TOL


def test_conv_full_batch_first_layer_and_relu_backward():
    """Batch 128 at conv1 (3025 CTAs, 49-way split wgrad) and a fused-ReLU layer's backward mask."""
    torch.manual_seed(7)
    x = torch.randn(128, 3, 224, 224, device=DEV)
    conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    ref = torch.nn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV).double()
    ref.load_state_dict({k: v.double() for k, v in conv.state_dict().items()})
    y64 = ref(x.double())
    y64.backward(dy.double())
    assert _err(y, y64.detach()) < TOL
    assert _err(conv.weight.grad, ref.weight.grad) < TOL
    assert _err(conv.bias.grad, ref.bias.grad) < 1e-5
    xr = torch.randn(8, 192, 13, 13, device=DEV, requires_grad=True)
    c3 = bnn.Conv2d(192, 384, 3, padding=1, relu=True).to(DEV)
    out = c3(xr)
    g = torch.randn_like(out)
    out.backward(g)
    x64 = xr.detach().double().requires_grad_()
    o64 = torch.relu(torch.nn.functional.conv2d(x64, c3.weight.double(), c3.bias.double(), padding=1))
    o64.backward(g.double())
    sure = (o64.detach().abs() > 1e-2)          # away from the ReLU kink
    assert float(((out > 0) != (o64 > 0))[sure].sum()) == 0
    rel_l2 = float((xr.grad.double() - x64.grad).norm() / x64.grad.norm())
    assert rel_l2 < 2e-2                          # a few ReLU gates flip near zero (TF32 forward)


@pytest.mark.parametrize("layer", [0, 1, 4])
def test_conv_relu_pool_fused_module_matches_float64(layer):
    """Conv2d(pool=True) = conv -> ReLU -> MaxPool2d(3,2) in one autograd node (the three pooled layers of the toy
    model); the bias gradient comes out of the pool-backward kernel."""
    Cin, H, Cout, K, s, p = LAYERS[layer]
    batch = 6
    torch.manual_seed(300 + layer)
    x = torch.randn(batch, Cin, H, H, device=DEV, requires_grad=(layer > 0))
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p, pool=True).to(DEV)
    out = conv(x)
    x64 = x.detach().double().requires_grad_(layer > 0)
    w64 = conv.weight.detach().double().requires_grad_()
    b64 = conv.bias.detach().double().requires_grad_()
    z64 = torch.nn.functional.conv2d(x64, w64, b64, stride=s, padding=p)
    o64 = torch.nn.functional.max_pool2d(torch.relu(z64), 3, 2)
    assert out.shape == o64.shape
    assert _err(out, o64.detach()) < TOL
    g = torch.randn_like(out)
    out.backward(g)
    o64.backward(g.double())

    def rel_l2(a, b):
        return float((a.double() - b).norm() / b.norm())
    # (1) end to end against the float64 network: a TF32 forward picks a different window maximum / ReLU gate where
    # two candidates are within rounding, so gradients agree in relative L2 only (same bound as the step test)
    assert rel_l2(conv.weight.grad, w64.grad) < 5e-2
    assert rel_l2(conv.bias.grad, b64.grad) < 5e-2
    if layer > 0:
        assert rel_l2(x.grad, x64.grad) < 5e-2
    # (2) exactness of the backward kernels given OUR forward's decisions: take the pool / ReLU routing from our own
    # conv output, do everything else in float64
    plain = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p).to(DEV)
    plain.load_state_dict(conv.state_dict())
    with torch.no_grad():
        z_ours = plain(x.detach())
    zl = z_ours.double().requires_grad_()
    torch.nn.functional.max_pool2d(torch.relu(zl), 3, 2).backward(g.double())
    dz = zl.grad                                   # gradient of the conv output under our routing
    x2 = x.detach().double().requires_grad_(layer > 0)
    w2 = conv.weight.detach().double().requires_grad_()
    b2 = conv.bias.detach().double().requires_grad_()
    torch.nn.functional.conv2d(x2, w2, b2, stride=s, padding=p).backward(dz)
    assert _err(conv.weight.grad, w2.grad) < TOL
    assert _err(conv.bias.grad, b2.grad) < 1e-5
    if layer > 0:
        assert _err(x.grad, x2.grad) < TOL


def test_space_to_depth_and_its_weight_form_exact():
    """b200ddp_space_to_depth + weight_prep mode 4|S<<8 against torch: pad -> pixel_unshuffle(4) for the input, and
    the stride-1 3x3 x 64-channel kernel rebuilt from the 11x11 stride-4 one; then the two convolutions agree in
    float64 (the identity the first layer's forward relies on)."""
    L = _lib.load()
    st = torch.cuda.current_stream(DEV).cuda_stream
    torch.manual_seed(11)
    N, C, H, S, pad, KH = 3, 3, 224, 4, 2, 11
    x = torch.randn(N, C, H, H, device=DEV)
    OH = (H + 2 * pad - KH) // S + 1
    KHp = (KH + S - 1) // S
    Hs = OH + KHp - 1
    out = torch.zeros(N, 64, Hs, Hs, device=DEV)
    _lib.check(L.b200ddp_space_to_depth(x.data_ptr(), out.data_ptr(), N, C, H, H, S, pad, 64, Hs, Hs, 0, st), "s2d")
    xp = torch.nn.functional.pad(x, (pad, S * Hs - H - pad, pad, S * Hs - H - pad))
    want = torch.nn.functional.pixel_unshuffle(xp, S)          # channel index (c*S + a)*S + b
    assert torch.equal(out[:, :C * S * S], want) and float(out[:, C * S * S:].abs().sum()) == 0
    w = torch.randn(64, C, KH, KH, device=DEV)
    ld = 64 * KHp * KHp
    wm = torch.empty(64, ld, device=DEV)
    _lib.check(L.b200ddp_conv2d_weight_prep(w.data_ptr(), wm.data_ptr(), 64, C, KH, KH, ld, 4 | (S << 8), 0, st), "prep4")
    wpad = torch.zeros(64, C, S * KHp, S * KHp, device=DEV)
    wpad[:, :, :KH, :KH] = w
    # w'[co][(c,a,b)][q][r] = wpad[co][c][S*q + a][S*r + b]
    w2 = wpad.reshape(64, C, KHp, S, KHp, S).permute(0, 1, 3, 5, 2, 4).reshape(64, C * S * S, KHp, KHp)
    w2 = torch.cat([w2, torch.zeros(64, 64 - C * S * S, KHp, KHp, device=DEV)], 1)
    want_m = w2.reshape(64, 2, 32, KHp * KHp).permute(0, 1, 3, 2).reshape(64, ld)     # channel-block-major
    assert torch.equal(wm, want_m)
    y_direct = torch.nn.functional.conv2d(x.double(), w.double(), stride=S, padding=pad)
    y_s2d = torch.nn.functional.conv2d(out.double(), w2.double())
    assert y_direct.shape == y_s2d.shape and float((y_direct - y_s2d).abs().max()) < 1e-9


def test_weight_prep_and_channel_sum_exact():
    L = _lib.load()
    st = torch.cuda.current_stream(DEV).cuda_stream
    w = torch.randn(7, 5, 3, 3, device=DEV)
    out = torch.empty(5, 7 * 9, device=DEV)
    _lib.check(L.b200ddp_conv2d_weight_prep(w.data_ptr(), out.data_ptr(), 7, 5, 3, 3, 63, 1, 0, st), "prep")
    want = w.flip(2, 3).permute(1, 0, 2, 3).reshape(5, 63)
    assert torch.equal(out, want)
    pad = torch.empty(7, 48, device=DEV)
    _lib.check(L.b200ddp_conv2d_weight_prep(w.data_ptr(), pad.data_ptr(), 7, 5, 3, 3, 48, 0, 0, st), "prep0")
    assert torch.equal(pad[:, :45], w.reshape(7, 45)) and float(pad[:, 45:].abs().sum()) == 0
    dy = torch.randn(9, 64, 55, 55, device=DEV)
    res = torch.empty(64, device=DEV)
    scratch = torch.zeros(int(L.b200ddp_channel_sum_scratch_bytes(9, 64)), dtype=torch.uint8, device=DEV)
    for _ in range(2):
        _lib.check(L.b200ddp_channel_sum(dy.data_ptr(), res.data_ptr(), 9, 64, 55 * 55, scratch.data_ptr(),
                                         scratch.numel(), 0, st), "chsum")
        assert torch.allclose(res.double(), dy.double().sum((0, 2, 3)), rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("layer", [1, 2])
def test_conv_table_order_path_on_wide_layers(layer, monkeypatch):
    """k_order 0 (torch's (ci,kh,kw) order + offset table) is what conv1 uses; run it on wider layers too."""
    monkeypatch.setattr(bnn, "_k_order", lambda cin: 0)
    Cin, H, Cout, K, s, p = LAYERS[layer]
    torch.manual_seed(layer)
    x = torch.randn(5, Cin, H, H, device=DEV, requires_grad=True)
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p).to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    x64 = x.detach().double().requires_grad_()
    w64 = conv.weight.detach().double().requires_grad_()
    y64 = torch.nn.functional.conv2d(x64, w64, conv.bias.detach().double(), stride=s, padding=p)
    y64.backward(dy.double())
    assert _err(y.detach(), y64.detach()) < TOL
    assert _err(conv.weight.grad, w64.grad) < TOL
    assert _err(x.grad, x64.grad) < TOL


# ---- bench-shape parity: every layer at batch 128, under the grid plans of both SM budgets -------
@pytest.fixture
def sm_budget(request):
    L = _lib.load()
    prev = L.b200ddp_set_sm_budget(request.param)
    yield request.param
    L.b200ddp_set_sm_budget(0 if prev >= 148 else prev)


@pytest.mark.parametrize("sm_budget", [148, 132], indirect=True)
@pytest.mark.parametrize("layer", [1, 2, 3, 4])
def test_conv_batch128_fwd_dgrad_wgrad_match_float64(layer, sm_budget):
    """conv2..conv5 forward / dgrad / wgrad at the bench's batch of 128, with the one-wave grid plans sized for all 148
    SMs (single GPU) and for the 132 SMs left beside a 16-channel collective (every multi-GPU run)."""
    Cin, H, Cout, K, s, p = LAYERS[layer]
    torch.manual_seed(500 + layer)
    x = torch.randn(128, Cin, H, H, device=DEV, requires_grad=True)
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p).to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    x64 = x.detach().double().requires_grad_()
    w64 = conv.weight.detach().double().requires_grad_()
    b64 = conv.bias.detach().double().requires_grad_()
    y64 = torch.nn.functional.conv2d(x64, w64, b64, stride=s, padding=p)
    y64.backward(dy.double())
    assert _err(y, y64.detach()) < TOL
    assert _err(x.grad, x64.grad) < TOL
    assert _err(conv.weight.grad, w64.grad) < TOL
    assert _err(conv.bias.grad, b64.grad) < 1e-5


@pytest.mark.parametrize("sm_budget", [132], indirect=True)
def test_conv1_batch128_under_collective_budget(sm_budget):
    torch.manual_seed(11)
    x = torch.randn(128, 3, 224, 224, device=DEV)
    conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    w64 = conv.weight.detach().double().requires_grad_()
    b64 = conv.bias.detach().double().requires_grad_()
    y64 = torch.nn.functional.conv2d(x.double(), w64, b64, stride=4, padding=2)
    y64.backward(dy.double())
    assert _err(y, y64.detach()) < TOL
    assert _err(conv.weight.grad, w64.grad) < TOL


# ---- strict mode: fp32 SIMT convolutions ------------------------------------------------------------
@pytest.mark.parametrize("layer", range(5))
def test_conv_strict_fp32_matches_float64(layer):
    Cin, H, Cout, K, s, p = LAYERS[layer]
    torch.manual_seed(700 + layer)
    x = torch.randn(3, Cin, H, H, device=DEV, requires_grad=True)
    conv = bnn.Conv2d(Cin, Cout, K, stride=s, padding=p, mode="fp32").to(DEV)
    y = conv(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    x64 = x.detach().double().requires_grad_()
    w64 = conv.weight.detach().double().requires_grad_()
    b64 = conv.bias.detach().double().requires_grad_()
    y64 = torch.nn.functional.conv2d(x64, w64, b64, stride=s, padding=p)
    y64.backward(dy.double())
    assert _err(y, y64.detach()) < 2e-6
    assert _err(x.grad, x64.grad) < 2e-6          # stride 4 included: the strict dgrad has no stride limit
    assert _err(conv.weight.grad, w64.grad) < 2e-6
    assert _err(conv.bias.grad, b64.grad) < 1e-5


def test_conv1_generic_kernel_still_matches_float64(monkeypatch):
    """The first layer through the generic implicit-GEMM kernel (k_order 0, offset table) instead of its own slab kernel."""
    monkeypatch.setenv("B200DDP_CONV1_SLAB", "0")
    torch.manual_seed(3)
    x = torch.randn(5, 3, 224, 224, device=DEV)
    conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
    y = conv(x)
    ref = torch.nn.functional.conv2d(x.double(), conv.weight.double(), conv.bias.double(), stride=4, padding=2)
    assert _err(y, ref) < TOL


def test_conv1_slab_kernel_edges_and_other_widths():
    """conv1.cu: image borders (TMA zero fill), the last single-row tile, non-finite inputs stay local, other input sizes."""
    torch.manual_seed(4)
    for (b, hw) in [(3, 224), (2, 160), (1, 67 * 4 - 4 + 8)]:
        x = torch.randn(b, 3, hw, hw, device=DEV)
        conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
        y = conv(x)
        ref = torch.nn.functional.conv2d(x.double(), conv.weight.double(), conv.bias.double(), stride=4, padding=2)
        assert y.shape == ref.shape and _err(y, ref) < TOL, (b, hw)
    # an infinite pixel must only reach the outputs whose 11x11 window really contains it (the kernel reads a 12th tap)
    x = torch.zeros(1, 3, 224, 224, device=DEV)
    x[0, 1, 100, 109] = float("inf")
    conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
    with torch.no_grad():
        conv.weight.abs_().add_(0.1)
    y = conv(x)
    ref = torch.nn.functional.conv2d(x, conv.weight, conv.bias, stride=4, padding=2)
    assert torch.equal(torch.isfinite(y), torch.isfinite(ref))
