"""SURVEY 8(f) rows on one B200: the evaluate() path (f1), the GPU input transform (f2) and the asynchronous
checkpoint / resume (f3), each against the oracle, the committed golden vectors from the real upstream code, and
size-independent properties at the bench's full sizes."""
import ctypes
import hashlib
import os

import numpy as np
import pytest
import torch
from torch import nn

import b200ddp
from b200ddp import _lib
from b200ddp.models import AlexNet
from oracle import ddp_oracle as O
from oracle import input_pipeline as ip
from oracle import reference_loop as rl

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


# ---- f1: evaluate() ---------------------------------------------------------------------------------------------------
def _head_eval(h, w, b, y, metrics, scratch, logits=None):
    L = _lib.load()
    _lib.check(L.b200ddp_head_eval(h.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), _lib.ptr(logits), h.shape[0],
                                   w.shape[0], w.shape[1], metrics.data_ptr(), scratch.data_ptr(), scratch.numel(), 0,
                                   _lib.cur_stream(DEV)), "head_eval")


def test_head_eval_matches_reference_evaluate_golden(golden_dir):
    """Through the C ABI: recorded hidden activations + head weights of the reference's own evaluate() run ->
    (sum of losses, correct, total); counts bit-exact, loss 1e-5 (golden from make_golden.gen_eval)."""
    g = np.load(os.path.join(golden_dir, "eval.npz"))
    bs = int(g["batch"])
    h = torch.from_numpy(g["hidden"]).to(DEV)
    w, b = torch.from_numpy(g["w"]).to(DEV), torch.from_numpy(g["b"]).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    metrics = torch.zeros(3, device=DEV)
    scratch = torch.zeros(int(_lib.load().b200ddp_eval_scratch_bytes(bs)), dtype=torch.uint8, device=DEV)
    logits = torch.empty(20, 10, device=DEV)
    for i in range(0, 20, bs):                         # 8, 8, 4: the last batch is ragged
        _head_eval(h[i:i + bs].contiguous(), w, b, y[i:i + bs].contiguous(), metrics, scratch, logits[i:i + bs])
    m = metrics.cpu().numpy()
    assert m[1] == float(g["correct"]) and m[2] == float(g["total"])
    assert abs(m[0] - float(g["total_test_loss"])) < 1e-5 * abs(float(g["total_test_loss"]))
    np.testing.assert_allclose(logits.cpu().numpy(), g["logits"], rtol=1e-4, atol=1e-5)
    # the logits-only entry point gives the same three sums
    m2 = torch.zeros(3, device=DEV)
    z = torch.from_numpy(g["logits"]).to(DEV)
    for i in range(0, 20, bs):
        _lib.check(_lib.load().b200ddp_logits_eval(z[i:i + bs].contiguous().data_ptr(), y[i:i + bs].contiguous().data_ptr(),
                                                   min(bs, 20 - i), 10, m2.data_ptr(), scratch.data_ptr(), scratch.numel(), 0,
                                                   _lib.cur_stream(DEV)), "logits_eval")
    assert m2[1] == m[1] and m2[2] == m[2] and abs(float(m2[0]) - m[0]) < 1e-4


def test_head_eval_ties_and_bad_labels_follow_torch():
    h = torch.zeros(3, 64, device=DEV)
    w = torch.zeros(10, 64, device=DEV)
    b = torch.tensor([1., 3., 3., 0., 0., 0., 0., 0., 0., 0.], device=DEV)       # tie between classes 1 and 2
    y = torch.tensor([1, 2, 12], device=DEV)
    metrics = torch.zeros(3, device=DEV)
    scratch = torch.zeros(int(_lib.load().b200ddp_eval_scratch_bytes(3)), dtype=torch.uint8, device=DEV)
    logits = torch.empty(3, 10, device=DEV)
    _head_eval(h, w, b, y, metrics, scratch, logits)
    pred = torch.max(logits, 1)[1]
    assert pred.tolist() == [1, 1, 1]                                             # first maximum, as torch.max
    assert float(metrics[1]) == float((pred == y).sum()) == 1.0                   # only row 0 is a hit
    assert float(metrics[2]) == 3.0
    ref = nn.functional.cross_entropy(logits[:2], y[:2], reduction="sum")         # the out-of-range label adds no loss
    assert abs(float(metrics[0]) - float(ref)) < 1e-5


def test_logits_eval_nan_wins_like_torch_max():
    z = torch.randn(5, 10, device=DEV)
    z[1, 7] = float("nan")
    z[3, 2] = float("nan")
    z[3, 6] = float("nan")                                                        # two NaNs: the first one is reported
    pred = torch.max(z, 1)[1]
    assert int(pred[1]) == 7 and int(pred[3]) == 2
    y = pred.clone()
    y[0] = (y[0] + 1) % 10
    m = torch.zeros(3, device=DEV)
    scratch = torch.zeros(int(_lib.load().b200ddp_eval_scratch_bytes(5)), dtype=torch.uint8, device=DEV)
    _lib.check(_lib.load().b200ddp_logits_eval(z.data_ptr(), y.data_ptr(), 5, 10, m.data_ptr(), scratch.data_ptr(),
                                               scratch.numel(), 0, _lib.cur_stream(DEV)), "logits_eval")
    assert float(m[1]) == 4.0 and float(m[2]) == 5.0 and not torch.isfinite(m[0])  # NaN rows make the loss NaN, as in torch


def test_evaluate_function_matches_reference_loop_on_the_toy_model():
    """b200ddp.evaluate (same signature as the reference's evaluate()) on AlexNet vs the restated reference loop
    (oracle/reference_loop.evaluate_epoch, asserted equal to the reference's own) over stock torch on the same GPU;
    the ragged last batch included. Strict fp32 mode so that the comparison is at 1e-5."""
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        tv = rl.make_model(seed=1234, dropout=0.5).to(DEV)
        ours = AlexNet(dropout=0.5, mode="fp32")
        ours.load_state_dict(tv.state_dict())
        ddp = b200ddp.DistributedDataParallel(ours.to(DEV), device_ids=[0])
        x, _ = rl.synthetic_pool(26, seed=5)
        with torch.no_grad():
            pred = tv.eval()(x.to(DEV)).argmax(1).cpu()
        y = torch.where(torch.arange(26) % 3 == 0, pred, (pred + 1) % 10)
        loader = torch.utils.data.DataLoader(torch.utils.data.TensorDataset(x, y), batch_size=8)
        want = rl.evaluate_epoch(tv, loader, nn.CrossEntropyLoss(), DEV)
        before = _lib.launch_count()
        got = b200ddp.evaluate(ddp, loader, b200ddp.CrossEntropyLoss(), DEV)
        assert _lib.launch_count() > before
        assert float(got[1]) == float(want[1]) == 9.0 and float(got[2]) == float(want[2]) == 26.0
        assert abs(float(got[0]) - float(want[0])) < 1e-5 * abs(float(want[0]))
        assert all(t.shape == (1,) and t.device.type == "cuda" for t in got)
        # one collective for all metrics (single rank: identity)
        a, b_, c = b200ddp.all_reduce_metrics(*got)
        assert float(c) == 26.0
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_inference_forward_writes_no_gate_or_argmax_bytes():
    """Under no_grad the Linear / pool layers are launched without their saved-for-backward outputs."""
    from b200ddp import nn as bnn
    lin = bnn.Linear(64, 64, relu=True, mode="fp32").to(DEV)
    x = torch.randn(4, 64, device=DEV)
    with torch.no_grad():
        y0 = lin(x)
    y1 = lin(x)
    assert torch.equal(y0, y1.detach()) and y1.grad_fn is not None
    pool = bnn.ReLUMaxPool2d()
    t = torch.randn(2, 3, 9, 9, device=DEV)
    with torch.no_grad():
        p0 = pool(t)
    assert torch.equal(p0, pool(t.requires_grad_()).detach())


# ---- f2: input transform -------------------------------------------------------------------------------------------------
def test_gpu_transform_is_bit_exact_with_torchvision_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "input_pipeline.npz"))
    tf = b200ddp.data.GpuTransform(224, train=True)
    imgs = torch.from_numpy(g["images"]).to(DEV)
    flips = torch.from_numpy(g["flips"]).to(DEV)
    out = tf(imgs, flips).cpu().numpy()
    assert out.shape == (4, 3, 224, 224) and out.dtype == np.float32
    for i in range(4):
        assert np.array_equal(out[i][:, :16, :], g["first_rows"][i])
        assert hashlib.sha256(np.ascontiguousarray(out[i]).tobytes()).hexdigest() == str(g["sha256"][i])


def test_gpu_transform_matches_oracle_on_random_and_ragged_inputs():
    rng = np.random.default_rng(5)
    for (b, h, w) in [(1, 32, 32), (3, 32, 32), (2, 17, 29)]:
        imgs = rng.integers(0, 256, (b, h, w, 3), dtype=np.uint8)
        flips = rng.integers(0, 2, (b,), dtype=np.uint8)
        size = 224 if h == 32 else 61
        tf = b200ddp.data.GpuTransform(size, train=True)
        got = tf(torch.from_numpy(imgs).to(DEV), torch.from_numpy(flips).to(DEV)).cpu().numpy()
        if h == w:
            assert np.array_equal(got, ip.transform_batch(imgs, flips, size))
        else:               # rectangular: each axis is resampled independently with its own table (oracle passes)
            for i in range(b):
                r = ip._pass(ip._pass(imgs[i], 1, size), 0, size)
                if flips[i]:
                    r = r[:, ::-1, :]
                t = np.ascontiguousarray(r.transpose(2, 0, 1)).astype(np.float32) / np.float32(255.0)
                ref = (t - np.asarray(ip.MEAN, np.float32).reshape(3, 1, 1)) / np.asarray(ip.STD, np.float32).reshape(3, 1, 1)
                assert np.array_equal(got[i], ref.astype(np.float32))
    # test-time transform: no flip at all
    tf = b200ddp.data.GpuTransform(224, train=False)
    imgs = rng.integers(0, 256, (2, 32, 32, 3), dtype=np.uint8)
    assert np.array_equal(tf(torch.from_numpy(imgs).to(DEV)).cpu().numpy(), ip.transform_batch(imgs, [0, 0]))


def test_gpu_transform_full_batch_properties():
    """Batch 128 (the bench's size): flipping twice is the identity on columns, a constant image maps to a constant,
    and the result is independent of how the batch is split into calls."""
    rng = np.random.default_rng(9)
    imgs = torch.from_numpy(rng.integers(0, 256, (128, 32, 32, 3), dtype=np.uint8)).to(DEV)
    imgs[5] = 77
    tf = b200ddp.data.GpuTransform(224, train=True)
    zeros, ones = torch.zeros(128, dtype=torch.uint8, device=DEV), torch.ones(128, dtype=torch.uint8, device=DEV)
    a, b = tf(imgs, zeros), tf(imgs, ones)
    assert torch.equal(a, b.flip(-1))
    for c in range(3):
        v = np.float32((np.float32(77) / np.float32(255) - np.float32(ip.MEAN[c])) / np.float32(ip.STD[c]))
        assert float(a[5, c].min()) == float(a[5, c].max()) == float(v)
    halves = torch.cat([tf(imgs[:50], zeros[:50]), tf(imgs[50:], zeros[50:])])
    assert torch.equal(a, halves)


def test_device_loader_delivers_sampler_order_one_batch_ahead():
    rng = np.random.default_rng(1)
    imgs = torch.from_numpy(rng.integers(0, 256, (50, 32, 32, 3), dtype=np.uint8))
    labels = torch.arange(50)
    sampler = b200ddp.DistributedSampler(torch.utils.data.TensorDataset(labels), num_replicas=2, rank=1, shuffle=True)
    sampler.set_epoch(3)
    want_idx = list(sampler)
    tf = b200ddp.data.GpuTransform(224, train=False)
    loader = b200ddp.data.DeviceLoader(imgs, labels, batch_size=8, device=DEV, sampler=sampler, transform=tf)
    assert len(loader) == 4
    seen = []
    for x, y in loader:
        assert x.is_cuda and x.dtype == torch.float32 and x.shape[1:] == (3, 224, 224)
        ref = ip.transform_batch(imgs[y.cpu()].numpy(), [0] * len(y))
        assert np.array_equal(x.cpu().numpy(), ref)
        seen += y.tolist()
    assert seen == want_idx


# ---- f3: checkpoint / resume ----------------------------------------------------------------------------------------------
def test_async_checkpoint_is_a_consistent_snapshot_and_resumes(tmp_path):
    torch.manual_seed(0)
    model = AlexNet(dropout=0.0, mode="tf32").to(DEV)
    ddp = b200ddp.DistributedDataParallel(model, device_ids=[0])
    opt = b200ddp.optim.Adam(ddp.parameters(), lr=1e-3)
    crit = b200ddp.CrossEntropyLoss()
    ds = torch.utils.data.TensorDataset(torch.arange(40))
    sampler = b200ddp.DistributedSampler(ds, num_replicas=1, rank=0, shuffle=True)
    sampler.set_epoch(4)
    x, y = rl.synthetic_pool(16, seed=3)
    x, y = x.to(DEV), y.to(DEV)

    def step():
        opt.zero_grad()
        loss = crit(ddp(x), y)
        loss.backward()
        opt.step()
        return float(loss)
    step(), step()
    want = {k: v.detach().clone() for k, v in ddp.state_dict().items()}
    ck = b200ddp.AsyncCheckpointer(ddp, optimizer=opt, sampler=sampler)
    path = str(tmp_path / "ckpt_0.pt")
    ck.save(path, epoch=4)
    l_next = step()                   # training goes on while the file is written; the file must hold the OLD values
    ck.wait()
    sd = torch.load(path, map_location="cpu")            # the reference's usage (README.md:51-52)
    assert list(sd) == list(want) and len(sd) == 16 and list(sd)[0] == "module.features.0.weight"
    for k in want:
        assert sd[k].dtype == torch.float32 and torch.equal(sd[k], want[k].cpu()), k
    # resume into a fresh model / optimizer / sampler and take the same next step
    torch.manual_seed(123)
    model2 = AlexNet(dropout=0.0, mode="tf32").to(DEV)
    ddp2 = b200ddp.DistributedDataParallel(model2, device_ids=[0])
    opt2 = b200ddp.optim.Adam(ddp2.parameters(), lr=5e-2)
    sampler2 = b200ddp.DistributedSampler(ds, num_replicas=1, rank=0, shuffle=True)
    info = b200ddp.load_checkpoint(path, ddp2, opt2, sampler2, map_location=str(DEV))
    assert info["epoch"] == 4 and list(sampler2) == list(sampler) and opt2.param_groups[0]["lr"] == 1e-3
    opt2.zero_grad()
    loss2 = crit(ddp2(x), y)
    loss2.backward()
    opt2.step()
    assert abs(float(loss2) - l_next) < 1e-6
    for (k, p), (_, q) in zip(model.named_parameters(), model2.named_parameters()):
        assert torch.equal(p, q), k          # same Adam moments and step count -> identical third step
