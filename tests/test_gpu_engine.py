"""T1 native step driver (csrc/engine.cu, b200ddp/engine.py): one C call for the whole forward and one for the whole
backward must give exactly what the per-layer module path gives — same kernels, same order — and must step aside
whenever its single workspace would not be safe."""
import pytest
import torch

import b200ddp
from b200ddp import _lib, engine
from b200ddp.models import AlexNet
from oracle import reference_loop as rl

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _run(monkeypatch, use_engine, steps=3, batch=16, dropout=0.5, ddp_wrap=True, seed=11):
    monkeypatch.setenv("B200DDP_ENGINE", "1" if use_engine else "0")
    torch.manual_seed(seed)
    model = AlexNet(dropout=dropout, mode="tf32").to(DEV)
    net = b200ddp.DistributedDataParallel(model, device_ids=[0]) if ddp_wrap else model
    opt = b200ddp.optim.Adam(net.parameters(), lr=1e-3)
    crit = b200ddp.CrossEntropyLoss()
    x, y = rl.synthetic_pool(steps * batch, seed=3)
    losses, first_grads, first_logits = [], None, None
    for s in range(steps):
        xb, yb = x[s * batch:(s + 1) * batch].to(DEV), y[s * batch:(s + 1) * batch].to(DEV)
        opt.zero_grad()
        out = net(xb)
        loss = crit(out, yb)
        loss.backward()
        if s == 0:
            torch.cuda.synchronize()
            first_logits = out.detach().clone()
            first_grads = [p.grad.detach().clone() for p in model.parameters()]
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    return model, losses, first_logits, first_grads


def test_engine_is_bit_identical_to_module_path(monkeypatch):
    before = _lib.launch_count()
    ma, la, za, ga = _run(monkeypatch, True)
    n_engine = _lib.launch_count() - before
    assert getattr(ma, "_engine_plans", None), "the native plan did not run"
    before = _lib.launch_count()
    mb, lb, zb, gb = _run(monkeypatch, False)
    n_module = _lib.launch_count() - before
    assert not getattr(mb, "_engine_plans", None)
    assert n_engine == n_module, (n_engine, n_module)            # the same kernels, launch for launch
    assert la == lb and torch.equal(za, zb)
    for (k, _), a, b in zip(ma.named_parameters(), ga, gb):
        assert torch.equal(a, b), k
    for (k, p), (_, q) in zip(ma.named_parameters(), mb.named_parameters()):
        assert torch.equal(p, q), k                               # after three fused Adam steps


def test_engine_without_ddp_and_without_dropout(monkeypatch):
    ma, la, za, ga = _run(monkeypatch, True, steps=2, dropout=0.0, ddp_wrap=False)
    mb, lb, zb, gb = _run(monkeypatch, False, steps=2, dropout=0.0, ddp_wrap=False)
    assert la == lb and torch.equal(za, zb)
    for a, b in zip(ga, gb):
        assert torch.equal(a, b)


def test_engine_steps_aside_when_two_graphs_are_alive(monkeypatch):
    """forward, forward, backward, backward: the second forward must not overwrite the first one's activations."""
    monkeypatch.setenv("B200DDP_ENGINE", "1")
    torch.manual_seed(2)
    model = AlexNet(dropout=0.0, mode="tf32").to(DEV)
    crit = b200ddp.CrossEntropyLoss()
    xa, ya = rl.synthetic_pool(8, seed=1)
    xb, yb = rl.synthetic_pool(8, seed=2)
    xa, ya, xb, yb = xa.to(DEV), ya.to(DEV), xb.to(DEV), yb.to(DEV)

    def grads(x, y):
        for p in model.parameters():
            p.grad = None
        crit(model(x), y).backward()
        torch.cuda.synchronize()
        return [p.grad.clone() for p in model.parameters()]
    ga, gb = grads(xa, ya), grads(xb, yb)
    for p in model.parameters():
        p.grad = None
    la = crit(model(xa), ya)          # native plan; its graph stays alive ...
    lb = crit(model(xb), yb)          # ... so this one has to take the module path
    with torch.no_grad():
        model(xb)                     # and an evaluation in between must leave the workspace alone, too
    la.backward()
    lb.backward()
    torch.cuda.synchronize()
    for p, a, b in zip(model.parameters(), ga, gb):
        assert torch.allclose(p.grad, a + b, rtol=1e-5, atol=1e-7)
    # a dropped graph frees the plan again
    out = model(xa)
    plan = next(iter(model._engine_plans.values()))
    assert plan.busy()
    del out
    assert not plan.busy()


def test_engine_inference_and_ragged_batches(monkeypatch):
    torch.manual_seed(5)
    model = AlexNet(dropout=0.5, mode="tf32").to(DEV).eval()
    x, _ = rl.synthetic_pool(13, seed=6)
    x = x.to(DEV)
    with torch.no_grad():
        monkeypatch.setenv("B200DDP_ENGINE", "1")
        za, ha = model(x), model.forward_hidden(x)
        zc = model(x[:5])
        monkeypatch.setenv("B200DDP_ENGINE", "0")
        zb, hb = model(x), model.forward_hidden(x)
    assert torch.equal(za, zb) and torch.equal(ha, hb) and torch.allclose(zc, zb[:5], rtol=1e-3, atol=1e-4)
    assert set(model._engine_plans) == {13, 5}


def test_engine_follows_parameters_into_the_arena_and_no_sync(monkeypatch):
    """A plan built before the DDP wrapper re-points the parameters must be rebuilt; accumulation under no_sync adds up."""
    monkeypatch.setenv("B200DDP_ENGINE", "1")
    torch.manual_seed(9)
    model = AlexNet(dropout=0.0, mode="tf32").to(DEV)
    crit = b200ddp.CrossEntropyLoss()
    xa, ya = rl.synthetic_pool(8, seed=1)
    xb, yb = rl.synthetic_pool(8, seed=2)
    xa, ya, xb, yb = xa.to(DEV), ya.to(DEV), xb.to(DEV), yb.to(DEV)
    crit(model(xa), ya).backward()                  # plan over the loose parameters
    ga = [p.grad.clone() for p in model.parameters()]
    ddp = b200ddp.DistributedDataParallel(model, device_ids=[0])
    for p in model.parameters():
        p.grad = None
    crit(ddp(xb), yb).backward()
    torch.cuda.synchronize()
    gb = [p.grad.clone() for p in model.parameters()]
    assert all(ddp.arena.grad_in_arena(i) for i in range(len(ddp.arena.params)))
    for p in model.parameters():
        p.grad = None
    with ddp.no_sync():
        crit(ddp(xa), ya).backward()
    crit(ddp(xb), yb).backward()
    torch.cuda.synchronize()
    for p, a, b in zip(model.parameters(), ga, gb):
        assert torch.allclose(p.grad, a + b, rtol=1e-5, atol=1e-7)


def test_stock_criterion_works_on_the_engine_output(monkeypatch):
    """the script's own nn.CrossEntropyLoss() may stay; the logits are an ordinary autograd tensor."""
    monkeypatch.setenv("B200DDP_ENGINE", "1")
    torch.manual_seed(6)
    model = AlexNet(dropout=0.0, mode="tf32").to(DEV)
    x, y = rl.synthetic_pool(8, seed=4)
    x, y = x.to(DEV), y.to(DEV)

    def grads(crit):
        for p in model.parameters():
            p.grad = None
        loss = crit(model(x), y)
        loss.backward()
        torch.cuda.synchronize()
        return float(loss), [p.grad.clone() for p in model.parameters()]
    la, ga = grads(b200ddp.CrossEntropyLoss())
    lb, gb = grads(torch.nn.CrossEntropyLoss())
    assert abs(la - lb) < 1e-6 * abs(lb)
    # the two loss kernels differ in the last bit of dlogits; TF32 rounding of the backward operands turns that into
    # ~1e-4 relative differences of individual products, so compare in rel-L2
    for a, b in zip(ga, gb):
        assert float((a.double() - b.double()).norm() / b.double().norm()) < 2e-3
