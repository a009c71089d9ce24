"""Per-rank bodies for the multi-process tests (spawned with torch.multiprocessing; gloo on CPU, nccl on GPU)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import warnings  # noqa: E402

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch import nn  # noqa: E402

warnings.filterwarnings("ignore", message=".*plumbing mode.*")


def _init(rank, world, port, backend="gloo"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    if backend == "nccl":
        torch.cuda.set_device(rank)
    dist.init_process_group(backend, rank=rank, world_size=world)


def _mlp(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(12, 300), nn.ReLU(), nn.Linear(300, 300), nn.ReLU(), nn.Linear(300, 5))


def _data(rank, step, n=6):
    g = torch.Generator().manual_seed(1000 + 17 * rank + step)
    return torch.randn(n, 12, generator=g), torch.randint(0, 5, (n,), generator=g)


def _gather_params(model, world):
    """All ranks' flattened parameters (collectives run on the parameters' own device: NCCL has no CPU path)."""
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    out = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(out, flat)
    return [t.cpu() for t in out]


def run(rank, world, port, case, out_dir, backend="gloo"):
    _init(rank, world, port, backend)
    try:
        globals()["case_" + case](rank, world, out_dir, torch.device("cuda", rank) if backend == "nccl" else torch.device("cpu"))
        dist.barrier()
        if rank == 0:
            open(os.path.join(out_dir, case + ".ok"), "w").write("ok")
    finally:
        dist.destroy_process_group()


# ---- cases -----------------------------------------------------------------------------------------------------
def case_mean_gradient(rank, world, out_dir, dev):
    import b200ddp
    twin = _mlp().to(dev)
    model = _mlp(seed=rank + 5).to(dev)                 # different init per rank: ctor must broadcast rank 0's
    ddp = b200ddp.DistributedDataParallel(model)
    ref0 = _mlp(seed=5).to(dev)
    for p, q in zip(model.parameters(), ref0.parameters()):
        assert torch.equal(p.detach().cpu(), q.detach().cpu()), "parameters were not broadcast from rank 0"
    twin.load_state_dict(model.state_dict())
    x, y = _data(rank, 0)
    x, y = x.to(dev), y.to(dev)
    nn.functional.cross_entropy(twin(x), y).backward()
    nn.functional.cross_entropy(ddp(x), y).backward()
    for p, q in zip(model.parameters(), twin.parameters()):
        local = q.grad.detach().clone()
        dist.all_reduce(local)
        assert torch.allclose(p.grad, local / world, rtol=1e-5, atol=1e-7)
        assert ddp.arena.grad_in_arena(ddp.arena.index_of(p))
    # evaluation under no_grad must not arm the reducer
    with torch.no_grad():
        ddp.eval()
        ddp(x)
        ddp.train()


def _stock_vs_ours(rank, world, dev, steps, ours_kwargs, stock_step):
    import b200ddp
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    a = _mlp().to(dev)
    b = _mlp().to(dev)
    ours = b200ddp.DistributedDataParallel(a, **ours_kwargs)
    stock = TorchDDP(b, device_ids=[dev.index] if dev.type == "cuda" else None)
    opt_a = b200ddp.optim.Adam(ours.parameters(), lr=1e-3)
    opt_b = torch.optim.Adam(stock.parameters(), lr=1e-3)
    for s in range(steps):
        x, y = _data(rank, s)
        x, y = x.to(dev), y.to(dev)
        opt_a.zero_grad()
        la = nn.functional.cross_entropy(ours(x), y)
        la.backward()
        opt_a.step()
        opt_b.zero_grad()
        stock_step(stock, x, y, rank, s)
        opt_b.step()
        for p, q in zip(a.parameters(), b.parameters()):
            assert torch.allclose(p.grad, q.grad, rtol=2e-5, atol=1e-7), f"grad mismatch at step {s}"
            assert torch.allclose(p, q, rtol=1e-5, atol=1e-6), f"param mismatch at step {s}"
    ps = _gather_params(a, world)
    assert all(torch.equal(ps[0], t) for t in ps), "ranks diverged"
    return ours, stock


def case_parity_with_stock_ddp(rank, world, out_dir, dev):
    def step(stock, x, y, rank, s):
        nn.functional.cross_entropy(stock(x), y).backward()
    ours, _ = _stock_vs_ours(rank, world, dev, 5, {}, step)
    assert ours.reducer.first_ready_order is not None


def _clip_restatement(kind, c):
    """stock DDP under no_sync + local clip + manual all-reduce + divide."""
    def step(stock, x, y, rank, s):
        with stock.no_sync():
            nn.functional.cross_entropy(stock(x), y).backward()
        params = [p for p in stock.parameters()]
        if kind == "norm":
            torch.nn.utils.clip_grad_norm_(params, c)
        else:
            torch.nn.utils.clip_grad_value_(params, c)
        for p in params:
            dist.all_reduce(p.grad)
            p.grad /= dist.get_world_size()
    return step


def case_clip_value_before_aggregation(rank, world, out_dir, dev):
    _stock_vs_ours(rank, world, dev, 3, {"clip_grad_value": 0.01}, _clip_restatement("value", 0.01))


def case_clip_norm_before_aggregation(rank, world, out_dir, dev):
    ours, _ = _stock_vs_ours(rank, world, dev, 3, {"clip_grad_norm": 0.05}, _clip_restatement("norm", 0.05))
    assert float(ours.reducer.last_local_norm) > 0.05      # the clip was active


def case_no_sync_accumulation(rank, world, out_dir, dev):
    import b200ddp
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    a, b = _mlp().to(dev), _mlp().to(dev)
    ours = b200ddp.DistributedDataParallel(a)
    stock = TorchDDP(b, device_ids=[dev.index] if dev.type == "cuda" else None)
    for model in (ours, stock):
        x0, y0 = _data(rank, 0)
        x1, y1 = _data(rank, 1)
        with model.no_sync():
            nn.functional.cross_entropy(model(x0.to(dev)), y0.to(dev)).backward()
        if model is ours:
            local_only = [p.grad.detach().clone() for p in a.parameters()]
        nn.functional.cross_entropy(model(x1.to(dev)), y1.to(dev)).backward()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=2e-5, atol=1e-7)
    # inside no_sync nothing was communicated: ranks held different gradients
    flat = torch.cat([g.reshape(-1) for g in local_only])
    out = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(out, flat)
    assert not torch.equal(out[0], out[1])


def case_fault_containment(rank, world, out_dir, dev):
    """README.md:92-95: one rank's bad gradient must not wreck the aggregate when clipping comes first."""
    import b200ddp
    x, y = _data(rank, 0)
    x, y = x.to(dev), y.to(dev)
    # (a) huge but finite gradient on rank 1, norm clipping
    m = _mlp().to(dev)
    ddp = b200ddp.DistributedDataParallel(m, clip_grad_norm=1.0)
    scale = 1e20 if rank == 1 else 1.0
    (nn.functional.cross_entropy(ddp(x), y) * scale).backward()
    total = torch.sqrt(sum((p.grad.double() ** 2).sum() for p in m.parameters()))
    assert torch.isfinite(total) and float(total) <= 1.0 + 1e-4
    # (b) NaN gradient on rank 1: stock semantics poison everybody, nonfinite='zero' contains it
    m2 = _mlp().to(dev)
    ddp2 = b200ddp.DistributedDataParallel(m2, clip_grad_norm=1.0, nonfinite="zero")
    loss = nn.functional.cross_entropy(ddp2(x), y)
    if rank == 1:
        loss = loss * float("nan")
    loss.backward()
    assert all(torch.isfinite(p.grad).all() for p in m2.parameters())
    assert any(float(p.grad.abs().sum()) > 0 for p in m2.parameters())
    m3 = _mlp().to(dev)
    ddp3 = b200ddp.DistributedDataParallel(m3, clip_grad_norm=1.0)
    loss = nn.functional.cross_entropy(ddp3(x), y)
    if rank == 1:
        loss = loss * float("nan")
    loss.backward()
    assert not all(torch.isfinite(p.grad).all() for p in m3.parameters())   # README's warning, reproduced


def case_checkpoint_rank0(rank, world, out_dir, dev):
    import b200ddp
    m = _mlp(seed=3).to(dev)
    ddp = b200ddp.DistributedDataParallel(m)
    path = os.path.join(out_dir, "ckpt_0.pt")
    if rank == 0:                                    # reference :217-221
        torch.save(ddp.state_dict(), path)
    dist.barrier()                                   # reference :223
    sd = torch.load(path, map_location="cpu")        # README.md:51-52
    assert all(k.startswith("module.") for k in sd) and len(sd) == 6
    fresh = b200ddp.DistributedDataParallel(_mlp(seed=9).to(dev))
    fresh.load_state_dict(sd)
    for (k, v), (_, w) in zip(fresh.state_dict().items(), ddp.state_dict().items()):
        assert torch.equal(v.cpu(), w.cpu()), k


def case_plumbing_alexnet(rank, world, out_dir, dev):
    """BASELINE config 1: the toy model through our surface on CPU/gloo (host logic only)."""
    import b200ddp
    from oracle import reference_loop as rl
    x, y = rl.synthetic_pool(32, seed=1, hw=224)
    ds = torch.utils.data.TensorDataset(x, y)
    sampler = b200ddp.DistributedSampler(ds, shuffle=True, num_replicas=world, rank=rank)
    loader = torch.utils.data.DataLoader(ds, batch_size=8, sampler=sampler)
    model = rl.make_model(seed=0)
    ddp = b200ddp.DistributedDataParallel(model)
    opt = b200ddp.optim.Adam(ddp.parameters(), lr=0.001)
    sampler.set_epoch(0)
    seen = []
    losses = []
    rl.train_epoch(ddp, loader, nn.CrossEntropyLoss(), opt, torch.device("cpu"), max_steps=2, losses=losses)
    seen = list(sampler)
    allseen = [None] * world
    dist.all_gather_object(allseen, seen)      # (gloo only: this case is the CPU plumbing config)
    assert not set(allseen[0]) & set(allseen[1])               # disjoint shards
    ps = _gather_params(model, world)
    assert torch.equal(ps[0], ps[1])                           # replicas stay identical
    assert ddp.bucket_sizes() == [67289128, 151011328, 9878784]
    if rank == 0:
        torch.save(ddp.state_dict(), os.path.join(out_dir, "ckpt_0.pt"))
    dist.barrier()
    sd = torch.load(os.path.join(out_dir, "ckpt_0.pt"), map_location="cpu")
    assert len(sd) == 16 and list(sd)[0] == "module.features.0.weight"


def case_accelerator(rank, world, out_dir, dev):
    import b200ddp
    from b200ddp.accelerate import Accelerator
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    acc = Accelerator(cpu=(dev.type == "cpu"))
    assert acc.num_processes == world and acc.is_local_main_process == (rank == 0)
    model = _mlp().to(acc.device)
    opt = b200ddp.optim.Adam(model.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(0)
    ds = torch.utils.data.TensorDataset(torch.randn(44, 12, generator=g), torch.randint(0, 5, (44,), generator=g))
    loader = torch.utils.data.DataLoader(ds, batch_size=4)
    model, opt, loader = acc.prepare(model, opt, loader)
    assert isinstance(model, b200ddp.DistributedDataParallel)
    n = 0
    for xb, yb in loader:
        assert xb.device.type == acc.device.type and xb.shape[0] == 4
        opt.zero_grad()
        loss = nn.functional.cross_entropy(model(xb), yb)
        acc.backward(loss)
        opt.step()
        n += 1
    assert n == len(loader) == 6                     # 11 batches over 2 ranks, evened out
    ps = _gather_params(model.module, world)
    assert torch.equal(ps[0], ps[1])
    acc.wait_for_everyone()
    acc.save_model(model, os.path.join(out_dir, "acc"))
    acc.wait_for_everyone()
    assert os.path.exists(os.path.join(out_dir, "acc", "model.safetensors"))


def case_alexnet_vs_stock_ddp(rank, world, out_dir, dev):
    """BASELINE config 3: toy model, DistributedSampler + set_epoch, our DDP/Adam/CE vs the stock stack."""
    import b200ddp
    from b200ddp.models import AlexNet
    from oracle import reference_loop as rl
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    # Stock arm with torch's default numerics (TF32 convolutions, fp32 Linear): what the reference script runs.
    # Ours: the default product path (TF32 tcgen05 convolutions and Linear layers).
    torch.backends.cudnn.allow_tf32 = True
    tv = rl.make_model(seed=1234, dropout=0.0).to(dev)
    ours = AlexNet(dropout=0.0, mode="tf32")
    ours.load_state_dict(tv.state_dict())
    ours = ours.to(dev)
    ddp = b200ddp.DistributedDataParallel(ours, device_ids=[rank])
    stock = TorchDDP(tv, device_ids=[rank])
    opt = b200ddp.optim.Adam(ddp.parameters(), lr=1e-4)
    opt_ref = torch.optim.Adam(stock.parameters(), lr=1e-4)
    x, y = rl.synthetic_pool(96, seed=7)
    ds = torch.utils.data.TensorDataset(x, y)
    sa = b200ddp.DistributedSampler(ds, shuffle=True, num_replicas=world, rank=rank)
    sb = torch.utils.data.DistributedSampler(ds, shuffle=True, num_replicas=world, rank=rank)
    for epoch in (0, 1):
        sa.set_epoch(epoch)
        sb.set_epoch(epoch)
        assert list(sa) == list(sb)
        la, lb = [], []
        rl.train_epoch(ddp, torch.utils.data.DataLoader(ds, batch_size=16, sampler=sa), b200ddp.CrossEntropyLoss(),
                       opt, dev, losses=la)
        rl.train_epoch(stock, torch.utils.data.DataLoader(ds, batch_size=16, sampler=sb), nn.CrossEntropyLoss(),
                       opt_ref, dev, losses=lb)
        for a, b in zip(la, lb):
            assert abs(a - b) < 3e-3 * abs(b), (la, lb)
    for (k, p), (_, q) in zip(ours.named_parameters(), tv.named_parameters()):
        rel = float((p.detach().double() - q.detach().double()).norm() / q.detach().double().norm())
        assert rel < 2e-2, (k, rel)        # 12 Adam steps: m/sqrt(v) turns tiny gradient differences into +-lr moves
    ps = _gather_params(ours, world)
    assert all(torch.equal(ps[0], t) for t in ps)
    assert ddp.bucket_sizes() == [67289128, 151011328, 9878784]


def case_syncbn_two_ranks(rank, world, out_dir, dev):
    """BASELINE config 4 (scaled down): convert_sync_batchnorm + DDP vs plain BN on the global batch."""
    import b200ddp
    torch.manual_seed(0)
    C = 64
    net = nn.Sequential(nn.Conv2d(3, C, 3, padding=1), nn.BatchNorm2d(C), nn.ReLU(), nn.Conv2d(C, 8, 3, padding=1))
    ref = nn.Sequential(nn.Conv2d(3, C, 3, padding=1), nn.BatchNorm2d(C), nn.ReLU(), nn.Conv2d(C, 8, 3, padding=1))
    ref.load_state_dict(net.state_dict())
    net = b200ddp.SyncBatchNorm.convert_sync_batchnorm(net).to(dev)       # README.md:79-81: convert, then wrap
    ddp = b200ddp.DistributedDataParallel(net, device_ids=[rank])
    ref = ref.to(dev)
    g = torch.Generator().manual_seed(5)
    xall = torch.randn(world * 6, 3, 20, 20, generator=g).to(dev)
    xloc = xall[rank * 6:(rank + 1) * 6]
    out = ddp(xloc)
    out.square().mean().backward()
    oref = ref(xall)
    (oref.square().mean()).backward()        # mean over the global batch == mean over ranks of local means
    assert torch.allclose(out, oref[rank * 6:(rank + 1) * 6], rtol=1e-3, atol=1e-4)
    assert torch.allclose(net[1].running_mean, ref[1].running_mean, rtol=1e-4, atol=1e-6)
    assert torch.allclose(net[1].running_var, ref[1].running_var, rtol=1e-4, atol=1e-6)
    for (k, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=2e-3, atol=1e-5), k


def case_own_allreduce_known_values(rank, world, out_dir, dev):
    """R4: the library's own all-reduce kernel over the registered arena (multimem / peer loads), through the C ABI,
    against exact expectations and against torch.distributed's NCCL all-reduce: whole arena, ranges at offsets, sizes
    that do not divide by the world size, many launches in a row (flag epochs), several CTA counts."""
    from b200ddp import _lib
    from b200ddp.comm import get_comm
    comm = get_comm(dev, None)
    n = 3 * 1024 * 1024 + 64 * 5
    buf = comm.arena_alloc(n)
    assert buf.is_cuda and buf.numel() == n and float(buf.abs().sum()) == 0.0
    peer = comm.set_algo("own")
    if comm.arena_mode < 1 or not peer:
        import pytest
        print("peer path unavailable on this box (arena_mode %d)" % comm.arena_mode)
        comm.set_algo("ring")
    idx = torch.arange(n, device=dev, dtype=torch.float32)
    for it, (start, cnt, ctas) in enumerate([(0, n, 64), (64, 4096, 8), (1 << 20, (1 << 20) + 64 * 3, 200), (0, n, 1),
                                             (n - 64, 64, 16)] * 3):
        comm.set_algo("own" if peer else "ring", ctas)
        _lib.load().b200ddp_comm_set_cta_shape(comm.handle, it % 2)       # both CTA shapes of the kernel
        buf.copy_((idx % 1000) + 7.0 * (rank + 1) + it)                 # small integers: sums are exact in fp32
        want = buf.clone()
        seg = want[start:start + cnt].clone()
        dist.all_reduce(seg)
        want[start:start + cnt] = seg / world
        torch.cuda.synchronize()
        comm.bucket_reduce_launch(buf[start:start + cnt].data_ptr(), cnt, 1.0 / world, dev)
        comm.bucket_reduce_join(dev)
        torch.cuda.synchronize()
        assert torch.equal(buf, want), (it, start, cnt, float((buf - want).abs().max()))
    # random data: equal to NCCL's sum up to fp32 summation order, and bit-identical on all ranks
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    buf.copy_(torch.randn(n, device=dev, generator=g))
    ref = buf.clone()
    dist.all_reduce(ref)
    ref /= world
    comm.bucket_reduce_launch(buf.data_ptr(), n, 1.0 / world, dev)
    comm.bucket_reduce_join(dev)
    torch.cuda.synchronize()
    assert torch.allclose(buf, ref, rtol=1e-6, atol=1e-6)
    mine = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(mine, buf)
    assert all(torch.equal(mine[0], t) for t in mine)
    with open(os.path.join(out_dir, f"arena_mode_rank{rank}.txt"), "w") as fh:
        fh.write(f"{comm.arena_mode} {int(peer)}")


def case_fused_adam_in_backward_is_bitwise(rank, world, out_dir, dev):
    """f4: per-bucket Adam launched as each bucket's all-reduce completes == the one-launch step after backward."""
    import b200ddp
    outs = []
    for overlap in (True, False):
        model = _mlp(seed=3).to(dev)
        ddp = b200ddp.DistributedDataParallel(model)
        opt = b200ddp.optim.Adam(ddp.parameters(), lr=1e-2, overlap_backward=overlap)
        for s in range(4):
            x, y = _data(rank, s)
            opt.zero_grad()
            nn.functional.cross_entropy(ddp(x.to(dev)), y.to(dev)).backward()
            if overlap and dev.type == "cuda":
                assert ddp.reducer.fused_applied
            opt.step()
        assert float(opt.state[next(iter(model.parameters()))]["step"]) == 4.0
        outs.append(torch.cat([p.detach().reshape(-1) for p in model.parameters()]).clone())
    assert torch.equal(outs[0], outs[1])
    ps = _gather_params(model, world)
    assert all(torch.equal(ps[0], t) for t in ps)


def case_small_collectives(rank, world, out_dir, dev):
    """B1: the SyncBatchNorm-sized all-gather / rank-ordered sum kernel over peer memory, 200 calls in a row with
    changing sizes (flag epochs, scratch parity), against torch.distributed."""
    from b200ddp.comm import get_comm
    os.environ["B200DDP_SMALL_COLLECTIVES"] = "own"          # the peer kernel at every world size (default: auto)
    comm = get_comm(dev, None)
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    used = 0
    for it in range(200):
        n = [1, 129, 513, 2048, 769][it % 5]
        send = torch.randn(n, device=dev, generator=g)
        if it % 2 == 0:
            out = torch.empty(world * n, device=dev)
            ok = comm.small_collective(send, out, sum_mode=False)
            ref = [torch.empty_like(send) for _ in range(world)]
            dist.all_gather(ref, send)
            if ok:
                assert torch.equal(out, torch.cat(ref)), it
        else:
            out = torch.empty(n, device=dev)
            ok = comm.small_collective(send, out, sum_mode=True, scale=0.5)
            ref = send.clone()
            dist.all_reduce(ref)
            if ok:
                assert torch.allclose(out, ref * 0.5, rtol=1e-6, atol=1e-6), it
                box = [torch.empty_like(out) for _ in range(world)]
                dist.all_gather(box, out)
                assert all(torch.equal(box[0], t) for t in box)          # rank-ordered sum: same bits everywhere
        used += int(ok)
    with open(os.path.join(out_dir, f"small_rank{rank}.txt"), "w") as fh:
        fh.write(str(used))


def case_metrics_one_collective(rank, world, out_dir, dev):
    """f1: all_reduce_metrics = the reference's five scalar all-reduces (torch.py:198-204) as one collective, in place."""
    import b200ddp
    a = torch.tensor([1.5 + rank], device=dev)
    b = torch.tensor([10.0 * (rank + 1)], device=dev)
    c = torch.tensor([3.0], device=dev)
    out = b200ddp.all_reduce_metrics(a, b, c)
    assert out[0] is a and float(a) == sum(1.5 + r for r in range(world))
    assert float(b) == sum(10.0 * (r + 1) for r in range(world)) and float(c) == 3.0 * world
