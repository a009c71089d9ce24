"""Generate tests/golden/* from the REAL upstream routines and the reference's own train() loop.

Run in the build container only (it imports /root/reference, which does not exist on the GPU box):

    python oracle/make_golden.py

Every fixture records inputs (or the seeds that regenerate them) and the outputs of the
third-party implementation the reference calls: torch.utils.data.DistributedSampler,
dist._compute_bucket_assignment_by_size, torch.optim.Adam, torch.nn.utils.clip_grad_*, torch
functional ops under autograd, and the reference's `train()` (multi-GPU-training-torch.py:104-133)
driven through the five patches of SURVEY.md 8(c).
"""
from __future__ import annotations

import hashlib
import importlib.util
import json
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import reference_loop as rl  # noqa: E402


class _Len:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def sha16(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.int64).tobytes()).hexdigest()[:16]


def gen_sampler():
    from torch.utils.data import DistributedSampler
    cases = []
    grid = [(50000, ws, e, r, False) for ws in (1, 2, 4, 8) for e in (0, 1) for r in sorted({0, ws - 1})]
    grid += [(10000, ws, e, r, False) for ws in (2, 8) for e in (0, 1) for r in (0, ws - 1)]
    grid += [(10, 4, 0, r, False) for r in range(4)] + [(10, 4, 3, r, True) for r in range(4)]
    grid += [(3, 8, 0, r, False) for r in (0, 5, 7)] + [(0 + 17, 5, 2, 4, True), (1, 1, 0, 0, False)]
    for n, ws, e, r, dl in grid:
        s = DistributedSampler(_Len(n), num_replicas=ws, rank=r, shuffle=True, seed=0, drop_last=dl)
        s.set_epoch(e)
        idx = np.array(list(s), dtype=np.int64)
        cases.append({"n": n, "world": ws, "epoch": e, "rank": r, "drop_last": dl, "seed": 0, "len": len(s),
                      "first": idx[:6].tolist(), "sha16": sha16(idx)})
    # unshuffled
    for n, ws, r in [(10, 4, 1), (50000, 8, 3)]:
        s = DistributedSampler(_Len(n), num_replicas=ws, rank=r, shuffle=False)
        idx = np.array(list(s), dtype=np.int64)
        cases.append({"n": n, "world": ws, "epoch": 0, "rank": r, "drop_last": False, "seed": 0, "shuffle": False,
                      "len": len(s), "first": idx[:6].tolist(), "sha16": sha16(idx)})
    # per-rank batch counts at bs 128
    counts = {}
    for ws in (1, 2, 4, 8):
        s = DistributedSampler(_Len(50000), num_replicas=ws, rank=0)
        n = len(s)
        counts[str(ws)] = {"batches": (n + 127) // 128, "last": n - (n - 1) // 128 * 128}
    json.dump({"cases": cases, "batch_counts_bs128": counts}, open(os.path.join(GOLD, "sampler.json"), "w"), indent=1)


def gen_buckets():
    out = []
    for hidden in (4096, 8192):
        m = rl.make_model(seed=0, hidden=hidden)
        params = list(reversed(list(m.named_parameters())))
        tensors = [p for _, p in params]
        idx, _ = dist._compute_bucket_assignment_by_size(tensors, [1 << 20, 25 << 20])
        out.append({"hidden": hidden, "names": [n for n, _ in params], "nbytes": [p.numel() * 4 for p in tensors],
                    "caps": [1 << 20, 25 << 20], "buckets": [list(b) for b in idx]})
    g = torch.Generator().manual_seed(3)
    sizes = torch.randint(1, 3_000_000, (40,), generator=g).tolist()
    tensors = [torch.empty(s) for s in sizes]
    for caps in ([1 << 20, 25 << 20], [4 << 20], [1 << 10, 1 << 20, 8 << 20]):
        idx, _ = dist._compute_bucket_assignment_by_size(tensors, caps)
        out.append({"hidden": None, "nbytes": [s * 4 for s in sizes], "caps": caps, "buckets": [list(b) for b in idx]})
    json.dump(out, open(os.path.join(GOLD, "buckets.json"), "w"))


def gen_adam():
    g = torch.Generator().manual_seed(11)
    n = 1003
    p0 = torch.randn(n, generator=g)
    grads = [torch.randn(n, generator=g) * s for s in (1.0, 0.1, 3.0)]
    out = {"p0": p0.numpy().copy()}
    for wd in (0.0, 0.01):
        p = nn.Parameter(p0.clone())
        opt = torch.optim.Adam([p], lr=1e-3, weight_decay=wd, foreach=False)
        for t, gr in enumerate(grads, 1):
            p.grad = gr.clone()
            opt.step()
            st = opt.state[p]
            tag = f"wd{wd}_t{t}"
            out[f"p_{tag}"] = p.detach().numpy().copy()
            out[f"m_{tag}"] = st["exp_avg"].numpy().copy()
            out[f"v_{tag}"] = st["exp_avg_sq"].numpy().copy()
    for t, gr in enumerate(grads, 1):
        out[f"g{t}"] = gr.numpy().copy()
    np.savez_compressed(os.path.join(GOLD, "adam.npz"), **out)


def gen_clip():
    g = torch.Generator().manual_seed(5)
    shapes = [(7, 5), (33,), (4, 3, 2), (1,)]
    grads = [torch.randn(s, generator=g) * 2 for s in shapes]
    out = {f"g{i}": t.numpy().copy() for i, t in enumerate(grads)}
    for mn in (1.0, 100.0):
        ps = [nn.Parameter(torch.zeros(s)) for s in shapes]
        for p, t in zip(ps, grads):
            p.grad = t.clone()
        total = torch.nn.utils.clip_grad_norm_(ps, mn)
        out[f"norm_total_{mn}"] = np.array(float(total))
        for i, p in enumerate(ps):
            out[f"norm_{mn}_g{i}"] = p.grad.numpy().copy()
    ps = [nn.Parameter(torch.zeros(s)) for s in shapes]
    for p, t in zip(ps, grads):
        p.grad = t.clone()
    torch.nn.utils.clip_grad_value_(ps, 0.5)
    for i, p in enumerate(ps):
        out[f"value_0.5_g{i}"] = p.grad.numpy().copy()
    np.savez_compressed(os.path.join(GOLD, "clip.npz"), **out)


def gen_layers():
    g = torch.Generator().manual_seed(21)
    out = {}
    # Linear + ReLU (dropout mask is ours, checked separately against Philox known answers)
    x = torch.randn(8, 36, generator=g, requires_grad=True)
    w = torch.randn(20, 36, generator=g, requires_grad=True)
    b = torch.randn(20, generator=g, requires_grad=True)
    dy = torch.randn(8, 20, generator=g)
    y = F.relu(F.linear(x, w, b))
    y.backward(dy)
    out.update(lin_x=x.detach().numpy(), lin_w=w.detach().numpy(), lin_b=b.detach().numpy(), lin_dy=dy.numpy(),
               lin_y=y.detach().numpy(), lin_dx=x.grad.numpy(), lin_dw=w.grad.numpy(), lin_db=b.grad.numpy())
    # cross entropy
    z = torch.randn(9, 10, generator=g, requires_grad=True)
    lab = torch.randint(0, 10, (9,), generator=g)
    loss = nn.CrossEntropyLoss()(z, lab)
    loss.backward()
    out.update(ce_z=z.detach().numpy(), ce_labels=lab.numpy(), ce_loss=np.array(loss.item(), dtype=np.float32),
               ce_dz=z.grad.numpy())
    # ReLU + MaxPool2d(3,2)
    xp = torch.randn(2, 3, 13, 15, generator=g, requires_grad=True)
    yp = F.max_pool2d(F.relu(xp), 3, 2)
    dyp = torch.randn(yp.shape, generator=g)
    yp.backward(dyp)
    out.update(pool_x=xp.detach().numpy(), pool_y=yp.detach().numpy(), pool_dy=dyp.numpy(), pool_dx=xp.grad.numpy())
    # SyncBN oracle = BatchNorm over the concatenated global batch (2 "ranks")
    C = 5
    xs = [torch.randn(3, C, 4, 6, generator=g) * 2 + 1 for _ in range(2)]
    wbn = torch.randn(C, generator=g).requires_grad_()
    bbn = torch.randn(C, generator=g).requires_grad_()
    rm, rv = torch.zeros(C), torch.ones(C)
    xa = torch.cat(xs).requires_grad_()
    ya = F.batch_norm(xa, rm, rv, wbn, bbn, training=True, momentum=0.1, eps=1e-5)
    yr = F.relu(ya)
    dya = torch.randn(ya.shape, generator=g)
    yr.backward(dya)
    out.update(bn_x0=xs[0].numpy(), bn_x1=xs[1].numpy(), bn_w=wbn.detach().numpy(), bn_b=bbn.detach().numpy(),
               bn_y=yr.detach().numpy(), bn_rm=rm.numpy(), bn_rv=rv.numpy(), bn_dy=dya.numpy(),
               bn_dx=xa.grad.numpy(), bn_gw=wbn.grad.numpy(), bn_gb=bbn.grad.numpy())
    np.savez_compressed(os.path.join(GOLD, "layers.npz"), **{k: np.ascontiguousarray(v) for k, v in out.items()})


# ---- the reference's own loop ------------------------------------------------------------------------------
def _load_reference():
    sys.path.insert(0, REF)
    spec = importlib.util.spec_from_file_location("ref_torch_script", os.path.join(REF, "multi-GPU-training-torch.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _RecordingCriterion:
    """Wraps nn.CrossEntropyLoss and records every loss the reference's train() computes."""

    def __init__(self):
        self.ce = nn.CrossEntropyLoss()
        self.values = []

    def __call__(self, out, lab):
        loss = self.ce(out, lab)
        self.values.append(float(loss.item()))
        return loss


def _ref_worker(rank, world, port, out_dir, n_samples, batch, lr, model_seed, data_seed):
    """SURVEY 8(c) harness: patches 1-5 around the reference's unmodified functions."""
    ref = _load_reference()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    dist.is_nccl_available = lambda: False                       # patch 1
    torch.cuda.set_device = lambda d: None                       # patch 2
    x, y = rl.synthetic_pool(n_samples, seed=data_seed)
    ds = torch.utils.data.TensorDataset(x, y)
    ref.load_datasets = lambda: (ds, ds)                         # patch 3
    torch.set_num_threads(2)
    # setup() hard-codes port 12355 (reference :31); it is fine for one job at a time
    ref.setup(rank, world)
    torch.manual_seed(42)
    ref.set_seed_based_on_rank(rank)
    from torch.utils.data import DataLoader, DistributedSampler
    # setup_dataloaders hard-codes batch 128 / 2 workers (:86-99); restate it with the test batch size
    sampler = DistributedSampler(ds, shuffle=True, num_replicas=world, rank=rank)
    loader = DataLoader(ds, batch_size=batch, sampler=sampler)
    model = rl.make_model(seed=model_seed, dropout=0.0)           # patch 4 (dropout off: mask RNG differs)
    from torch.nn.parallel import DistributedDataParallel as DDP
    ddp = DDP(model)                                             # patch 5
    crit = _RecordingCriterion()
    opt = torch.optim.Adam(ddp.parameters(), lr=lr)
    sampler.set_epoch(0)
    total, n = ref.train(ddp, loader, crit, opt, torch.device("cpu"))    # the reference's own hot loop
    sd = ddp.state_dict()
    torch.save({"losses": crit.values, "total": float(total.item()), "n": float(n.item()),
                "param_sums": {k: float(v.double().sum()) for k, v in sd.items()},
                "param_l2": {k: float(v.double().norm()) for k, v in sd.items()},
                "keys": list(sd.keys())}, os.path.join(out_dir, f"ref_rank{rank}.pt"))
    dist.barrier()
    ref.cleanup()


def gen_trace():
    out = {"recipe": {"model_seed": 1234, "data_seed": 7, "n_samples": 64, "dropout": 0.0,
                      "note": "reference train() on CPU/gloo through the 5 patches of SURVEY 8(c); "
                              "sampler epoch 0, shuffle=True, seed 0"},
           "runs": []}
    for world, batch, lr in [(1, 16, 1e-3), (1, 16, 1e-4), (2, 8, 1e-3), (2, 8, 1e-4)]:
        with tempfile.TemporaryDirectory() as td:
            mp.spawn(_ref_worker, args=(world, 12355, td, 64, batch, lr, 1234, 7), nprocs=world, join=True)
            ref_ranks = [torch.load(os.path.join(td, f"ref_rank{r}.pt")) for r in range(world)]
        with tempfile.TemporaryDirectory() as td:
            mp.spawn(rl.ddp_worker, args=(world, 12361, td, 64, batch, lr, 64 // (world * batch), 1234, 7, 0.0),
                     nprocs=world, join=True)
            port_ranks = [torch.load(os.path.join(td, f"trace_rank{r}.pt")) for r in range(world)]
        for r in range(world):
            a, b = ref_ranks[r]["losses"], port_ranks[r]["losses"]
            assert len(a) == len(b) and max(abs(x - y) for x, y in zip(a, b)) < 2e-3 * max(1.0, max(a)), (a, b)
        out["runs"].append({"world": world, "batch": batch, "lr": lr,
                            "losses": [r["losses"] for r in ref_ranks],
                            "param_sums": ref_ranks[0]["param_sums"], "param_l2": ref_ranks[0]["param_l2"],
                            "keys": ref_ranks[0]["keys"],
                            "restatement_max_abs_loss_diff": max(
                                max(abs(x - y) for x, y in zip(ref_ranks[r]["losses"], port_ranks[r]["losses"]))
                                for r in range(world))})
        print("trace", world, batch, lr, out["runs"][-1]["losses"], flush=True)
    json.dump(out, open(os.path.join(GOLD, "trace.json"), "w"), indent=1)


def gen_eval():
    """The reference's own evaluate() (multi-GPU-training-torch.py:136-153) on CPU: 20 samples in batches of 8 (last one
    ragged), AlexNet-10 seed 1234 in eval mode. Records the hidden activations entering the head, the head's weights,
    logits and labels per batch, and evaluate()'s three totals; asserts the restatement returns the same."""
    ref = _load_reference()
    model = rl.make_model(seed=1234, dropout=0.5)
    x, y = rl.synthetic_pool(20, seed=11)
    model.eval()
    with torch.no_grad():                             # an untrained model predicts one class: make every other label a hit
        pred = model(x).argmax(1)
    y = torch.where(torch.arange(20) % 2 == 0, pred, (pred + 1) % 10)
    ds = torch.utils.data.TensorDataset(x, y)
    loader = torch.utils.data.DataLoader(ds, batch_size=8)
    hid, logits = [], []
    def _grab(m, i, o):
        hid.append(i[0].detach().numpy().copy())
        logits.append(o.detach().numpy().copy())
    h1 = model.classifier[6].register_forward_hook(_grab)
    crit = nn.CrossEntropyLoss()
    tl, c, n = ref.evaluate(model, loader, crit, torch.device("cpu"))
    h1.remove()
    tl2, c2, n2 = rl.evaluate_epoch(model, loader, crit, torch.device("cpu"))
    assert float(tl) == float(tl2) and float(c) == float(c2) and float(n) == float(n2)
    np.savez_compressed(os.path.join(GOLD, "eval.npz"),
                        hidden=np.concatenate(hid), logits=np.concatenate(logits), labels=y.numpy(),
                        w=model.classifier[6].weight.detach().numpy(), b=model.classifier[6].bias.detach().numpy(),
                        batch=np.int64(8), total_test_loss=np.float32(tl.item()), correct=np.float32(c.item()),
                        total=np.float32(n.item()))
    print("eval", float(tl), float(c), float(n))


def gen_input_pipeline():
    """The real transform of reference data_and_toy_model.py:13-36 (torchvision 0.26 + Pillow) on CIFAR-shaped uint8
    images: the uint8 result of Resize(224), the first 16 rows of the final float tensor, and a sha256 of every full
    float output (the GPU kernel is bit-exact, so a hash pins the full size)."""
    import torchvision.transforms as T
    from PIL import Image
    mean, std = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)
    rng = np.random.default_rng(32)
    imgs = rng.integers(0, 256, (4, 32, 32, 3), dtype=np.uint8)
    imgs[2] = 255                                     # saturated image
    imgs[3, :, :16] = 0                               # hard vertical edge
    flips = np.array([0, 1, 0, 1], np.uint8)
    tail = T.Compose([T.ToTensor(), T.Normalize(mean, std)])
    resized, rows, hashes = [], [], []
    for i in range(4):
        im = T.Resize(224)(Image.fromarray(imgs[i]))
        resized.append(np.asarray(im).copy())
        if flips[i]:
            im = T.functional.hflip(im)                # RandomHorizontalFlip's action when it fires
        out = tail(im).numpy()
        rows.append(out[:, :16, :].copy())
        hashes.append(hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest())
    import PIL
    import torchvision
    np.savez_compressed(os.path.join(GOLD, "input_pipeline.npz"), images=imgs, flips=flips, resized_u8=np.stack(resized),
                        first_rows=np.stack(rows), sha256=np.array(hashes),
                        versions=np.array([f"torchvision {torchvision.__version__}", f"Pillow {PIL.__version__}"]))


def gen_philox():
    kat = [{"ctr": [0, 0, 0, 0], "key": [0, 0], "out": [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]},
           {"ctr": [0xffffffff] * 4, "key": [0xffffffff] * 2, "out": [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]},
           {"ctr": [0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], "key": [0xa4093822, 0x299f31d0],
            "out": [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]}]
    json.dump({"source": "Random123 v1.14 kat_vectors, philox4x32 10 rounds (Salmon et al., SC'11)", "kat": kat},
              open(os.path.join(GOLD, "philox.json"), "w"), indent=1)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or ["sampler", "buckets", "adam", "clip", "layers", "philox", "trace", "eval", "input_pipeline"]
    for w in which:
        globals()["gen_" + w]()
        print("wrote", w, flush=True)
