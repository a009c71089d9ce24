"""Host-side logic that runs without a GPU: the sampler, the native bucket planner, the C-ABI surface,
arenas, state-dict layout, the Accelerator shim's batch sharding, and Adam's plumbing arithmetic."""
import ctypes
import hashlib
import json
import os
import re

import numpy as np
import pytest
import torch
from torch import nn

import b200ddp
from b200ddp import _lib
from b200ddp.accelerate import Accelerator, BatchSamplerShard
from b200ddp.arena import FlatArena, arena_of
from b200ddp.reducer import plan_buckets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _Len:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


# ---- C ABI ------------------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "b200ddp.h")).read()
    declared = sorted(set(re.findall(r"\b(b200ddp_\w+)\s*\(", header)))
    assert len(declared) >= 35
    lib = ctypes.CDLL(_lib.lib_path())
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b200ddp.h but not exported"
    # and the Python binding covers exactly the declared surface
    assert sorted(_lib._PROTOS) == declared


def test_library_loads_and_reports_errors_without_gpu():
    L = _lib.load()
    assert L.b200ddp_version() >= 100
    rc = L.b200ddp_plan_buckets(None, None, -1, None, 0, None, None)
    assert rc == 1 and b"plan_buckets" in L.b200ddp_last_error()


def test_kernel_entry_points_refuse_cpu_tensors():
    x = torch.randn(4, 8)
    with pytest.raises(_lib.B200Error):
        b200ddp.nn.Linear(8, 4)(x)
    with pytest.raises(_lib.B200Error):
        b200ddp.CrossEntropyLoss()(torch.randn(4, 10), torch.zeros(4, dtype=torch.int64))


# ---- S1 sampler ---------------------------------------------------------------------------------------------
def test_sampler_bit_exact_against_golden(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "sampler.json")))
    for c in g["cases"]:
        s = b200ddp.DistributedSampler(_Len(c["n"]), num_replicas=c["world"], rank=c["rank"],
                                       shuffle=c.get("shuffle", True), seed=c["seed"], drop_last=c["drop_last"])
        s.set_epoch(c["epoch"])
        idx = np.array(list(s), dtype=np.int64)
        assert len(s) == c["len"] == len(idx)
        assert idx[:6].tolist() == c["first"]
        assert hashlib.sha256(idx.tobytes()).hexdigest()[:16] == c["sha16"]


@pytest.mark.parametrize("n,ws,drop_last", [(50000, 8, False), (101, 4, True), (7, 3, False), (5, 8, False)])
def test_sampler_equals_torch_sampler(n, ws, drop_last):
    from torch.utils.data import DistributedSampler as TorchSampler
    for epoch in (0, 3):
        seen = []
        for r in range(ws):
            a = b200ddp.DistributedSampler(_Len(n), num_replicas=ws, rank=r, drop_last=drop_last, seed=5)
            b = TorchSampler(_Len(n), num_replicas=ws, rank=r, drop_last=drop_last, seed=5)
            a.set_epoch(epoch)
            b.set_epoch(epoch)
            assert list(a) == list(b) and len(a) == len(b)
            seen += list(a)
        if not drop_last and n >= ws:
            assert set(seen) == set(range(n))       # shards cover the dataset


def test_sampler_set_epoch_changes_order_and_validates_rank():
    s = b200ddp.DistributedSampler(_Len(1000), num_replicas=2, rank=0)
    e0 = list(s)
    assert list(s) == e0                  # without set_epoch every epoch repeats (README pitfall)
    s.set_epoch(1)
    assert list(s) != e0
    with pytest.raises(ValueError):
        b200ddp.DistributedSampler(_Len(10), num_replicas=2, rank=2)
    with pytest.raises(RuntimeError):
        b200ddp.DistributedSampler(_Len(10))     # no process group


def test_sampler_works_inside_dataloader():
    ds = torch.utils.data.TensorDataset(torch.arange(100).float())
    s = b200ddp.DistributedSampler(ds, num_replicas=4, rank=1)
    dl = torch.utils.data.DataLoader(ds, batch_size=8, sampler=s)
    got = torch.cat([b[0] for b in dl]).long().tolist()
    assert got == list(s) and len(dl) == 4


# ---- R1 bucket planner (native) ----------------------------------------------------------------------------
def test_native_planner_matches_upstream_golden(golden_dir):
    for case in json.load(open(os.path.join(golden_dir, "buckets.json"))):
        assert plan_buckets(case["nbytes"], case["caps"]) == case["buckets"]


def test_native_planner_matches_torch_binding_random():
    import torch.distributed as dist
    g = torch.Generator().manual_seed(9)
    for _ in range(5):
        sizes = torch.randint(1, 2_000_000, (int(torch.randint(1, 60, (1,), generator=g)),), generator=g).tolist()
        tensors = [torch.empty(s) for s in sizes]
        idx, _ = dist._compute_bucket_assignment_by_size(tensors, [1 << 20, 25 << 20])
        assert plan_buckets([s * 4 for s in sizes], [1 << 20, 25 << 20]) == [list(b) for b in idx]
    assert plan_buckets([], [1 << 20]) == []


# ---- arenas / wrapper layout --------------------------------------------------------------------------------
def _tiny():
    torch.manual_seed(0)
    return nn.Sequential(nn.Linear(10, 7), nn.ReLU(), nn.Linear(7, 3))


def test_arena_keeps_parameter_identity_and_values():
    m = _tiny()
    before = {k: v.clone() for k, v in m.state_dict().items()}
    params = list(m.parameters())
    a = FlatArena(list(reversed(params)))
    assert [id(p) for p in m.parameters()] == [id(p) for p in params]
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k])
    for i, p in enumerate(a.params):
        assert p.data_ptr() == a.p_flat.data_ptr() + 4 * a.offsets[i]
        assert a.offsets[i] % 64 == 0
    assert arena_of(params) is a and arena_of(params[:2]) is None
    a.p_flat.zero_()
    assert all(float(p.detach().abs().sum()) == 0 for p in m.parameters())


def test_ddp_wrapper_state_dict_layout_and_checkpoint(tmp_path):
    import torchvision
    torch.manual_seed(1)
    net = torchvision.models.alexnet(weights=None)
    net.classifier[6] = nn.Linear(4096, 10)
    import b200ddp.distributed as D
    D._warned_plumbing = False
    with pytest.warns(UserWarning, match='plumbing'):
        ddp = b200ddp.DistributedDataParallel(net)
    sd = ddp.state_dict()
    assert len(sd) == 16 and list(sd)[0] == "module.features.0.weight"
    assert all(k.startswith("module.") and v.dtype == torch.float32 for k, v in sd.items())
    assert ddp.bucket_sizes() == [67289128, 151011328, 9878784]
    path = tmp_path / "ckpt_0.pt"
    torch.save(sd, path)
    loaded = torch.load(path, map_location="cpu")
    assert list(loaded) == list(sd)
    fresh = torchvision.models.alexnet(weights=None)
    fresh.classifier[6] = nn.Linear(4096, 10)
    fresh.load_state_dict({k[len("module."):]: v for k, v in loaded.items()})
    for (k, a), (_, b) in zip(fresh.state_dict().items(), net.state_dict().items()):
        assert torch.equal(a, b), k
    ddp.load_state_dict(loaded)
    assert [id(p) for p in ddp.parameters()] == [id(p) for p in net.parameters()]
    assert ddp.module is net
    ddp.eval()
    assert not net.training
    ddp.train()
    assert net.training


def test_our_alexnet_is_state_dict_compatible_with_torchvision():
    import torchvision
    from b200ddp.models import AlexNet
    tv = torchvision.models.alexnet(weights=None)
    tv.classifier[6] = nn.Linear(4096, 10)
    ours = AlexNet()
    assert [(k, tuple(v.shape)) for k, v in ours.state_dict().items()] == \
           [(k, tuple(v.shape)) for k, v in tv.state_dict().items()]
    ours.load_state_dict(tv.state_dict())
    assert sum(p.numel() for p in ours.parameters()) == 57_044_810
    wide = AlexNet(hidden=8192)
    assert sum(p.numel() for p in wide.parameters()) == 145_195_850 or sum(p.numel() for p in wide.parameters()) > 1.4e8


def test_single_process_backward_on_cpu_fills_arena_grads():
    m = _tiny()
    ddp = b200ddp.DistributedDataParallel(m)
    x = torch.randn(5, 10)
    ref = _tiny()
    ref(x).sum().backward()
    ddp(x).sum().backward()
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad)
        i = ddp.arena.index_of(p)
        assert ddp.arena.grad_in_arena(i)
    assert ddp.reducer.first_ready_order is not None


def test_convert_sync_batchnorm_shares_parameters_and_buffers():
    net = nn.Sequential(nn.Conv2d(3, 4, 3), nn.BatchNorm2d(4), nn.ReLU(), nn.Sequential(nn.BatchNorm2d(4)))
    net.eval()
    bn0 = net[1]
    out = b200ddp.SyncBatchNorm.convert_sync_batchnorm(net)
    assert isinstance(out[1], b200ddp.SyncBatchNorm) and isinstance(out[3][0], b200ddp.SyncBatchNorm)
    assert out[1].weight is bn0.weight and out[1].bias is bn0.bias
    assert out[1].running_mean is bn0.running_mean and out[1].num_batches_tracked is bn0.num_batches_tracked
    assert out[1].training is False
    assert set(out[1].state_dict()) == {"weight", "bias", "running_mean", "running_var", "num_batches_tracked"}
    with pytest.raises(ValueError):
        out[1](torch.randn(2, 4, 5, 5))          # CPU input raises, as upstream does


# ---- O1 Adam plumbing arithmetic ---------------------------------------------------------------------------
def test_adam_host_arithmetic_matches_torch_adam(golden_dir):
    g = np.load(os.path.join(golden_dir, "adam.npz"))
    p = nn.Parameter(torch.from_numpy(g["p0"].copy()))
    opt = b200ddp.optim.Adam([p], lr=1e-3)
    for t in (1, 2, 3):
        p.grad = torch.from_numpy(g[f"g{t}"].copy())
        opt.step()
        np.testing.assert_allclose(p.detach().numpy(), g[f"p_wd0.0_t{t}"], rtol=2e-6, atol=2e-7)
    sd = opt.state_dict()
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and float(sd["state"][0]["step"]) == 3.0
    assert sd["param_groups"][0]["lr"] == 1e-3 and sd["param_groups"][0]["betas"] == (0.9, 0.999)
    opt2 = b200ddp.optim.Adam([p], lr=1e-3)
    opt2.load_state_dict(sd)
    assert float(opt2.state[p]["step"]) == 3.0
    opt.zero_grad()
    assert p.grad is None


# ---- Accelerator shim ------------------------------------------------------------------------------------------
def test_batch_sampler_shard_partitions_batches():
    from torch.utils.data import BatchSampler, SequentialSampler
    bs = BatchSampler(SequentialSampler(range(22)), 4, False)
    shards = [list(BatchSamplerShard(bs, 2, r)) for r in range(2)]
    assert shards[0] == [[0, 1, 2, 3], [8, 9, 10, 11], [16, 17, 18, 19]]
    assert shards[1] == [[4, 5, 6, 7], [12, 13, 14, 15], [20, 21, 0, 1]]
    assert [len(BatchSamplerShard(bs, 2, r)) for r in range(2)] == [3, 3]
    bs = BatchSampler(SequentialSampler(range(24)), 4, False)
    shards = [list(BatchSamplerShard(bs, 3, r)) for r in range(3)]
    assert sorted(i for s in shards for b in s for i in b) == list(range(24))
    assert all(len(s) == 2 for s in shards)


def test_accelerator_single_process(tmp_path, monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    acc = Accelerator(cpu=True)
    assert acc.device.type == "cpu" and acc.is_local_main_process and acc.num_processes == 1
    m = _tiny()
    opt = b200ddp.optim.Adam(m.parameters(), lr=1e-3)
    ds = torch.utils.data.TensorDataset(torch.randn(16, 10), torch.randint(0, 3, (16,)))
    dl = torch.utils.data.DataLoader(ds, batch_size=4)
    m2, opt2, dl2 = acc.prepare(m, opt, dl)
    assert m2 is m and opt2 is opt and len(dl2) == 4
    for x, y in dl2:
        opt2.zero_grad()
        loss = nn.functional.cross_entropy(m2(x), y)
        acc.backward(loss)
        opt2.step()
    acc.wait_for_everyone()
    acc.save_model(m2, str(tmp_path))
    from safetensors.torch import load_file
    sd = load_file(str(tmp_path / "model.safetensors"))
    assert set(sd) == set(m.state_dict()) and not any(k.startswith("module.") for k in sd)


def test_sm_budget_drives_the_one_wave_grid_plans():
    """b200ddp_set_sm_budget: the wgrad planner splits the pixel range so that tiles x splits fits the SMs it may
    count on; the DDP wrapper lowers the budget by the number of NCCL channels. Host-only
    entry points, no GPU needed: the workspace query exposes the split count."""
    L = _lib.load()
    prev = L.b200ddp_set_sm_budget(148)
    try:
        # conv2 of the toy model: Cin=64, 27x27, Cout=192, 5x5: K = 1600 -> 7 row tiles of 256, one channel tile
        full = int(L.b200ddp_conv2d_wgrad_workspace_bytes(128, 64, 27, 27, 192, 5, 5, 1, 2))
        assert full == (148 // 7) * 192 * 1600 * 4
        assert L.b200ddp_set_sm_budget(132) == 148
        less = int(L.b200ddp_conv2d_wgrad_workspace_bytes(128, 64, 27, 27, 192, 5, 5, 1, 2))
        assert less == (132 // 7) * 192 * 1600 * 4
        assert L.b200ddp_set_sm_budget(0) == 132          # 0 = all SMs again
        assert int(L.b200ddp_conv2d_wgrad_workspace_bytes(128, 64, 27, 27, 192, 5, 5, 1, 2)) == full
        assert L.b200ddp_set_sm_budget(4) == 148          # clamped to a sane minimum
        assert L.b200ddp_set_sm_budget(148) == 16
    finally:
        L.b200ddp_set_sm_budget(prev)


def test_space_to_depth_geometry_and_opt_in(monkeypatch):
    from b200ddp import nn as bnn
    # conv1 of the toy model: 224 -> 55 outputs, 11x11 stride 4 becomes 3x3 over a 57x57 grid
    assert bnn._s2d_geometry(224, 224, 11, 11, 4, 2) == (55, 55, 3, 3, 57, 57)
    monkeypatch.delenv("B200DDP_CONV1_S2D", raising=False)
    assert not bnn._use_s2d(3, 11, 11, 4)                 # measured: no net gain, so off unless asked for
    monkeypatch.setenv("B200DDP_CONV1_S2D", "1")
    assert bnn._use_s2d(3, 11, 11, 4)
    assert not bnn._use_s2d(64, 5, 5, 1) and not bnn._use_s2d(16, 11, 11, 4)     # stride 1 / too many channels


# ---- round 2 host logic ----------------------------------------------------------------------------------------
def test_convert_sync_batchnorm_folds_relu_and_leaves_no_eager_activation():
    import b200ddp
    from b200ddp import nn as bnn
    from b200ddp.models import AlexNet
    m = b200ddp.SyncBatchNorm.convert_sync_batchnorm(AlexNet(bn=True))
    kinds = [type(l).__name__ for l in m.features]
    assert "ReLU" not in kinds and "MaxPool2d" not in kinds and kinds.count("SyncBatchNorm") == 5
    assert all(l.fuse_relu for l in m.features if isinstance(l, bnn.SyncBatchNorm))
    assert all(not l.relu for l in m.features if isinstance(l, bnn.ReLUMaxPool2d))      # the ReLU moved into the BN kernels
    again = b200ddp.SyncBatchNorm.convert_sync_batchnorm(m)                              # idempotent
    assert [type(l).__name__ for l in again.features] == kinds


def test_checkpoint_round_trip_on_cpu_keeps_reference_file_format(tmp_path):
    """AsyncCheckpointer / load_checkpoint on a CPU model (no arena, no side stream): the primary file is what
    torch.save(model.state_dict()) of the reference writes (:217-221); the side-car restores Adam and the sampler."""
    import b200ddp
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))
    opt = b200ddp.optim.Adam(net.parameters(), lr=3e-3)
    ds = torch.utils.data.TensorDataset(torch.arange(20))
    sampler = b200ddp.DistributedSampler(ds, num_replicas=2, rank=1, shuffle=True, seed=5)
    sampler.set_epoch(7)
    x, y = torch.randn(4, 6), torch.tensor([0, 1, 2, 0])
    for _ in range(2):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(net(x), y).backward()
        opt.step()
    path = str(tmp_path / "ckpt_5.pt")
    ck = b200ddp.AsyncCheckpointer(net, optimizer=opt, sampler=sampler)
    ck.save(path, epoch=5, extra={"note": "x"})
    ck.wait()
    sd = torch.load(path, map_location="cpu")
    assert list(sd) == list(net.state_dict()) and all(torch.equal(sd[k], v) for k, v in net.state_dict().items())
    net2 = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))
    opt2 = b200ddp.optim.Adam(net2.parameters(), lr=1.0)
    sampler2 = b200ddp.DistributedSampler(ds, num_replicas=2, rank=1, shuffle=True, seed=0)
    info = b200ddp.load_checkpoint(path, net2, opt2, sampler2)
    assert info == {"epoch": 5, "extra": {"note": "x"}} and list(sampler2) == list(sampler)
    assert opt2.param_groups[0]["lr"] == 3e-3
    for a, b in ((net, opt), (net2, opt2)):
        b.zero_grad()
        torch.nn.functional.cross_entropy(a(x), y).backward()
        b.step()
    for p, q in zip(net.parameters(), net2.parameters()):
        assert torch.equal(p, q)


def test_engine_applies_only_to_the_standard_cuda_tf32_model():
    from b200ddp import engine
    from b200ddp.models import AlexNet
    x = torch.zeros(2, 3, 224, 224)
    assert not engine.applicable(AlexNet(), x)                      # CPU tensors: never
    assert engine.enabled()


def test_gpu_transform_coefficient_table_shape():
    import ctypes
    L = _lib.load()
    tab = (ctypes.c_int * (5 * 224))()
    assert L.b200ddp_resize_coeffs(32, 224, tab) == 0
    t = list(tab)
    assert all(1 <= c <= 3 for c in t[224:448]) and t[0] == 0 and t[223] <= 31
    assert all(sum(t[448 + 3 * i:451 + 3 * i]) in range((1 << 22) - 2, (1 << 22) + 3) for i in range(224))   # weights sum to 1.0
