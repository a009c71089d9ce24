"""Multi-GPU (NCCL over NVLink) parity: the same cases as tests/test_gloo_ws2.py on 2 B200s, plus the toy
model against stock torch DistributedDataParallel. Skipped on boxes with fewer than 2 GPUs."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from tests import _workers
from tests.test_gloo_ws2 import CASES, _free_port

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _spawn(fn, case, tmp_path, world=2):
    mp.spawn(fn, args=(world, _free_port(), case, str(tmp_path), "nccl"), nprocs=world, join=True)
    assert os.path.exists(os.path.join(str(tmp_path), case + ".ok"))


@pytest.mark.parametrize("case", CASES)
def test_nccl(case, tmp_path):
    _spawn(_workers.run, case, tmp_path)


def test_alexnet_ddp_matches_stock_ddp(tmp_path):
    _spawn(_workers.run, "alexnet_vs_stock_ddp", tmp_path)


def test_syncbn_two_ranks_matches_global_batch(tmp_path):
    _spawn(_workers.run, "syncbn_two_ranks", tmp_path)


def test_own_allreduce_kernel_known_values(tmp_path):
    _spawn(_workers.run, "own_allreduce_known_values", tmp_path)
    print("arena mode / peer path:", open(os.path.join(str(tmp_path), "arena_mode_rank0.txt")).read())


def test_small_collectives_over_peer_memory(tmp_path):
    _spawn(_workers.run, "small_collectives", tmp_path)
    print("small collectives served by the peer kernel:", open(os.path.join(str(tmp_path), "small_rank0.txt")).read(), "of 200")
