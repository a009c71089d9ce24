"""world_size-2 tests of the N>1 host logic on CPU (gloo): wrapper, reducer, sampler, optimizer, shim.
The same cases run over NCCL on 2 GPUs in tests/test_gpu_ddp.py."""
import os
import socket

import pytest
import torch.multiprocessing as mp

from tests import _workers


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(case, tmp_path, world=2, backend="gloo"):
    mp.spawn(_workers.run, args=(world, _free_port(), case, str(tmp_path), backend), nprocs=world, join=True)
    assert os.path.exists(os.path.join(str(tmp_path), case + ".ok"))


CASES = ["mean_gradient", "parity_with_stock_ddp", "clip_value_before_aggregation", "clip_norm_before_aggregation",
         "no_sync_accumulation", "fault_containment", "checkpoint_rank0", "accelerator", "fused_adam_in_backward_is_bitwise", "metrics_one_collective"]


@pytest.mark.parametrize("case", CASES)
def test_gloo(case, tmp_path):
    _spawn(case, tmp_path)


def test_gloo_plumbing_alexnet_config1(tmp_path):
    """BASELINE.json configs[0]: toy model via the reference's loop on CPU/gloo world_size=2."""
    _spawn("plumbing_alexnet", tmp_path)
