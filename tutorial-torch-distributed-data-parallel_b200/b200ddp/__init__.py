"""b200ddp — B200-native DDP training step for the toy model of
annalena-k/tutorial-torch-distributed-data-parallel, behind the surface the tutorial's scripts use.

    from b200ddp import DistributedDataParallel as DDP, DistributedSampler    # multi-GPU-training-torch.py:17,19
    from b200ddp import optim, nn as bnn                                      # optim.Adam, bnn.CrossEntropyLoss
    from b200ddp.accelerate import Accelerator                                # multi-GPU-training-accelerate.py:19
    from b200ddp.models import load_model                                     # data_and_toy_model.py:41-45
"""
import os as _os

# The gradient all-reduce runs concurrently with the convolution backward; NCCL's persistent kernel owns one SM per
# channel for its whole duration. A fixed channel count and one-wave grids planned for the remaining SMs
# (b200ddp_set_sm_budget, see comm.py) keep both sides at full speed. Measured on 8 x B200 (ring, PreMulSum, 228 MB per
# step): 8 channels 3.89 ms/step, 12: 2.94, 16: 2.88, 24: 2.92, 32: 2.95 (one GPU: 2.55); on 2 GPUs 8 and 16 tie.
# Must be in the environment before the first NCCL communicator of the process is created.
_os.environ.setdefault("NCCL_MAX_NCHANNELS", "16")

from . import _lib, checkpoint, data, nn, optim  # noqa: F401,E402
from .arena import FlatArena  # noqa: F401
from .checkpoint import AsyncCheckpointer, load_checkpoint  # noqa: F401
from .distributed import DistributedDataParallel  # noqa: F401
from .evaluate import all_reduce_metrics, evaluate  # noqa: F401
from .nn import CrossEntropyLoss, SyncBatchNorm  # noqa: F401
from .optim import Adam  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401

convert_sync_batchnorm = SyncBatchNorm.convert_sync_batchnorm


def reserve_memory_pool(gigabytes: float = 24.0, device=None) -> None:
    """Have torch's caching allocator hold one block of `gigabytes` from the start, which it then carves up for the
    step's activations and workspaces. Without it the allocator grows by cudaMalloc calls during the first few hundred
    steps, and a cudaMalloc of a multi-hundred-MB segment stalls the launching thread for 30-65 ms (measured: single
    steps of 31 and 65 ms among 2.34 ms ones). 180 GB of HBM per B200 make the reservation cheap."""
    import torch
    if gigabytes <= 0 or not torch.cuda.is_available():
        return
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    blk = torch.empty(int(gigabytes * (1 << 30)), dtype=torch.uint8, device=dev)
    del blk

__all__ = ["reserve_memory_pool", "DistributedDataParallel", "DistributedSampler", "SyncBatchNorm", "convert_sync_batchnorm", "Adam",
           "CrossEntropyLoss", "FlatArena", "nn", "optim", "evaluate", "all_reduce_metrics", "data", "checkpoint",
           "AsyncCheckpointer", "load_checkpoint"]
__version__ = "0.1.0"
