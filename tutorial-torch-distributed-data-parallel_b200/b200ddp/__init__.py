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

from . import _lib, nn, optim  # noqa: F401,E402
from .arena import FlatArena  # noqa: F401
from .distributed import DistributedDataParallel  # noqa: F401
from .nn import CrossEntropyLoss, SyncBatchNorm  # noqa: F401
from .optim import Adam  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401

convert_sync_batchnorm = SyncBatchNorm.convert_sync_batchnorm

__all__ = ["DistributedDataParallel", "DistributedSampler", "SyncBatchNorm", "convert_sync_batchnorm", "Adam",
           "CrossEntropyLoss", "FlatArena", "nn", "optim"]
__version__ = "0.1.0"
