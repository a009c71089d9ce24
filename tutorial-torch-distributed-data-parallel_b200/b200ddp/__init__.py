"""b200ddp — B200-native DDP training step for the toy model of
annalena-k/tutorial-torch-distributed-data-parallel, behind the surface the tutorial's scripts use.

    from b200ddp import DistributedDataParallel as DDP, DistributedSampler    # multi-GPU-training-torch.py:17,19
    from b200ddp import optim, nn as bnn                                      # optim.Adam, bnn.CrossEntropyLoss
    from b200ddp.accelerate import Accelerator                                # multi-GPU-training-accelerate.py:19
    from b200ddp.models import load_model                                     # data_and_toy_model.py:41-45
"""
from . import _lib, nn, optim  # noqa: F401
from .arena import FlatArena  # noqa: F401
from .distributed import DistributedDataParallel  # noqa: F401
from .nn import CrossEntropyLoss, SyncBatchNorm  # noqa: F401
from .optim import Adam  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401

convert_sync_batchnorm = SyncBatchNorm.convert_sync_batchnorm

__all__ = ["DistributedDataParallel", "DistributedSampler", "SyncBatchNorm", "convert_sync_batchnorm", "Adam",
           "CrossEntropyLoss", "FlatArena", "nn", "optim"]
__version__ = "0.1.0"
