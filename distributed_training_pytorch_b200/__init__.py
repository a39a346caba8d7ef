"""distributed_training_pytorch_b200 -- the DDP training step of the
ammunk/distributed-training-pytorch toy model as hand-written sm_100a kernels behind the
reference's own front-end call shapes (include/ddp_b200.h).

Front end (drop-in names):
    from distributed_training_pytorch_b200 import (
        DistributedDataParallel as DDP, DataLoader, DistributedSampler, nn, optim,
        ToyModel, ToyData, get_args)
"""
from types import SimpleNamespace

from . import _lib
from ._lib import DdpB200Error
from .argument_parser import get_args
from .data import DataLoader, DistributedSampler
from .ddp import LOCAL_ONLY, DistributedDataParallel
from .loss import MSELoss
from .optim import SGD, Adam
from .toy_model_and_data import MultiGPUModel, ToyData, ToyModel

# `nn.MSELoss()` / `optim.Adam(...)` spellings used by the demos (demo.py:80-86)
nn = SimpleNamespace(MSELoss=MSELoss)
optim = SimpleNamespace(Adam=Adam, SGD=SGD)

__all__ = ["DistributedDataParallel", "DataLoader", "DistributedSampler", "MSELoss", "Adam", "SGD", "nn", "optim",
           "ToyModel", "ToyData", "MultiGPUModel", "get_args", "DdpB200Error", "LOCAL_ONLY"]
