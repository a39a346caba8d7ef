"""The toy model and the synthetic dataset of the reference, as this package's front end.

ToyModel mirrors toy_model_and_data.py:8-25 (five Linear layers 2->H->H->H->H->1 with
LeakyReLU(0.01) between them, H = 10) and is still an nn.Module, so parameter
initialisation consumes the torch RNG exactly like the reference (kaiming_uniform_ +
uniform bias per nn.Linear, torch:nn/modules/linear.py:117-128) and state_dict keys stay
`layers.{0,2,4,6,8}.{weight,bias}`.  The arithmetic never runs through these modules:
DistributedDataParallel (ddp.py) flattens the parameters into one HBM buffer and runs the
sm_100a kernels on it.

ToyData mirrors toy_model_and_data.py:27-36: z ~ N(0,1) (one randn(512) call),
x_i = [z_i, z_i], y_i = 0.5*eps_i + z_i^2 with one randn(1) call per sample (the call
pattern matters: it fixes the RNG stream, hence the bit-exact dataset of tests/golden).
"""
import os

import torch
import torch.nn as nn
from torch.utils.data import Dataset

DEFAULT_HIDDEN = 10


def _hidden_from_env(hidden):
    if hidden is not None:
        return int(hidden)
    return int(os.environ.get("DDPB_HIDDEN", DEFAULT_HIDDEN))     # additive knob (config 5: widened model)


def _mlp(dims, act_last):
    mods = []
    for i in range(len(dims) - 1):
        mods.append(nn.Linear(dims[i], dims[i + 1]))
        if i + 2 < len(dims) or act_last:
            mods.append(nn.LeakyReLU())
    return nn.Sequential(*mods)


class ToyModel(nn.Module):
    """2 -> H -> H -> H -> H -> 1 MLP; H defaults to the reference's 10."""

    def __init__(self, hidden=None):
        super().__init__()
        self.hidden = _hidden_from_env(hidden)
        h = self.hidden
        self.dims = (2, h, h, h, h, 1)
        self.layers = _mlp(self.dims, act_last=False)

    def forward(self, x):
        raise RuntimeError(
            "ToyModel computes only through distributed_training_pytorch_b200.DistributedDataParallel "
            "(sm_100a kernels); wrap it first. There is no eager/CPU forward in this package.")


class MultiGPUModel(nn.Module):
    """The two-stage split of demo_one_model_multi_gpu.py:17-42: layers0 (2->H->H, both
    activated) on dev0, layers1 (H->H->H->1) on dev1."""

    def __init__(self, dev0, dev1, hidden=None):
        super().__init__()
        self.dev0, self.dev1 = dev0, dev1
        self.hidden = _hidden_from_env(hidden)
        h = self.hidden
        self.dims0, self.dims1 = (2, h, h), (h, h, h, 1)
        self.layers0 = _mlp(self.dims0, act_last=True)
        self.layers1 = _mlp(self.dims1, act_last=False)
        self.layers0.to(self.dev0)
        self.layers1.to(self.dev1)

    def forward(self, x):
        raise RuntimeError("MultiGPUModel computes only through DistributedDataParallel(device_ids=None)")


class ToyData(Dataset):
    """512 samples; dense copies (x_dense [n,2], y_dense [n,1]) feed the native loaders."""

    def __init__(self, n=512):
        z = torch.randn(n)
        noise = torch.cat([torch.randn(1) for _ in range(n)])      # one draw per sample, in order
        self.x_dense = torch.stack([z, z], dim=1).contiguous()
        self.y_dense = (noise * 0.5 + z ** 2).reshape(n, 1).contiguous()
        self.data = list(self.x_dense.unbind(0))
        self.y = list(self.y_dense.unbind(0))

    def __getitem__(self, idx):
        return self.data[idx], self.y[idx]

    def __len__(self):
        return len(self.data)
