"""Helpers for the -m gpu tests: call the C ABI on torch-owned device buffers."""

import numpy as np
import torch

from distributed_training_pytorch_b200 import _lib

P = _lib.TOY_PARAMS


def dev(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda()


def new_state():
    return torch.zeros(4, dtype=torch.int64, device="cuda")


_TABLES = {}


def bias_table(lr=1e-3, beta1=0.9, beta2=0.999):
    key = (lr, beta1, beta2)
    if key not in _TABLES:
        _TABLES[key] = torch.from_numpy(_lib.adam_bias_table(lr, beta1, beta2)).cuda()
    return _TABLES[key]


class ToyHarness:
    """Device buffers of one toy model + a prefilled ddpb_toy_job."""

    def __init__(self, flat_params, opt=_lib.OPT_ADAM, lr=1e-3, momentum=0.0, ring=64):
        self.params = dev(flat_params)
        self.grads = torch.zeros(P + 1, device="cuda")
        self.m = torch.zeros(P, device="cuda")
        self.v = torch.zeros(P, device="cuda")
        self.state = new_state()
        self.ring = torch.zeros(ring, device="cuda")
        j = self.job = _lib.ToyJob()
        j.params, j.grads, j.m, j.v = (t.data_ptr() for t in (self.params, self.grads, self.m, self.v))
        j.state, j.loss_ring, j.loss_ring_n = self.state.data_ptr(), self.ring.data_ptr(), ring
        j.opt, j.lr, j.beta1, j.beta2, j.eps, j.momentum = opt, lr, 0.9, 0.999, 1e-8, momentum
        self.table = bias_table(lr)
        j.bias_table, j.bias_table_n = self.table.data_ptr(), self.table.shape[0]
        j.idx_rows, j.idx_stride = 1, 0
        self.seq = 0
        self.keep = None

    def bind(self, x, y, mode, idx=None, idx_rows=1, idx_stride=0, B=None, out=None, staged=True):
        j = self.job
        self.keep = (x, y, idx, out)
        j.x, j.y = x.data_ptr(), (y.data_ptr() if y is not None else None)
        j.idx = idx.data_ptr() if idx is not None else None
        j.idx_rows, j.idx_stride = idx_rows, idx_stride
        j.n_data = x.shape[0] if (idx is not None and staged) else 0
        j.B = int(B if B is not None else x.shape[0])
        j.mode = mode
        j.out = out.data_ptr() if out is not None else None

    def launch(self, others=()):
        jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
        jobs[0] = self.job
        for i, o in enumerate(others, 1):
            jobs[i] = o.job
        n = 1 + len(others)
        _lib.check(_lib.lib().ddpb_toy_step(jobs, n, torch.cuda.current_stream().cuda_stream), "ddpb_toy_step")
        slots = []
        for h in (self,) + tuple(others):
            slots.append(h.seq % h.job.loss_ring_n)
            if h.job.mode != _lib.MODE_FWD:
                h.seq += 1
        return slots

    def loss(self, slot):
        return float(self.ring[slot].item())
