#!/usr/bin/env python3
"""Development aid: where does the host time of one end-to-end iteration go?"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_training_pytorch_b200 import DataLoader, DistributedSampler, ToyData, ToyModel, nn, optim  # noqa: E402
from distributed_training_pytorch_b200 import DistributedDataParallel as DDP  # noqa: E402

torch.manual_seed(1234)
mx, my, data = ToyModel(), ToyModel(), ToyData()
model_X, model_Y = DDP(mx.cuda(), device_ids=[0]), DDP(my.cuda(), device_ids=[0])
sampler = DistributedSampler(data, num_replicas=1, rank=0, shuffle=True)
loader = DataLoader(data, batch_size=256, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler)
opt_X, opt_Y = optim.Adam(model_X.parameters(), lr=1e-3), optim.Adam(model_Y.parameters(), lr=1e-3)
loss = nn.MSELoss()


def run(iters, marks=None):
    it, epoch = 0, 0
    while it < iters:
        sampler.set_epoch(epoch)
        t0 = time.perf_counter()
        for x, t in loader:
            t1 = time.perf_counter()
            opt_X.zero_grad(set_to_none=True); opt_Y.zero_grad(set_to_none=True)
            x, t = x.cuda(), t.cuda()
            t2 = time.perf_counter()
            l_X = loss(model_X(x), t); l_X.backward()
            l_Y = loss(model_Y(x), t); l_Y.backward()
            opt_X.step(); opt_Y.step()
            t3 = time.perf_counter()
            a = l_X.detach().cpu()
            t4 = time.perf_counter()
            b = l_Y.detach().cpu()
            t5 = time.perf_counter()
            if marks is not None:
                marks.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4))
            it += 1
            if it == iters:
                break
            t0 = time.perf_counter()
        epoch += 1


run(200)
marks = []
t = time.perf_counter()
run(4000, marks)
dt = time.perf_counter() - t
print(f"wall per iteration: {dt / 4000 * 1e6:.1f} us")
import numpy as np
m = np.median(np.asarray(marks), axis=0) * 1e6
print("median us: loader %.1f | .cuda x2 %.1f | describe (fwd/loss/bwd/step x2) %.1f | first .cpu (launch+wait) %.1f | second .cpu %.1f" % tuple(m))
pr = cProfile.Profile()
pr.enable(); run(2000); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
