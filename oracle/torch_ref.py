#!/usr/bin/env python3
"""Stock-PyTorch restatement of the reference's CPU path -- TEST / BASELINE INFRASTRUCTURE.

/root/reference does not exist on the GPU box, and its demo.py hard-codes CUDA calls, so
the "reference's gloo CPU path"  is restated here with exactly the
library objects the reference uses: nn.Linear / nn.LeakyReLU (toy_model_and_data.py:13-21),
nn.MSELoss (demo.py:86), optim.Adam(lr=1e-3) (demo.py:80-81), DistributedDataParallel over
a gloo process group (demo.py:70,72), DistributedSampler + DataLoader(batch_size=256)
(demo.py:144-148), and the loss all-reduce on a gloo side group (demo.py:84,114-118).
tests/test_torch_ref.py checks that it reproduces tests/golden (produced by the real
reference) BIT-EXACTLY, so timing it is timing the reference's arithmetic.

Only bench.py (cpu_baseline, --impl reference) and tests/ may run this file.

CLI (one process per rank; rank 0 prints one JSON line):
    python oracle/torch_ref.py --world W --rank R --port P --iters K --warmup N [--seed S]
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn.parallel import DistributedDataParallel
from torch.utils.data import DataLoader, Dataset, DistributedSampler


def make_model(hidden=10):
    widths = [2, hidden, hidden, hidden, hidden, 1]
    mods = []
    for i in range(5):
        mods.append(nn.Linear(widths[i], widths[i + 1]))
        if i < 4:
            mods.append(nn.LeakyReLU())
    holder = nn.Module()
    holder.layers = nn.Sequential(*mods)
    holder.forward = lambda x: holder.layers(x)
    return holder


class SyntheticPairs(Dataset):
    """x_i = [z_i, z_i], y_i = 0.5 eps_i + z_i^2 (toy_model_and_data.py:27-36), same RNG call order."""

    def __init__(self, n=512):
        z = torch.randn(n)
        self.data = [torch.tensor([float(v)] * 2) for v in z]
        self.y = [torch.randn(1) * 0.5 + v[0] ** 2 for v in self.data]

    def __getitem__(self, i):
        return self.data[i], self.y[i]

    def __len__(self):
        return len(self.data)


def run(world, rank, port, iters, warmup=0, seed=1234, mode="distributed", record=False):
    """Runs warmup+iters iterations; returns dict(seconds=<timed part>, losses, params)."""
    torch.set_num_threads(int(os.environ.get("OMP_NUM_THREADS", "1")))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.manual_seed(seed)
    nets = [make_model(), make_model()]
    data = SyntheticPairs()
    models = [DistributedDataParallel(n) for n in nets]
    opts = [torch.optim.Adam(m.parameters(), lr=1e-3) for m in models]
    if mode == "distributed":
        sampler = DistributedSampler(data, shuffle=True)
        loader = DataLoader(data, batch_size=256, shuffle=False, sampler=sampler)
    else:
        sampler, loader = None, DataLoader(data, batch_size=256, shuffle=False)
    side = dist.new_group(backend="gloo")
    criterion = nn.MSELoss()
    losses, done, epoch, t0 = [], 0, 0, None
    total = warmup + iters
    while done < total:
        if sampler is not None:
            sampler.set_epoch(epoch)
        for x, y in loader:
            if done == warmup:
                dist.barrier()
                t0 = time.perf_counter()
            for o in opts:
                o.zero_grad(set_to_none=True)
            outs = [m(x) for m in models]
            vals = []
            for out in outs:
                l = criterion(out, y)
                l.backward()
                vals.append(l)
            for o in opts:
                o.step()
            b = y.size(0)
            logged = []
            for l in vals:
                t = l.detach().cpu() * b
                dist.all_reduce(t, group=side, op=dist.ReduceOp.SUM)
                logged.append(float(t / (b * world)))
            if record:
                losses.append(logged)
            done += 1
            if done == total:
                break
        epoch += 1
    dist.barrier()
    seconds = time.perf_counter() - t0
    flat = [torch.cat([p.detach().reshape(-1) for p in n.parameters()]).numpy() for n in nets]
    dist.destroy_process_group()
    return dict(seconds=seconds, losses=losses, params=flat, global_batch=256 * (1 if world == 1 else 2))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--port", type=int, required=True)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    res = run(a.world, a.rank, a.port, a.iters, a.warmup, a.seed)
    if a.rank == 0:
        print(json.dumps({"seconds": res["seconds"], "iters": a.iters, "world": a.world}))


if __name__ == "__main__":
    main()
