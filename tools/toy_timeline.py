#!/usr/bin/env python3
"""Development aid: per-phase SM-clock timeline of toy_step_kernel in the hot (L2-warm) loop.

Builds a -DDDPB_TIMELINE variant of the library into gpurun_out/ (the product library is not
touched), runs 200 fused steps of the two-model job and prints, for the last step, the cycle
stamps of every warp of every CTA relative to the earliest stamp of its cluster.
Compute warps 0-3: 0 start, 1 smem zeroed, 2 weights staged, 3 layer 0 done (inputs arrived),
4 before layer 4, 5 before weight gradients, 6 before the push, 7 end.
Optimizer warps 4-5: 0 start, 1 smem zeroed, 2 staged, 3 partials arrived, 4 end.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_training_pytorch_b200 import _lib, build  # noqa: E402

out = os.path.join(ROOT, "gpurun_out", "libddp_b200_timeline.so")
os.makedirs(os.path.dirname(out), exist_ok=True)
subprocess.run(["nvcc"] + build.NVCC_FLAGS + ["-DDDPB_TIMELINE", "-o", out] + build.sources() + ["-ldl"], check=True)
lib = ctypes.CDLL(out)
lib.ddpb_toy_step.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]

g = np.load(os.path.join(ROOT, "tests", "golden", "ref_w1.npz"))
d = np.load(os.path.join(ROOT, "tests", "golden", "dataset.npz"))
X, Y = torch.from_numpy(d["x"]).cuda(), torch.from_numpy(d["y"]).cuda()
rows = torch.randint(0, 512, (64, 256), dtype=torch.int32, device="cuda")
table = torch.from_numpy(_lib.adam_bias_table(1e-3, 0.9, 0.999)).cuda()
stamps = torch.zeros(2 * 8 * 64, dtype=torch.int64, device="cuda")
jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
keep = []
for i, key in enumerate(("init_X", "init_Y")):
    bufs = dict(p=torch.from_numpy(g[key]).cuda(), g=torch.zeros(372, device="cuda"), m=torch.zeros(371, device="cuda"),
                v=torch.zeros(371, device="cuda"), s=torch.zeros(4, dtype=torch.int64, device="cuda"),
                r=torch.zeros(1024, device="cuda"))
    keep.append(bufs)
    j = jobs[i]
    j.params, j.grads, j.m, j.v, j.state = (bufs[k].data_ptr() for k in "pgmvs")
    j.loss_ring, j.loss_ring_n = bufs["r"].data_ptr(), 1024
    j.bias_table, j.bias_table_n = table.data_ptr(), table.shape[0]
    j.x, j.y, j.idx, j.idx_rows, j.idx_stride = X.data_ptr(), Y.data_ptr(), rows.data_ptr(), int(os.environ.get('ROWS', 1)), 256
    j.n_data = int(os.environ.get('NDATA', 512))
    j.B, j.mode, j.opt = int(os.environ.get("B", 256)), _lib.MODE_STEP, _lib.OPT_ADAM
    j.lr, j.beta1, j.beta2, j.eps = 1e-3, 0.9, 0.999, 1e-8
    j.out = stamps.data_ptr()
stream = torch.cuda.current_stream().cuda_stream
for _ in range(200):
    assert lib.ddpb_toy_step(jobs, 2, stream) == 0
torch.cuda.synchronize()
t = stamps.cpu().numpy().reshape(2, 8, 8, 8)          # model, cta, warp, slot
for model in range(2):
    base = t[model][t[model] > 0].min()
    print(f"model {model}: cycles relative to the cluster's first stamp")
    for cta in (0, 3, 7):
        for warp in range(6):
            rel = [int(v - base) if v > 0 else -1 for v in t[model, cta, warp, :8]]
            print(f"  cta {cta} warp {warp}: {rel}")
print("loss", float(keep[0]["r"][199]), float(keep[1]["r"][199]))
