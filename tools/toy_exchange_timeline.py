#!/usr/bin/env python3
"""Development aid: where does the in-kernel gradient exchange of toy_step_kernel spend its time on W GPUs?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 tools/toy_exchange_timeline.py

Builds a -DDDPB_XCHG_TIMELINE variant of the library into gpurun_out/ (the product library is not touched)
and runs 400 back-to-back fused steps of ONE toy model per rank through the C ABI (peer windows over IPC, the
sampler's shard of each epoch), a host barrier every 100 steps.  The thread that owns bucket element 0 stamps
%globaltimer at kernel entry, when the local gradients are reduced, when its exchange stores are issued, when
every rank's ticket has been seen, and at the end.  Rank 0 prints one JSON line per rank: median / 10th
percentile / minimum of each phase in us.  The MINIMUM of the wait phase is the NVLink round trip (a step
in which every peer had already delivered); median minus minimum is launch skew between the processes."""
import ctypes
import json
import os
import socket
import subprocess
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_training_pytorch_b200 import _lib, build  # noqa: E402

STEPS = 400


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    out = os.path.join(ROOT, "gpurun_out", "libddp_b200_xchg_timeline.so")
    if rank == 0:
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(["nvcc"] + build.NVCC_FLAGS + ["-DDDPB_XCHG_TIMELINE", "-o", out] + build.sources() + ["-ldl"], check=True)
    dist.barrier()
    lib = ctypes.CDLL(out)
    lib.ddpb_toy_step.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.ddpb_peer_alloc.argtypes = [ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
    lib.ddpb_peer_open.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.ddpb_sampler_epochs_i32.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint64,
                                            ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64]

    mine, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
    assert lib.ddpb_peer_alloc(_lib.XCHG_WINDOW_BYTES, ctypes.byref(mine), handle) == 0
    handles = [None] * world
    dist.all_gather_object(handles, (socket.gethostname(), handle.raw))
    windows = []
    for r, (_, raw) in enumerate(handles):
        if r == rank:
            windows.append(mine.value)
        else:
            ptr = ctypes.c_void_p()
            assert lib.ddpb_peer_open(raw, ctypes.byref(ptr)) == 0
            windows.append(ptr.value)

    g = np.load(os.path.join(ROOT, "tests", "golden", "ref_w1.npz"))
    d = np.load(os.path.join(ROOT, "tests", "golden", "dataset.npz"))
    X, Y = torch.from_numpy(d["x"]).cuda(), torch.from_numpy(d["y"]).cuda()
    B = -(-512 // world)
    rows_h = np.empty((STEPS, B), np.int32)
    assert lib.ddpb_sampler_epochs_i32(512, world, rank, 1, 0, 0, STEPS, 0, rows_h.ctypes.data, B) == 0
    rows = torch.from_numpy(rows_h).cuda()
    table = torch.from_numpy(_lib.adam_bias_table(1e-3, 0.9, 0.999)).cuda()
    stamps = torch.zeros(STEPS, 8, dtype=torch.int64, device="cuda")
    bufs = dict(p=torch.from_numpy(g["init_X"]).cuda(), g=torch.zeros(372, device="cuda"), m=torch.zeros(371, device="cuda"),
                v=torch.zeros(371, device="cuda"), s=torch.zeros(4, dtype=torch.int64, device="cuda"),
                r=torch.zeros(1024, device="cuda"))
    jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
    j = jobs[0]
    j.params, j.grads, j.m, j.v, j.state = (bufs[k].data_ptr() for k in "pgmvs")
    j.loss_ring, j.loss_ring_n = bufs["r"].data_ptr(), 1024
    j.bias_table, j.bias_table_n = table.data_ptr(), table.shape[0]
    j.x, j.y, j.idx_rows, j.idx_stride, j.n_data = X.data_ptr(), Y.data_ptr(), 1, B, 512
    j.B, j.mode, j.opt = B, _lib.MODE_STEP, _lib.OPT_ADAM
    j.lr, j.beta1, j.beta2, j.eps = 1e-3, 0.9, 0.999, 1e-8
    j.world, j.rank, j.inv_world = world, rank, 1.0 / world
    for r, w in enumerate(windows):
        j.peer_data[r] = w
    stream = torch.cuda.current_stream().cuda_stream
    for step in range(STEPS):
        if step % 100 == 0:
            torch.cuda.synchronize()
            dist.barrier()
        j.idx = rows.data_ptr() + step * B * 4
        j.out = stamps.data_ptr() + step * 64
        assert lib.ddpb_toy_step(jobs, 1, stream) == 0
    torch.cuda.synchronize()
    assert int(bufs["s"][2]) == 0, "a peer timed out"
    t = stamps.cpu().numpy().astype(np.float64)[20:]             # skip the warm-up steps
    phases = {"compute_local_grads": t[:, 1] - t[:, 0], "issue_stores": t[:, 2] - t[:, 1], "wait_all_tickets": t[:, 3] - t[:, 2],
              "adam_and_stores": t[:, 4] - t[:, 3], "kernel": t[:, 4] - t[:, 0]}
    rec = {"rank": rank, "world": world, "local_batch": B, "steps": int(t.shape[0]),
           "us": {k: {"median": round(float(np.median(v)) / 1e3, 2), "p10": round(float(np.percentile(v, 10)) / 1e3, 2),
                      "min": round(float(v.min()) / 1e3, 2)} for k, v in phases.items()},
           "polling_rounds_median": float(np.median(t[:, 5]))}
    recs = [None] * world
    dist.all_gather_object(recs, rec)
    if rank == 0:
        for r in recs:
            print(json.dumps(r), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
