#!/usr/bin/env python3
"""Development aid: time the three layer GEMMs of the widened model alone (CUDA events, L2 flushed,
median) for the kernel family / K split selected by DDPB_GEMM_KERNEL / DDPB_GEMM_SPLIT.

    python tools/gemm_probe.py [--hidden 4096] [--batch 256] [--reps 15]"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_training_pytorch_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, nargs="+", default=[4096])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=15)
    a = ap.parse_args()
    L = _lib.lib()
    handle = ctypes.CDLL(_lib.LIB_PATH)
    clusters = {}
    for s in (1, 2, 4):
        n = ctypes.c_int()
        if handle.ddpb_gemm_max_active_clusters(s, ctypes.byref(n)) == 0:
            clusters[s] = n.value
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for H in a.hidden:
        B, ld = a.batch, 256
        x = torch.randn(B, H, device="cuda")
        W = torch.randn(H, H, device="cuda") / H ** 0.5
        b = torch.randn(H, device="cuda")
        out, out_t = torch.empty(B, H, device="cuda"), torch.zeros(H + 1, ld, device="cuda")
        dz, dz_t = torch.randn(B, H, device="cuda"), torch.randn(H, ld, device="cuda")
        dW, db = torch.empty(H, H, device="cuda"), torch.empty(H, device="cuda")

        def timed(fn):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.reps)]
            for e0, e1 in ev:
                flush.fill_(1)
                e0.record(); _lib.check(fn()); e1.record()
            torch.cuda.synchronize()
            t = sorted(e0.elapsed_time(e1) for e0, e1 in ev)
            return t[len(t) // 2] * 1e3
        k = {"fwd": timed(lambda: L.ddpb_linear_fwd(x.data_ptr(), W.data_ptr(), b.data_ptr(), out.data_ptr(), out_t.data_ptr(), B, H, H, ld, 1, st)),
             "dgrad": timed(lambda: L.ddpb_linear_dgrad(dz.data_ptr(), W.data_ptr(), x.data_ptr(), out.data_ptr(), out_t.data_ptr(), B, H, H, ld, st)),
             "wgrad": timed(lambda: L.ddpb_linear_wgrad(dz_t.data_ptr(), out_t.data_ptr(), dW.data_ptr(), db.data_ptr(), B, H, H, ld, st))}
        print(json.dumps({"hidden": H, "batch": B, "kernel": os.environ.get("DDPB_GEMM_KERNEL", "pair"),
                          "split": os.environ.get("DDPB_GEMM_SPLIT", "auto"), "us": {n: round(v, 1) for n, v in k.items()},
                          "hbm_frac": {n: round(4.0 * H * H / (v * 1e-6) / 1e9 / 6545.3, 3) for n, v in k.items()},
                          "max_active_clusters": clusters}), flush=True)


if __name__ == "__main__":
    main()
