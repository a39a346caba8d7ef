#!/usr/bin/env python3
"""Widened-model measurements (BASELINE.json configs[4]): step time and per-kernel roofline
fractions of the tcgen05 GEMMs and the fused Adam pass, CUDA events, one GPU.

    python tools/bench_wide.py [--hidden 1024 4096 8192] [--batch 256] [--steps 20]
Prints one JSON line per width."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import distributed_training_pytorch_b200 as b200  # noqa: E402
from distributed_training_pytorch_b200 import _lib  # noqa: E402


def timed(fn, reps, flush):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        if flush is not None:
            flush.fill_(1)
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2]


def sweep(a, world):
    """Several ranks (torchrun): step time of the widened model per bucket cap; the NVLink side of the
    roofline is the all-reduce of 4*P bytes (bus factor 2(W-1)/W) against 770 GB/s measured per direction."""
    import torch.distributed as dist
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    side = dist.new_group(backend="gloo")
    caps = a.bucket_caps or [1, 5, 25, 100, 400, 1e9]
    for H in a.hidden:
        B = max(1, min(a.batch, 512 // world))
        for cap in caps:
            torch.manual_seed(0)
            model = b200.DistributedDataParallel(b200.ToyModel(hidden=H).cuda(), device_ids=[local], bucket_cap_mb=(cap or None))   # 0 = the package's default plan
            opt = b200.optim.Adam(model.parameters(), lr=1e-3)
            loss_fn = b200.nn.MSELoss()
            g = torch.Generator(device="cuda").manual_seed(rank)
            z = torch.randn(B, device="cuda", generator=g)
            x = torch.stack([z, z], 1).contiguous(); y = (z * z).view(B, 1).contiguous()

            def step():
                l = loss_fn(model(x), y); l.backward(); opt.step()
            for _ in range(3):
                step()
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                step()
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=side)
            P = model._backend.P
            # each bucket's fused all-reduce + Adam kernel on its own (nothing else on the GPU): is the
            # kernel at wire speed, or is it the overlap with backward that costs?
            alone = None
            be = model._backend
            if be.fused_exchange:
                # DDPB_BUCKET_DIAG: comma-separated diagnostic modes of the bucket kernel (0 = the real kernel;
                # 1 = read the local buffer instead of the peers'; 2 = do not push parameters; 3 = both; results of
                # modes != 0 are wrong on purpose) -- each measured in this same process
                alone = {}
                cs = torch.cuda.current_stream().cuda_stream
                for diag in [int(v) for v in os.environ.get("DDPB_BUCKET_DIAG", "0").split(",")]:
                    rows = []
                    for index, (layer, start, n) in enumerate(be.buckets):
                        be.bucket_jobs[index].reserved = diag
                        ts = []
                        for rep in range(5):
                            dist.barrier(); torch.cuda.synchronize()
                            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                            be.ticket += 1
                            a0.record(); be._fused_update(index, cs, True, False); a1.record(); torch.cuda.synchronize()
                            ts.append(a0.elapsed_time(a1) * 1e3)
                        tk = torch.tensor([sorted(ts)[2]], dtype=torch.float64)
                        dist.all_reduce(tk, op=dist.ReduceOp.MAX, group=side)
                        rows.append({"layer": layer, "floats": n, "us": round(float(tk), 1),
                                     "nvlink_in_gbs": round(4.0 * n * (world - 1) / world / float(tk) / 1e3, 1)})
                        be.bucket_jobs[index].reserved = 0
                    alone[f"diag{diag}"] = rows
            if rank == 0:
                ms = float(t)
                wire = 2.0 * (world - 1) / world * 4.0 * P
                print(json.dumps({"hidden": H, "world": world, "local_batch": B, "bucket_cap_mb": cap,
                                  "buckets": len(model._backend.buckets), "ms_per_step": ms,
                                  "bucket_kernels_alone": alone,
                                  "allreduce": "fused peer-window reduce+Adam kernel" if model._backend.fused_exchange
                                  else "ncclAllReduce + fused avg+Adam",
                                  "samples_per_s": B * world / (ms * 1e-3),
                                  "allreduce_wire_bytes_per_gpu": wire,
                                  "nvlink_floor_ms": wire / 770e9 * 1e3, "hbm_floor_ms": 40.0 * P / 6545.3e9 * 1e3}), flush=True)
            model._backend.close()                           # frees the peer windows of this model (collective)
            del model, opt
            torch.cuda.empty_cache()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, nargs="+", default=[1024, 4096, 8192])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--bucket-caps", type=float, nargs="+", default=None,
                    help="under torchrun: sweep bucket_cap_mb (step time per cap, max over ranks)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return sweep(a, world)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
    L = _lib.lib()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for H in a.hidden:
        B = a.batch
        torch.manual_seed(0)
        model = b200.DistributedDataParallel(b200.ToyModel(hidden=H).cuda(), device_ids=[0])
        opt = b200.optim.Adam(model.parameters(), lr=1e-3)
        loss_fn = b200.nn.MSELoss()
        z = torch.randn(B, device="cuda")
        x = torch.stack([z, z], 1).contiguous(); y = (z * z).view(B, 1).contiguous()

        def step():
            l = loss_fn(model(x), y); l.backward(); opt.step()
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        ms_step = timed(step, a.steps, None)
        be = model._backend
        st = torch.cuda.current_stream().cuda_stream
        w, bias = be._w(2)
        gw, gb = be._g(2)
        P = be.P
        k = {}
        k["fwd"] = timed(lambda: L.ddpb_linear_fwd(be.act[0].data_ptr(), w, bias, be.act[1].data_ptr(), be.act_t[1].data_ptr(), B, H, H, be.ld, 1, st), a.steps, flush)
        k["dgrad"] = timed(lambda: L.ddpb_linear_dgrad(be.dz[0].data_ptr(), w, be.act[0].data_ptr(), be.dz[1].data_ptr(), be.dz_t[1].data_ptr(), B, H, H, be.ld, st), a.steps, flush)
        k["wgrad"] = timed(lambda: L.ddpb_linear_wgrad(be.dz_t[0].data_ptr(), be.act_t[0].data_ptr(), gw, gb, B, H, H, be.ld, st), a.steps, flush)
        k["adam"] = timed(lambda: be._update(0, P, st, False), a.steps, flush)
        wbytes = 4.0 * H * H
        flops = 2.0 * B * H * H
        line = {"hidden": H, "batch": B, "params": P, "ms_per_step": ms_step, "samples_per_s": B / (ms_step * 1e-3),
                "kernels_ms": k,
                "gemm_hbm_frac": {n: wbytes / (k[n] * 1e-3) / 1e9 / peaks["hbm_gbs"] for n in ("fwd", "dgrad", "wgrad")},
                "gemm_tflops": {n: flops / (k[n] * 1e-3) / 1e12 for n in ("fwd", "dgrad", "wgrad")},
                "adam_gbs": 28.0 * P / (k["adam"] * 1e-3) / 1e9, "adam_hbm_frac": 28.0 * P / (k["adam"] * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "step_floor_ms": (40.0 * P / (peaks["hbm_gbs"] * 1e9)) * 1e3,
                "note": "kernel times: median of CUDA-event pairs with a 256 MiB L2 flush before each; algorithmic bytes 4*H*H per GEMM, 28*P for Adam; step floor = 40 B/param"}
        print(json.dumps(line), flush=True)
        del model, opt, be
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
