#!/usr/bin/env python3
"""Headline benchmark: toy-model train samples/sec (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (`config.workload`): demo.py's iteration -- TWO 371-parameter toy models, the
512-sample synthetic dataset, DistributedSampler shards, local batch 256/256/128/64 at
W = 1/2/4/8, Adam lr 1e-3 (reference demo.py:80-129).  One "step" = one iteration of that
loop (both models: forward, MSE, backward, gradient all-reduce when W > 1, Adam).

Numbers in the JSON line (rank 0 prints exactly one):
  value     samples/s with the dataset and the sampler's index rows resident in HBM
            (C-ABI launches, no host copies), CUDA events per step, L2 flushed between
            timed steps, max over ranks.
  hot_l2    the same steps back to back without flushes (one event pair around all K).
  e2e       samples/s through the public front end (DataLoader -> .cuda() H2D -> model /
            loss / backward / step -> loss.detach().cpu() D2H) from pinned host batches.
  roofline  dominant kernel (toy_step_kernel) against the measured HBM peak; see DESIGN.md
            for the algorithmic bytes.  The kernel is latency-bound by construction.
  cpu_baseline  oracle/torch_ref.py (the reference's stock-PyTorch CPU/gloo path, verified
            bit-identical to the reference run) timed on the host for a bounded sample.

`--impl reference` times that CPU path as the arm the driver compares against.
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "toy-model train samples/sec (device-timed, max over ranks)"
UNIT = "samples/s"
DATASET = 512
BATCH = 256
P = 371
H = 10


def local_batch(world):
    return min(BATCH, -(-DATASET // world))


def global_batch(world):
    return local_batch(world) * world


def workload(world):
    return {"workload": "demo.py: 2 x ToyModel(2-10-10-10-10-1), ToyData(512), DistributedSampler, Adam lr=1e-3",
            "models": 2, "params_per_model": P, "dataset": DATASET, "local_batch": local_batch(world),
            "global_batch": global_batch(world), "world": world}


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


# --------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, windows):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, sm_max, reasons = [], None, set()
        for t, line in self.lines:
            if not any(a <= t <= b for a, b in windows):
                continue
            f = [c.strip() for c in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); sm_max = float(f[2])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": sm_max, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------
# reference arm / cpu baseline: oracle/torch_ref.py on the host cores
# --------------------------------------------------------------------------------------
def run_cpu_reference(world, iters, warmup):
    """Spawns the W-rank gloo CPU job (one process per rank, OMP_NUM_THREADS=1 like the
    reference's launch scripts) and returns the seconds rank 0 measured for `iters` steps."""
    port = free_port()
    # Not the launcher's environment: under torch.distributed.run, TORCHELASTIC_USE_AGENT_STORE makes
    # init_process_group("tcp://...") expect the agent's store instead of starting its own, and the
    # job would wait for ever.
    launcher = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK", "GROUP_WORLD_SIZE", "ROLE_RANK",
                "ROLE_WORLD_SIZE", "ROLE_NAME", "MASTER_ADDR", "MASTER_PORT")
    env = {k: v for k, v in os.environ.items() if k not in launcher and not k.startswith(("TORCHELASTIC_", "TORCH_NCCL_"))}
    env.update(OMP_NUM_THREADS="1", MKL_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    script = os.path.join(ROOT, "oracle", "torch_ref.py")
    procs = [subprocess.Popen([sys.executable, script, "--world", str(world), "--rank", str(r), "--port", str(port),
                               "--iters", str(iters), "--warmup", str(warmup)], env=env, cwd=ROOT,
                              stdout=subprocess.PIPE if r == 0 else subprocess.DEVNULL, stderr=subprocess.DEVNULL,
                              text=True) for r in range(world)]
    out, _ = procs[0].communicate(timeout=900)
    for p in procs[1:]:
        p.wait(timeout=60)
    line = [l for l in out.splitlines() if l.startswith("{")][-1]
    return json.loads(line)["seconds"]


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    # bounded sample: each "step" is one iteration of the same loop; ~2.5 ms each on one core
    seconds = run_cpu_reference(world, args.steps, args.warmup)
    value = args.steps * global_batch(world) / seconds
    cores = os.cpu_count()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * seconds / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload(world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": world, "kind": "port",
                             "sample": f"{args.steps} iterations of the same loop, {world} gloo rank(s) x 1 thread "
                                       f"(host has {cores} cores), stock torch CPU path (oracle/torch_ref.py, "
                                       "bit-identical to the reference run)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# --------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------
def ours(args):
    import torch
    import torch.distributed as dist

    from distributed_training_pytorch_b200 import DataLoader, DistributedSampler, ToyData, ToyModel, _lib, nn, optim
    from distributed_training_pytorch_b200 import DistributedDataParallel as DDP
    from distributed_training_pytorch_b200.engine import engine_for

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{free_port()}", rank=0, world_size=1)
    side = dist.new_group(backend="gloo")
    eng = engine_for(local_rank)
    K, W = args.steps, args.warmup
    lb, gb = local_batch(world), global_batch(world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=side)
        return float(t)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # 2x the 126 MB L2

    def build(device_resident):
        torch.manual_seed(1234)
        mx, my, data = ToyModel(), ToyModel(), ToyData()
        model_X, model_Y = DDP(mx.cuda(), device_ids=[local_rank]), DDP(my.cuda(), device_ids=[local_rank])
        sampler = DistributedSampler(data, shuffle=True)
        loader = DataLoader(data, batch_size=BATCH, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler,
                            device_resident=device_resident)
        opts = (optim.Adam(model_X.parameters(), lr=1e-3), optim.Adam(model_Y.parameters(), lr=1e-3))
        return model_X, model_Y, sampler, loader, opts

    def iteration_stream(sampler, loader):
        epoch = 0
        while True:
            sampler.set_epoch(epoch)
            for batch in loader:
                yield batch
            epoch += 1

    loss_fn = nn.MSELoss()

    def one_step(models, opts, batch, read_loss):
        data, target = batch
        for o in opts:
            o.zero_grad(set_to_none=True)
        data, target = data.cuda(), target.cuda()
        losses = []
        for m in models:
            l = loss_fn(m(data), target)
            l.backward()
            losses.append(l)
        for o in opts:
            o.step()
        if read_loss:
            return [l.detach().cpu() for l in losses]
        eng.flush()
        return None

    clock = ClockSampler(local_rank)
    clock.start()
    windows = []

    # ---- value: HBM-resident inputs, per-step events, L2 flushed between timed steps
    model_X, model_Y, sampler, loader, opts = build(device_resident=True)
    batches = iteration_stream(sampler, loader)
    for _ in range(max(W, 3)):
        one_step((model_X, model_Y), opts, next(batches), read_loss=False)
    barrier()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    launches0 = eng.launches
    t_begin = time.perf_counter()
    for i in range(K):
        flush_buf.fill_(i & 0xff)                     # evict L2 (outside the timed region)
        batch = next(batches)
        starts[i].record()
        one_step((model_X, model_Y), opts, batch, read_loss=False)
        stops[i].record()
    barrier()
    windows.append((t_begin, time.perf_counter()))
    launches = eng.launches - launches0
    cold_ms = sum(s.elapsed_time(e) for s, e in zip(starts, stops))
    cold_ms = max_over_ranks(cold_ms)

    # ---- hot_l2: same steps back to back, straight through the C ABI as ONE captured CUDA graph
    # of K fused launches (each node carries its sampler row; step counters live on the device),
    # so the host is not the bottleneck.  With several ranks the gradient exchange is inside
    # the kernel (peer windows), so the same graph works; the NCCL path goes through the front end.
    direct = None
    fused = world == 1 or model_X._backend.fused_exchange
    if fused:
        eng.flush()
        n_rows = 1024
        per_epoch = len(sampler) // lb
        rows = torch.from_numpy(sampler.epochs_i32(0, n_rows // per_epoch).reshape(n_rows, lb)).cuda()
        jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
        for i, m in enumerate((model_X, model_Y)):
            j = m._backend.job
            j.loss_ring = m._backend.loss_dev.data_ptr()          # nobody reads losses in this loop: keep them in HBM
            j.x, j.y, j.idx = loader._dev[0].data_ptr(), loader._dev[1].data_ptr(), rows.data_ptr()
            j.idx_rows, j.idx_stride, j.B, j.mode, j.n_data = n_rows, lb, lb, _lib.MODE_STEP, DATASET
            jobs[i] = j
        stream = torch.cuda.current_stream().cuda_stream
        toy_step = _lib.lib().ddpb_toy_step

        # capture K back-to-back steps once; the replay is what is timed
        import ctypes
        graph = ctypes.c_void_p()
        L = _lib.lib()
        side_stream = torch.cuda.Stream()
        with torch.cuda.stream(side_stream):
            cap = side_stream.cuda_stream
            for _ in range(20):
                _lib.check(toy_step(jobs, 2, cap), "ddpb_toy_step")
            side_stream.synchronize()
            _lib.check(L.ddpb_graph_begin(cap), "ddpb_graph_begin")
            for i in range(K):
                for k in range(2):           # every graph node carries its own sampler row: no device-side lookup
                    jobs[k].idx, jobs[k].idx_rows = rows.data_ptr() + (i % n_rows) * lb * 4, 1
                _lib.check(toy_step(jobs, 2, cap), "ddpb_toy_step")
            _lib.check(L.ddpb_graph_end(cap, ctypes.byref(graph)), "ddpb_graph_end")
            _lib.check(L.ddpb_graph_launch(graph, cap), "ddpb_graph_launch")      # warm the instantiated graph
            side_stream.synchronize()

        def direct():
            _lib.check(L.ddpb_graph_launch(graph, stream), "ddpb_graph_launch")
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin = time.perf_counter()
    ev0.record()
    if direct is not None:
        direct()                                      # one graph launch = K fused steps
    else:
        for i in range(K):
            one_step((model_X, model_Y), opts, next(batches), read_loss=False)
    ev1.record()
    barrier()
    windows.append((t_begin, time.perf_counter()))
    hot_ms = max_over_ranks(ev0.elapsed_time(ev1))

    # ---- e2e: pinned host batches through the public front end, loss read back every step
    model_X, model_Y, sampler, loader, opts = build(device_resident=False)
    batches = iteration_stream(sampler, loader)
    for _ in range(max(W, 3)):
        one_step((model_X, model_Y), opts, next(batches), read_loss=True)
    barrier()
    t_begin = time.perf_counter()
    e2e_ms = 0.0
    last = None
    for i in range(K):
        flush_buf.fill_(i & 0xff)
        torch.cuda.synchronize()
        ev0.record()
        last = one_step((model_X, model_Y), opts, next(batches), read_loss=True)
        ev1.record()
        ev1.synchronize()
        e2e_ms += ev0.elapsed_time(ev1)
    barrier()
    windows.append((t_begin, time.perf_counter()))
    e2e_ms = max_over_ranks(e2e_ms)

    # keep the GPU busy long enough for nvidia-smi to see the clocks this workload runs at
    t_begin = time.perf_counter()
    while time.perf_counter() - t_begin < 1.0:
        for _ in range(200):
            one_step((model_X, model_Y), opts, next(batches), read_loss=False)
        torch.cuda.synchronize()
    windows.append((t_begin, time.perf_counter()))
    clocks = clock.stop(windows)

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs, burst)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # algorithmic bytes per launch: per model 4P (bucket) + 28P (Adam) + 48*B*H + 8*B (activations)
    alg_bytes = 2 * (4 * P + 28 * P + 48 * lb * H + 8 * lb)
    kernel_ms = hot_ms / K if fused else None
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms else None

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        iters = 4000                                        # ~10-15 s of single-core work
        secs = run_cpu_reference(1, iters, 20)
        cpu_baseline = {"value": iters * gb / secs, "unit": UNIT, "cores": 1, "kind": "port",
                        "sample": f"{iters} iterations of the same two-model loop (B=256), 1 gloo rank x 1 thread, "
                                  "stock torch CPU path (oracle/torch_ref.py, bit-identical to the reference run)"}

    if rank == 0:
        line = {"metric": METRIC, "value": K * gb / (cold_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": max(W, 3), "ms_per_step": cold_ms / K, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(workload(world), timing="CUDA events per step; L2 flushed (256 MiB fill) between timed "
                                                       "steps; max over ranks",
                               parallelism=f"dp{world}",
                               allreduce=("none (single rank)" if world == 1 else
                                          "fused one-shot exchange over NVLink peer windows inside the step kernel"
                                          if model_X._backend.fused_exchange else "ncclAllReduce per bucket + fused avg+Adam"),
                               scaling_note="one 512-sample epoch is the fixed job: 2 steps at W=1, 1 step at W>=2"),
                "hot_l2": {"value": K * gb / (hot_ms * 1e-3), "unit": UNIT, "ms_per_step": hot_ms / K,
                           "timing": "one CUDA-event pair around K back-to-back steps, no flush"
                                     + ("; the K launches are one captured CUDA graph" if fused else "")},
                "e2e": {"value": K * gb / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K,
                        "h2d_bytes_per_step": lb * 12, "d2h_bytes_per_step": 8},
                "gpu_launches": launches,
                "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": (achieved / peak) if achieved else None,
                             # dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full,
                             # profiles/r01_v3_toy_step_raw.csv (cold caches; warm launches hit L2)
                             "traffic": 50432 if lb == BATCH else None,
                             "kernel": "toy_step_kernel", "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "note": "latency-bound by construction: 2 clusters x 8 CTAs, 273 KB and 1 MFLOP per launch"},
                "cpu_baseline": cpu_baseline,
                "loss_after_bench": [float(v) for v in last] if last else None}
        emit(line)
    eng.destroy_comm()
    dist.destroy_process_group()


_RESULT_FD = None


def emit(line):
    """The one JSON line, written to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    # Libraries write to stdout too (NCCL prints its version banner there when NCCL_DEBUG is set on
    # the box): keep fd 1 for the result line only and send everything else to stderr.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
