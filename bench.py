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
import gc
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
class Ctx:
    pass


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of toy_step_kernel, from the newest
    committed `ncu --set full` raw page under profiles/ (cold caches: ncu flushes between replays)."""
    import csv
    import glob
    pages = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_toy_step_raw.csv")))
    if not pages:
        return None, None
    page = pages[-1]
    try:
        rows = list(csv.reader(open(page)))
        hdr, units = rows[0], rows[1]
        name_col = hdr.index("Kernel Name")
        body = [r for r in rows[2:] if "toy_step_kernel" in r[name_col]]
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        total = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(name)
            total += sum(float(r[i]) for r in body) / len(body) * scale[units[i]]
        return total, os.path.relpath(page, ROOT)
    except Exception:
        return None, os.path.relpath(page, ROOT)


def golden_parity(c):
    """Outside every timed region: the drop-in demo.py loop for 200 iterations from seed 1234 on THIS
    world size, once per gradient route, against the reference's own run at the same world size
    (tests/golden/ref_w{W}.npz, minted by tests/golden/make_golden.py from /root/reference/demo.py).
    Tolerances are those of tests/test_gpu_multi.py: logged losses rel 2e-5, parameters abs 2e-5 after
    200 steps, replicas bit-identical."""
    import hashlib

    import numpy as np
    import torch
    import torch.distributed as dist

    import demo
    from distributed_training_pytorch_b200 import DataLoader, DistributedSampler, ToyData, ToyModel
    from distributed_training_pytorch_b200 import DistributedDataParallel as DDP

    path = os.path.join(ROOT, "tests", "golden", f"ref_w{c.world}.npz")
    if not os.path.exists(path):
        return {"w": c.world, "ok": None, "skipped": f"no golden file for world {c.world}"}
    g = np.load(path)
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    routes = ["single"] if c.world == 1 else ["p2p", "nccl"]
    out = {"w": c.world, "steps": 200, "golden": os.path.relpath(path, ROOT),
           "tolerance": {"loss_rel": 2e-5, "param_abs": 2e-5, "replicas": "bit-identical"}}
    saved = {k: os.environ.get(k) for k in ("DDPB_ALLREDUCE", "DDPB_ITERATIONS", "DDPB_NO_WANDB", "DDPB_FUSED_LOSS")}
    ok_all = True
    for route in routes:
        os.environ.update(DDPB_ALLREDUCE="p2p" if route == "single" else route, DDPB_ITERATIONS="200", DDPB_NO_WANDB="1")
        os.environ.pop("DDPB_FUSED_LOSS", None)
        torch.manual_seed(1234)
        mx, my = ToyModel(), ToyModel()
        data = ToyData()
        model_X, model_Y = DDP(mx.cuda(), device_ids=[c.local_rank]), DDP(my.cuda(), device_ids=[c.local_rank])
        sampler = DistributedSampler(data, shuffle=True)
        loader = DataLoader(data, batch_size=BATCH, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler)
        logged = []
        demo.training_demo(model_X, model_Y, loader, sampler, c.world, on_loss=lambda it, lx, ly: logged.append((lx, ly)))
        px, py = model_X.flat_parameters().cpu().numpy(), model_Y.flat_parameters().cpu().numpy()
        digest = hashlib.sha256(px.tobytes() + py.tobytes()).hexdigest()
        digests = [None] * c.world
        dist.all_gather_object(digests, digest, group=c.side)
        rec = {"fused_exchange": bool(model_X._backend.fused_exchange), "replicas_identical": len(set(digests)) == 1,
               "params_sha256": digest[:16]}
        if c.rank == 0:
            losses = np.asarray(logged, np.float32)
            rec["loss_rel_max"] = float(np.max(np.abs(losses - ref) / ref)) if losses.shape == ref.shape else None
            rec["param_abs_max"] = float(max(np.max(np.abs(px - g["mid_X"])), np.max(np.abs(py - g["mid_Y"]))))
            rec["ok"] = bool(rec["loss_rel_max"] is not None and rec["loss_rel_max"] <= 2e-5 and rec["param_abs_max"] <= 2e-5
                             and rec["replicas_identical"] and (route != "p2p" or rec["fused_exchange"]))
        flags = [rec.get("ok")]
        dist.broadcast_object_list(flags, src=0, group=c.side)
        rec["ok"] = flags[0]
        ok_all = ok_all and bool(flags[0])
        out[route] = rec
        del model_X, model_Y
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v
    out["ok"] = ok_all
    return out


def incumbent_torch(c, hidden, batch, steps, allow_tf32):
    """The GPU incumbent: the reference's own stock-PyTorch modules (nn.Linear /
    LeakyReLU / MSELoss / optim.Adam, torch DDP over NCCL when W > 1) on the SAME GPU, inputs resident.
    Returns ms per iteration of one model's step (max over ranks)."""
    import torch
    import torch.nn as tnn
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = bool(allow_tf32)
    try:
        torch.manual_seed(0)
        dims = (2, hidden, hidden, hidden, hidden, 1)
        layers = []
        for i in range(5):
            layers.append(tnn.Linear(dims[i], dims[i + 1]))
            if i < 4:
                layers.append(tnn.LeakyReLU())
        model = tnn.Sequential(*layers).cuda()
        if c.world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[c.local_rank])
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
        loss_fn = tnn.MSELoss()
        z = torch.randn(batch, device="cuda")
        x, y = torch.stack([z, z], 1).contiguous(), (z * z).view(batch, 1).contiguous()

        def step():
            opt.zero_grad(set_to_none=True)
            loss_fn(model(x), y).backward()
            opt.step()
        for _ in range(5):
            step()
        c.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        c.barrier()
        return c.max_over_ranks(e0.elapsed_time(e1)) / steps
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def wide_record(c, peaks, hidden=None):
    """BASELINE.json configs[4]: the widened toy model (hidden width 4096 by default), local batch 256 on one GPU
    and 512 / W on W GPUs (demo.py's 512-sample epoch), through the front end; per-kernel medians
    (CUDA events, L2 flushed) on one GPU; the stock-PyTorch incumbent beside it."""
    import torch

    import distributed_training_pytorch_b200 as b200
    from distributed_training_pytorch_b200 import _lib
    H = hidden or int(os.environ.get("DDPB_BENCH_WIDE_HIDDEN", "4096"))
    B = BATCH if c.world == 1 else max(1, DATASET // c.world)
    steps = 30
    torch.manual_seed(0)
    model = b200.DistributedDataParallel(b200.ToyModel(hidden=H).cuda(), device_ids=[c.local_rank])
    opt = b200.optim.Adam(model.parameters(), lr=1e-3)
    loss_fn = b200.nn.MSELoss()
    z = torch.randn(B, device="cuda")
    x, y = torch.stack([z, z], 1).contiguous(), (z * z).view(B, 1).contiguous()

    def step():
        opt.zero_grad(set_to_none=True)
        l = loss_fn(model(x), y)
        l.backward()
        opt.step()
        return l
    for _ in range(5):
        step()
    c.barrier()
    t_begin = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        last = step()
    e1.record()
    c.barrier()
    ms = c.max_over_ranks(e0.elapsed_time(e1)) / steps
    final_loss = float(last.detach().cpu())
    weak = None
    if c.world > 1:                      # weak scaling of the same model: every rank keeps the reference's batch of 256
        zw = torch.randn(BATCH, device="cuda")
        x, y = torch.stack([zw, zw], 1).contiguous(), (zw * zw).view(BATCH, 1).contiguous()
        for _ in range(3):
            step()
        c.barrier()
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        c.barrier()
        wms = c.max_over_ranks(e0.elapsed_time(e1)) / steps
        weak = {"local_batch": BATCH, "global_batch": BATCH * c.world, "ms_per_step": wms,
                "value": BATCH * c.world / (wms * 1e-3), "unit": UNIT, "scaling": "weak"}
    be = model._backend
    P = be.P
    hbm = peaks["hbm_gbs"]
    rec = {"workload": f"widened ToyModel(hidden={H}), Adam lr=1e-3, local batch {B}, {c.world} GPU(s)", "hidden": H,
           "params": P, "local_batch": B, "global_batch": B * c.world, "dtype": "tf32 multiply, fp32 accumulate/storage",
           "steps": steps, "ms_per_step": ms, "value": B * c.world / (ms * 1e-3), "unit": UNIT,
           "allreduce": ("none" if c.world == 1 else "fused peer-window reduce+Adam kernel per bucket" if be.fused_exchange
                         else "ncclAllReduce + fused avg+Adam"),
           "buckets": len(be.buckets), "loss_after": final_loss, "weak": weak,
           "floors_ms": {"hbm_40B_per_param": 40.0 * P / (hbm * 1e9) * 1e3,
                         "nvlink_allreduce": (2.0 * (c.world - 1) / c.world * 4.0 * P / 900e9 * 1e3) if c.world > 1 else 0.0}}
    if c.world == 1:
        L = _lib.lib()
        st = torch.cuda.current_stream().cuda_stream
        w, bias = be._w(2)
        gw, gb = be._g(2)

        def timed(fn, reps=15):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
            for a, b in ev:
                c.flush_buf.fill_(1)
                a.record(); fn(); b.record()
            torch.cuda.synchronize()
            t = sorted(a.elapsed_time(b) for a, b in ev)
            return t[len(t) // 2]
        k = {"fwd": timed(lambda: L.ddpb_linear_fwd(be.act[0].data_ptr(), w, bias, be.act[1].data_ptr(), be.act_t[1].data_ptr(), B, H, H, be.ld, 1, st)),
             "dgrad": timed(lambda: L.ddpb_linear_dgrad(be.dz[0].data_ptr(), w, be.act[0].data_ptr(), be.dz[1].data_ptr(), be.dz_t[1].data_ptr(), B, H, H, be.ld, st)),
             "wgrad": timed(lambda: L.ddpb_linear_wgrad(be.dz_t[0].data_ptr(), be.act_t[0].data_ptr(), gw, gb, B, H, H, be.ld, st)),
             "adam": timed(lambda: be._update(0, P, st, False))}
        wbytes = 4.0 * H * H
        rec["kernels"] = {
            name: {"ms": k[name], "bound": "hbm", "algorithmic_bytes": wbytes, "achieved_gbs": wbytes / (k[name] * 1e-3) / 1e9,
                   "frac": wbytes / (k[name] * 1e-3) / 1e9 / hbm, "tflops": 2.0 * B * H * H / (k[name] * 1e-3) / 1e12}
            for name in ("fwd", "dgrad", "wgrad")}
        rec["kernels"]["adam"] = {"ms": k["adam"], "bound": "hbm", "algorithmic_bytes": 28.0 * P,
                                  "achieved_gbs": 28.0 * P / (k["adam"] * 1e-3) / 1e9, "frac": 28.0 * P / (k["adam"] * 1e-3) / 1e9 / hbm}
        rec["kernels_note"] = ("median of CUDA-event pairs with a 256 MiB L2 flush before each launch (includes the "
                               "launch gap); algorithmic bytes 4*H*H per GEMM (the weight stream), 28*P for Adam")
    c.windows.append((t_begin, time.perf_counter()))
    if hasattr(be, "close"):
        be.close()
    del model, opt, be
    torch.cuda.empty_cache()
    # the incumbent: stock PyTorch on the same GPU(s), same shapes
    inc = {}
    for name, tf32 in (("fp32", False), ("tf32", True)):
        t_begin = time.perf_counter()
        t = incumbent_torch(c, H, B, 20, tf32)
        c.windows.append((t_begin, time.perf_counter()))
        inc[name] = {"ms_per_step": t, "value": B * c.world / (t * 1e-3),
                     "what": f"torch {torch.__version__} nn.Linear/LeakyReLU/MSELoss/optim.Adam"
                             f"{' + torch DDP (NCCL)' if c.world > 1 else ''}, allow_tf32={tf32}"}
    rec["incumbent_torch_gpu"] = inc
    rec["speedup_vs_incumbent"] = {k: v["ms_per_step"] / ms for k, v in inc.items()}
    torch.cuda.empty_cache()
    return rec


def pipeline_record(c):
    """BASELINE.json configs[3]: MultiGPUModel split over two GPUs of one process (rank 0 only, GPUs 0/1;
    needs W >= 2 so that a second GPU exists).  Step time through the front end, the latency split of
    the three launches, and the stock-PyTorch incumbent (the reference's .to(dev1) version)."""
    import torch
    if torch.cuda.device_count() < 2:
        return None
    c.barrier()
    rec = None
    if c.rank == 0:
        import torch.nn as tnn

        import distributed_training_pytorch_b200 as b200
        d0, d1 = 0, 1
        torch.manual_seed(1234)
        # this record runs on rank 0 alone: a one-process pipeline, no collectives
        model = b200.DistributedDataParallel(b200.MultiGPUModel(d0, d1), device_ids=None, process_group=b200.LOCAL_ONLY)
        opt = b200.optim.Adam(model.parameters(), lr=1e-3)
        loss_fn = b200.nn.MSELoss()
        z = torch.randn(BATCH)
        x = torch.stack([z, z], 1).contiguous().to(d0)
        y = (z * z).view(BATCH, 1).contiguous().to(d1)

        def step(read):
            opt.zero_grad(set_to_none=True)
            l = loss_fn(model(x), y)
            l.backward()
            opt.step()
            return float(l.detach().cpu()) if read else None
        for _ in range(20):
            step(True)
        steps = 300
        for d in (d0, d1):
            torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        for _ in range(steps):
            step(True)
        wall_read = (time.perf_counter() - t0) / steps * 1e3
        with torch.cuda.device(d0):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for d in (d0, d1):
                torch.cuda.synchronize(d)
            e0.record()
            for _ in range(steps):
                step(False)
            e1.record()
            for d in (d0, d1):
                torch.cuda.synchronize(d)
            dev_ms = e0.elapsed_time(e1) / steps
        # latency split of one iteration: events around the three launches
        be = model._backend
        split = be.timed_split(x, y, 50) if hasattr(be, "timed_split") else None
        # incumbent: the reference's model, verbatim structure (demo_one_model_multi_gpu.py:17-42)
        torch.manual_seed(1234)
        l0 = tnn.Sequential(tnn.Linear(2, 10), tnn.LeakyReLU(), tnn.Linear(10, 10), tnn.LeakyReLU()).to(d0)
        l1 = tnn.Sequential(tnn.Linear(10, 10), tnn.LeakyReLU(), tnn.Linear(10, 10), tnn.LeakyReLU(), tnn.Linear(10, 1)).to(d1)
        topt = torch.optim.Adam(list(l0.parameters()) + list(l1.parameters()), lr=1e-3)
        tl = tnn.MSELoss()

        def tstep():
            topt.zero_grad(set_to_none=True)
            loss = tl(l1(l0(x).to(d1)), y)
            loss.backward()
            topt.step()
            return float(loss.detach().cpu())
        for _ in range(20):
            tstep()
        t0 = time.perf_counter()
        for _ in range(100):
            tstep()
        inc_ms = (time.perf_counter() - t0) / 100 * 1e3
        rec = {"workload": "demo_one_model_multi_gpu.py: MultiGPUModel (2->10->10 | 10->10->10->1) on GPUs 0/1 of one process, "
                           "B=256, Adam", "ms_per_step_device": dev_ms, "ms_per_step_with_loss_read": wall_read,
               "value": BATCH / (wall_read * 1e-3), "unit": UNIT, "launches_per_step": 3, "latency_split_us": split,
               "incumbent_torch_gpu": {"ms_per_step_with_loss_read": inc_ms, "value": BATCH / (inc_ms * 1e-3),
                                       "what": f"torch {torch.__version__} modules with .to(dev1) between the stages, loss read every step"},
               "speedup_vs_incumbent": inc_ms / wall_read,
               "timing": "device: one CUDA-event pair on GPU 0 around 300 steps without loss reads; with_loss_read: host wall "
                         "clock around 300 steps that read the loss every step (the reference's loop logs it every step)"}
    # the other ranks wait on the HOST (gloo): an NCCL barrier would park a spinning kernel on GPU 1, which
    # rank 0's second stage is using
    import torch.distributed as dist
    dist.barrier(group=c.side)
    return rec


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
    K, W = args.steps, max(args.warmup, 3)
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
    c = Ctx()
    c.world, c.rank, c.local_rank, c.side, c.barrier, c.max_over_ranks, c.flush_buf = world, rank, local_rank, side, barrier, max_over_ranks, flush_buf
    c.windows = windows = []

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peaks = json.load(open(peaks_path))
        peak, peak_src = peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs, burst)"
    else:
        peaks = {"hbm_gbs": 6650.0}
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"

    # ---- parity first (outside every timed region): a bench line without it is not printed green
    parity = None if args.no_parity else golden_parity(c)

    def build(device_resident, n_data=DATASET):
        torch.manual_seed(1234)
        mx, my, data = ToyModel(), ToyModel(), ToyData(n_data)
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

    def one_step(models, opts, batch, read):
        """One iteration of demo.py's loop.  read: None (no loss read), "local" (loss.detach().cpu()),
        "fused" (the all-rank mean out of the gradient exchange), "gloo" (the reference's two host collectives)."""
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
        if read is None:
            eng.flush()
            return None
        if read == "local" or world == 1:
            return [l.detach().cpu() for l in losses]
        if read == "fused":
            return [l.detach().global_mean() for l in losses]
        out = []
        bs = target.size(0)
        for l in losses:
            t = l.detach().cpu() * bs
            dist.all_reduce(t, group=side, op=dist.ReduceOp.SUM)
            out.append(t / (bs * world))
        return out

    clock = ClockSampler(local_rank)
    clock.start()

    def cold_leg(n_data, steps):
        """HBM-resident inputs, per-step events, L2 flushed between timed steps."""
        model_X, model_Y, sampler, loader, opts = build(True, n_data)
        batches = iteration_stream(sampler, loader)
        for _ in range(W - 1):                            # the W-th warm-up step follows the barrier below
            one_step((model_X, model_Y), opts, next(batches), None)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        stops = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        gc.collect()                                      # no cyclic-garbage pass of the interpreter inside the timed steps
        barrier()                                         # all ranks enter the timed region together (a rank that is late
        # into step 0 is time its peers spend waiting inside the kernel); the last warm-up step runs after the host
        # barrier and lines the ranks up to microseconds: the gradient exchange itself makes each wait for all
        one_step((model_X, model_Y), opts, next(batches), None)
        launches0 = eng.launches
        t_begin = time.perf_counter()
        for i in range(steps):
            flush_buf.fill_(i & 0xff)                     # evict L2 (outside the timed region)
            batch = next(batches)
            starts[i].record()
            one_step((model_X, model_Y), opts, batch, None)
            stops[i].record()
        barrier()
        windows.append((t_begin, time.perf_counter()))
        launches = eng.launches - launches0
        raw = [s.elapsed_time(e) for s, e in zip(starts, stops)]
        worst = max(range(steps), key=raw.__getitem__)
        per_step = sorted(raw)
        ms = max_over_ranks(sum(per_step))
        median = per_step[len(per_step) // 2]
        stats = {"median_ms": max_over_ranks(median), "max_ms": max_over_ranks(per_step[-1]),
                 "steps_over_5x_median": int(max_over_ranks(float(sum(1 for t in per_step if t > 5 * median)))),
                 "slowest_step_index_rank0": worst}
        return ms, launches, (model_X, model_Y, sampler, loader, opts, batches), stats

    # The device comes out of idle (clocks, memory and PCIe link power states): 0.3 s of plain memory traffic before
    # the first timed leg, so that a short run (--steps 20) measures the same machine state as a long one.  These are
    # not steps of the workload; every leg still does its own W warm-up steps.
    t_begin = time.perf_counter()
    while time.perf_counter() - t_begin < 0.3:
        for _ in range(50):
            flush_buf.fill_(1)
        torch.cuda.synchronize()

    # ---- value: strong scaling of demo.py's 512-sample epoch
    cold_ms, launches, kept, cold_stats = cold_leg(DATASET, K)
    model_X, model_Y, sampler, loader, opts, batches = kept

    # ---- hot_l2: same steps back to back, straight through the C ABI as ONE captured CUDA graph
    # of K fused launches (each node carries its sampler row; step counters live on the device),
    # so the host is not the bottleneck.  With several ranks the gradient exchange is inside
    # the kernel (peer windows), so the same graph works; the NCCL path goes through the front end.
    direct = None
    fused = world == 1 or model_X._backend.fused_exchange
    if fused:
        import ctypes
        eng.flush()
        n_rows = 1024
        per_epoch = len(sampler) // lb
        rows = torch.from_numpy(sampler.epochs_i32(0, n_rows // per_epoch).reshape(n_rows, lb)).cuda()
        jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
        for i, m in enumerate((model_X, model_Y)):
            j = _lib.ToyJob.from_buffer_copy(m._backend.job)
            j.loss_ring = m._backend.loss_dev.data_ptr()          # nobody reads losses in this loop: keep them in HBM
            j.global_loss_ring = None
            j.x, j.y, j.idx = loader._dev[0].data_ptr(), loader._dev[1].data_ptr(), rows.data_ptr()
            j.idx_rows, j.idx_stride, j.B, j.mode, j.n_data = n_rows, lb, lb, _lib.MODE_STEP, DATASET
            jobs[i] = j
        stream = torch.cuda.current_stream().cuda_stream
        L = _lib.lib()
        toy_step = L.ddpb_toy_step
        graph = ctypes.c_void_p()
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
            one_step((model_X, model_Y), opts, next(batches), None)
    ev1.record()
    barrier()
    windows.append((t_begin, time.perf_counter()))
    hot_ms = max_over_ranks(ev0.elapsed_time(ev1))
    del kept, model_X, model_Y, sampler, loader, opts, batches

    # ---- weak scaling (secondary): the dataset grows with W (256 * W samples), so every rank
    # keeps B = 256 and the exchange is the only thing that changes with W
    weak = None
    if world > 1:
        Kw = min(K, 500)
        weak_ms, _, kept, _ = cold_leg(BATCH * world, Kw)
        weak = {"value": Kw * BATCH * world / (weak_ms * 1e-3), "unit": UNIT, "ms_per_step": weak_ms / Kw, "steps": Kw,
                "dataset": BATCH * world, "local_batch": BATCH, "global_batch": BATCH * world, "scaling": "weak",
                "timing": "as `value`: CUDA events per step, L2 flushed, max over ranks"}
        del kept

    # ---- e2e: pinned host batches through the public front end, the step's loss read back every step.
    # W > 1: the loss that is read is the one demo.py logs, the mean over all ranks -- by default out of
    # the gradient exchange (bucket element 371), and as `e2e_gloo_loss` the reference's way (local
    # loss.cpu() + two gloo all-reduces per step, demo.py:112-121).
    def e2e_leg(read, steps, n_data=DATASET):
        model_X, model_Y, sampler, loader, opts = build(False, n_data)
        batches = iteration_stream(sampler, loader)
        for _ in range(W - 1):
            one_step((model_X, model_Y), opts, next(batches), read)
        barrier()
        one_step((model_X, model_Y), opts, next(batches), read)    # the W-th warm-up step: lines the ranks up after the host barrier
        t_begin = time.perf_counter()
        total, wall, last = 0.0, 0.0, None
        for i in range(steps):
            flush_buf.fill_(i & 0xff)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ev0.record()
            last = one_step((model_X, model_Y), opts, next(batches), read)
            ev1.record()
            ev1.synchronize()
            wall += time.perf_counter() - t0
            total += ev0.elapsed_time(ev1)
        barrier()
        windows.append((t_begin, time.perf_counter()))
        return max_over_ranks(total), max_over_ranks(wall * 1e3), last, (model_X, model_Y, opts, batches)

    read_mode = "local" if world == 1 else ("gloo" if os.environ.get("DDPB_FUSED_LOSS", "1") == "0" else "fused")
    e2e_ms, e2e_wall_ms, last, kept = e2e_leg(read_mode, K)
    e2e_gloo = None
    if world > 1 and read_mode == "fused":
        Kg = min(K, 300)
        g_ms, _, _, kept2 = e2e_leg("gloo", Kg)
        e2e_gloo = {"value": Kg * gb / (g_ms * 1e-3), "unit": UNIT, "ms_per_step": g_ms / Kg, "steps": Kg,
                    "loss_reduction": "loss.detach().cpu() * B + 2 gloo all_reduce per step (reference demo.py:112-121)"}
        del kept2
    e2e_weak = None
    if world > 1:                      # the same end-to-end loop with B = 256 on every rank (dataset 256 * W)
        Kw = min(K, 500)
        w_ms, _, _, kept2 = e2e_leg(read_mode, Kw, BATCH * world)
        e2e_weak = {"value": Kw * BATCH * world / (w_ms * 1e-3), "unit": UNIT, "ms_per_step": w_ms / Kw, "steps": Kw,
                    "local_batch": BATCH, "global_batch": BATCH * world, "scaling": "weak"}
        del kept2

    # keep the GPU busy long enough for nvidia-smi to see the clocks this workload runs at
    model_X, model_Y, opts, batches = kept
    t_begin = time.perf_counter()
    while time.perf_counter() - t_begin < 1.0:
        for _ in range(200):
            one_step((model_X, model_Y), opts, next(batches), None)
        torch.cuda.synchronize()
    windows.append((t_begin, time.perf_counter()))
    del kept, model_X, model_Y, opts, batches

    # ---- secondary configurations (extra keys; same clock sampler)
    wide = None if args.no_wide else wide_record(c, peaks)
    # one GPU: the same record at H = 8192, where the layers run on the CTA-pair kernels (csrc/gemm_pair.cu)
    wide_8192 = None if (args.no_wide or world > 1) else wide_record(c, peaks, 8192)
    pipeline = None if (args.no_pipeline or world < 2) else pipeline_record(c)
    barrier()
    incumbent = None
    if not args.no_incumbent:
        t_begin = time.perf_counter()
        t = incumbent_torch(c, H, lb, 100, False)
        windows.append((t_begin, time.perf_counter()))
        incumbent = {"ms_per_step": 2 * t, "value": gb / (2 * t * 1e-3), "unit": UNIT,
                     "what": f"torch {torch.__version__} eager: 2 x (nn.Linear/LeakyReLU MLP, MSELoss, backward, optim.Adam)"
                             f"{' under torch DDP (NCCL)' if world > 1 else ''}, resident inputs, no loss read; one model "
                             "timed, doubled for demo.py's two", "compare_with": "hot_l2 / value (resident inputs)"}
    clocks = clock.stop(windows)

    # algorithmic bytes per launch: per model 4P (bucket) + 28P (Adam) + 48*B*H + 8*B (activations)
    alg_bytes = 2 * (4 * P + 28 * P + 48 * lb * H + 8 * lb)
    kernel_ms = cold_ms / K                             # the same timing `value` is computed from
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    hot_kernel_ms = hot_ms / K if fused else None
    traffic, traffic_src = ncu_traffic() if lb == BATCH else (None, None)

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        iters = 4000                                        # ~10-15 s of single-core work
        secs = run_cpu_reference(1, iters, 20)
        cpu_baseline = {"value": iters * gb / secs, "unit": UNIT, "cores": 1, "kind": "port",
                        "sample": f"{iters} iterations of the same two-model loop (B=256), 1 gloo rank x 1 thread, "
                                  "stock torch CPU path (oracle/torch_ref.py, bit-identical to the reference run)"}

    if rank == 0:
        be_fused = world > 1 and fused
        line = {"metric": METRIC, "value": K * gb / (cold_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": cold_ms / K, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload(world),
                "value_per_step": dict(cold_stats, note="per-step CUDA-event times of the `value` leg, max over ranks; `value` "
                                                        "uses their SUM, so a single host stall on any rank (the peers wait for "
                                                        "it inside the kernel) counts in full"),
                "notes": {"timing": "CUDA events per step; L2 flushed (256 MiB fill) between timed steps; max over ranks",
                          "parallelism": f"dp{world}",
                          "allreduce": ("none (single rank)" if world == 1 else
                                        "fused one-shot exchange over NVLink peer windows inside the step kernel"
                                        if be_fused else "ncclAllReduce per bucket + fused avg+Adam"),
                          "scaling_note": "one 512-sample epoch is the fixed job: 2 steps at W=1, 1 step at W>=2",
                          "preconditioning": "0.3 s of plain device memory traffic (no workload steps) before the first "
                                             "timed leg brings the device out of its idle power states; every leg then "
                                             "does its own W warm-up steps, the last of them after the barrier"},
                "hot_l2": {"value": K * gb / (hot_ms * 1e-3), "unit": UNIT, "ms_per_step": hot_ms / K,
                           "timing": "one CUDA-event pair around K back-to-back steps, no flush"
                                     + ("; the K launches are one captured CUDA graph" if fused else "")},
                "e2e": {"value": K * gb / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K,
                        "host_wall_ms_per_step": e2e_wall_ms / K,
                        "h2d_bytes_per_step": lb * 12, "d2h_bytes_per_step": 8 if read_mode == "local" else 16,
                        "h2d": "the step kernel loads the batch (x | y rows) from the loader's pinned staging block over "
                               "PCIe (mapped host memory); no copy-engine operation",
                        "d2h": "the kernel stores each model's loss into a pinned ring the host polls",
                        "loss_read": {"local": "loss.detach().cpu() of both models, every step",
                                      "fused": "loss.detach().global_mean() of both models, every step: the all-rank mean "
                                               "out of the gradient exchange (no host collective)",
                                      "gloo": "loss.detach().cpu() + 2 gloo all_reduce, every step"}[read_mode]},
                "e2e_gloo_loss": e2e_gloo,
                "e2e_weak": e2e_weak,
                "weak": weak,
                "gpu_launches": launches,
                "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic,
                             "traffic_source": (f"{traffic_src}: dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                                                "ncu --set full (cold caches)") if traffic_src else None,
                             "kernel": "toy_step_kernel", "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "launch_ms": kernel_ms, "launch_ms_source": "the per-step CUDA events `value` is computed from "
                                                                         "(cold L2, one launch per step)",
                             "hot": ({"launch_ms": hot_kernel_ms, "achieved": alg_bytes / (hot_kernel_ms * 1e-3) / 1e9,
                                      "frac": alg_bytes / (hot_kernel_ms * 1e-3) / 1e9 / peak,
                                      "source": "K steps replayed as one CUDA graph, warm L2"} if hot_kernel_ms else None),
                             "note": "latency-bound by construction: 2 clusters x 8 CTAs, 273 KB and 1 MFLOP per launch"},
                "cpu_baseline": cpu_baseline,
                "parity": parity,
                "incumbent_torch_gpu": incumbent,
                "wide": wide,
                "wide_h8192": wide_8192,
                "pipeline": pipeline,
                "loss_after_bench": [float(v) for v in last] if last else None}
        emit(line)
    eng.destroy_comm()
    dist.destroy_process_group()
    if parity is not None and parity.get("ok") is False:
        raise SystemExit("golden parity check FAILED (see the `parity` record)")


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
    ap.add_argument("--no-parity", dest="no_parity", action="store_true", help="skip the golden-fixture check")
    ap.add_argument("--no-wide", dest="no_wide", action="store_true", help="skip the widened-model record")
    ap.add_argument("--no-pipeline", dest="no_pipeline", action="store_true", help="skip the model-parallel record")
    ap.add_argument("--no-incumbent", dest="no_incumbent", action="store_true", help="skip the stock-PyTorch GPU records")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
