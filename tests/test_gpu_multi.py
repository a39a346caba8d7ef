"""Multi-GPU parity: the drop-in demo.py under torchrun (one process per GPU, NCCL rendezvous)
against the reference's own W-rank run (tests/golden/ref_w{W}.npz, seed 1234).

Both gradient paths are covered: the fused one-shot exchange over NVLink peer windows inside
the step kernel (default) and ncclAllReduce of the flat bucket + the fused average+Adam pass.
Tolerances as in test_gpu_toy_step.py (fp32 path): loss curve rel 2e-5 and parameters abs 2e-5
after 200 steps; replicas must stay BIT-identical (fixed summation order on every rank)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_demo(tmp_path, world, allreduce, extra_args=("--backend=nccl", "--torchrun", "--dry_run"), **extra_env):
    env = dict(os.environ, DDPB_ITERATIONS="200", DDPB_SEED="1234", DDPB_NO_WANDB="1", DDPB_ALLREDUCE=allreduce,
               OMP_NUM_THREADS="1", **extra_env)
    os.makedirs(tmp_path, exist_ok=True)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "_demo_worker.py"), str(tmp_path)] + list(extra_args)
    proc = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, (proc.stdout + proc.stderr)[-4000:]
    return [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]


def _worlds():
    n = torch.cuda.device_count()
    return [w for w in (2, 4, 8) if w <= n]


@pytest.mark.parametrize("allreduce", ["p2p", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_demo_matches_reference_run(tmp_path, golden_dir, world, allreduce):
    if world not in _worlds():
        pytest.skip(f"needs {world} GPUs")
    ranks = _run_demo(tmp_path, world, allreduce)
    g = np.load(os.path.join(golden_dir, f"ref_w{world}.npz"))
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    losses = ranks[0]["losses"]
    assert losses.shape == (200, 2)
    assert np.max(np.abs(losses - ref) / ref) <= 2e-5
    assert np.max(np.abs(ranks[0]["px"] - g["mid_X"])) <= 2e-5 and np.max(np.abs(ranks[0]["py"] - g["mid_Y"])) <= 2e-5
    for r in ranks[1:]:
        assert np.array_equal(r["px"], ranks[0]["px"]) and np.array_equal(r["py"], ranks[0]["py"])


def _check_against_reference_run(ranks, g):
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    losses = ranks[0]["losses"]
    assert losses.shape == (200, 2)
    assert np.max(np.abs(losses - ref) / ref) <= 2e-5
    assert np.max(np.abs(ranks[0]["px"] - g["mid_X"])) <= 2e-5 and np.max(np.abs(ranks[0]["py"] - g["mid_Y"])) <= 2e-5
    for r in ranks[1:]:
        assert np.array_equal(r["px"], ranks[0]["px"]) and np.array_equal(r["py"], ranks[0]["py"])


@pytest.mark.parametrize("allreduce", ["p2p", "nccl"])
@pytest.mark.parametrize("world", [2, 4])
def test_ranks_sharing_one_gpu_match_reference_run(tmp_path, golden_dir, world, allreduce):
    """W processes on ONE GPU (DDPB_SHARE_GPUS=1, gloo rendezvous -- NCCL refuses two ranks per device):
    the W-rank arithmetic of demo.py without needing W GPUs.  `p2p` is the fused in-kernel exchange
    through IPC-mapped peer windows (the processes' kernels take turns on the GPU, each polling for
    the other's tickets); `nccl` here means "all-reduce the bucket outside the kernel, then the fused
    average+Adam pass", with the bucket carried by the gloo group.  Same golden fixtures, same
    tolerances as the one-GPU-per-rank test above."""
    ranks = _run_demo(tmp_path, world, allreduce, ("--backend=gloo", "--torchrun", "--dry_run"), DDPB_SHARE_GPUS="1",
                      CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
    _check_against_reference_run(ranks, np.load(os.path.join(golden_dir, f"ref_w{world}.npz")))


@pytest.mark.parametrize("allreduce", ["p2p", "nccl"])
def test_fused_loss_equals_the_gloo_reduction(tmp_path, golden_dir, allreduce):
    """DDPB_FUSED_LOSS=1: the logged all-rank mean loss comes out of the gradient exchange (bucket
    element 371) one iteration late instead of two gloo all-reduces per step (demo.py:112-121).  Same
    training trajectory, and the logged values agree with the gloo-reduced ones to 1e-6 (relative)."""
    one_gpu = dict(DDPB_SHARE_GPUS="1", CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
    args = ("--backend=gloo", "--torchrun", "--dry_run")
    if torch.cuda.device_count() >= 2:
        one_gpu, args = {}, ("--backend=nccl", "--torchrun", "--dry_run")
    gloo = _run_demo(os.path.join(tmp_path, "gloo"), 2, allreduce, args, **one_gpu)
    fused = _run_demo(os.path.join(tmp_path, "fused"), 2, allreduce, args, DDPB_FUSED_LOSS="1", **one_gpu)
    assert fused[0]["losses"].shape == (200, 2)
    assert np.max(np.abs(fused[0]["losses"] - gloo[0]["losses"]) / gloo[0]["losses"]) <= 1e-6
    for a, b in zip(gloo, fused):
        assert np.array_equal(a["px"], b["px"]) and np.array_equal(a["py"], b["py"])
    _check_against_reference_run(fused, np.load(os.path.join(golden_dir, "ref_w2.npz")))


def test_single_rank_demo_entry_point(tmp_path, golden_dir):
    ranks = _run_demo(tmp_path, 1, "p2p")
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    assert np.max(np.abs(ranks[0]["losses"] - ref) / ref) <= 2e-5
    assert np.max(np.abs(ranks[0]["px"] - g["mid_X"])) <= 2e-5


def test_standard_dataloader_mode(tmp_path, golden_dir):
    """--dataloader standard: every rank walks the whole dataset in order (demo.py:149-154); pinned
    by the reference's own W=2 run in that mode (tests/golden/ref_w2_standard.npz)."""
    if 2 not in _worlds():
        pytest.skip("needs 2 GPUs")
    ranks = _run_demo(tmp_path, 2, "p2p", ("--backend=nccl", "--torchrun", "--dry_run", "--dataloader=standard"))
    assert np.array_equal(ranks[0]["px"], ranks[1]["px"]) and np.array_equal(ranks[0]["py"], ranks[1]["py"])
    g = np.load(os.path.join(golden_dir, "ref_w2_standard.npz"))
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    assert np.max(np.abs(ranks[0]["losses"] - ref) / ref) <= 2e-5
    assert np.max(np.abs(ranks[0]["px"] - g["mid_X"])) <= 2e-5 and np.max(np.abs(ranks[0]["py"] - g["mid_Y"])) <= 2e-5


def test_mpiexec_entry_point_with_the_compat_launcher(tmp_path, golden_dir):
    """demo_assume_started_with_mpiexec.py started by tools/mpiexec_compat.py (no MPI in the image):
    the logged values are each rank's LOCAL loss (the reference reduces temporaries), the model
    trajectory is the same as demo.py's."""
    world = 2 if 2 in _worlds() else 1
    env = dict(os.environ, DDPB_ITERATIONS="200", DDPB_SEED="1234", DDPB_TEST_ENTRY="mpiexec", OMP_NUM_THREADS="1",
               TASKS_PER_NODE=str(world))
    cmd = [sys.executable, os.path.join(ROOT, "tools", "mpiexec_compat.py"), "-n", str(world), sys.executable,
           os.path.join(ROOT, "tests", "_demo_worker.py"), str(tmp_path), "--backend=nccl"]
    proc = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, (proc.stdout + proc.stderr)[-4000:]
    g = np.load(os.path.join(golden_dir, f"ref_w{world}.npz"))
    ranks = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    assert np.max(np.abs(ranks[0]["px"] - g["mid_X"])) <= 2e-5 and np.max(np.abs(ranks[0]["py"] - g["mid_Y"])) <= 2e-5
    mean = np.mean([r["losses"] for r in ranks], axis=0)          # equal shards: mean of local losses = logged value
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    assert np.max(np.abs(mean - ref) / ref) <= 2e-5


def _widened_run(tmp_path, world, cap_mb, allreduce, **extra_env):
    """The widened model trained on `world` GPUs against an fp32 PyTorch model trained on the
    concatenated batch (equal shards: the average of the shard gradients IS the gradient of the whole
    batch), within the tf32 tolerance."""
    H, B, steps = 128, 64, 12
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "_wide_worker.py"), str(tmp_path),
           str(H), str(B), str(steps), str(cap_mb)]
    proc = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600,
                          env=dict(os.environ, OMP_NUM_THREADS="1", DDPB_ALLREDUCE=allreduce, **extra_env))
    assert proc.returncode == 0, (proc.stdout + proc.stderr)[-4000:]
    ranks = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    r0 = ranks[0]
    assert int(r0["fused"]) == int(allreduce == "p2p")                  # no silent fallback to the other path
    assert np.any(r0["exp_avg"] != 0)
    for r in ranks[1:]:
        assert np.array_equal(r0["exp_avg"], r["exp_avg"])              # gathered optimizer state
        assert np.array_equal(r0["init"], r["init"])                    # broadcast from rank 0
        assert np.array_equal(r0["final"], r["final"])                  # replicas stay identical
    # fp32 reference on the concatenated batch
    torch.backends.cuda.matmul.allow_tf32 = False
    dims = (2, H, H, H, H, 1)
    layers = []
    for i in range(5):
        layers.append(torch.nn.Linear(dims[i], dims[i + 1]))
        if i < 4:
            layers.append(torch.nn.LeakyReLU())
    twin = torch.nn.Sequential(*layers).cuda()
    flat, off = torch.from_numpy(r0["init"]).cuda(), 0
    with torch.no_grad():
        for p in twin.parameters():
            p.copy_(flat[off:off + p.numel()].view_as(p)); off += p.numel()
    opt = torch.optim.Adam(twin.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(7)
    for step in range(steps):
        z = torch.randn(world * B, generator=g)
        y = (0.5 * torch.randn(world * B, generator=g) + z * z).view(-1, 1).cuda()
        x = torch.stack([z, z], 1).cuda()
        opt.zero_grad()
        per_sample = (twin(x) - y) ** 2
        per_sample.mean().backward()
        opt.step()
        tol = 4e-3 if step == 0 else 5e-2          # same weights at step 0; afterwards the two Adam runs drift apart
        for r in range(world):
            shard = per_sample[r * B:(r + 1) * B].mean().item()
            assert abs(ranks[r]["losses"][step] - shard) <= tol * shard, (step, r)
    final_t = torch.cat([p.detach().reshape(-1) for p in twin.parameters()]).cpu().numpy()
    delta = np.abs(r0["final"] - final_t)
    assert delta.max() <= 2 * steps * 1e-3 and delta.mean() <= 5e-4
    return r0


@pytest.mark.parametrize("allreduce", ["p2p", "nccl"])
@pytest.mark.parametrize("cap_mb", [25, 0.05])
def test_widened_model_data_parallel(tmp_path, cap_mb, allreduce):
    """Widened model on 2 GPUs: rank 0's weights are broadcast, gradients are averaged bucket by
    bucket (one bucket per layer at 25 MiB, many at 0.05 MiB) by the fused peer-window kernel or by
    ncclAllReduce, replicas stay bit-identical, and the run tracks the fp32 model."""
    if 2 not in _worlds():
        pytest.skip("needs 2 GPUs")
    r0 = _widened_run(tmp_path, 2, cap_mb, allreduce)
    assert int(r0["buckets"]) == (5 if cap_mb == 25 else 8)          # 16512-float hidden layers cut at 13104 floats


@pytest.mark.parametrize("allreduce,streams", [("p2p", "1"), ("p2p", "2"), ("nccl", "1")])
def test_widened_model_default_bucket_plan(tmp_path, allreduce, streams):
    """No bucket cap named: three buckets (the thin first / last layers ride with their neighbours), on one
    or two communication streams (DDPB_COMM_STREAMS; with two, the step counter is advanced once after all
    buckets).  Same checks as above: broadcast, bit-identical replicas, gathered optimizer state, fp32 tracking."""
    if 2 not in _worlds():
        pytest.skip("needs 2 GPUs")
    r0 = _widened_run(tmp_path, 2, 0, allreduce, DDPB_COMM_STREAMS=streams)
    assert int(r0["buckets"]) == 3


@pytest.mark.parametrize("world", [4, 8])
def test_widened_model_fused_exchange_more_ranks(tmp_path, world):
    """The fused bucket kernel with 4 and 8 ranks (shards of a quarter / an eighth of each bucket, the
    scalar tail of the last layer's bucket on the last rank)."""
    if world not in _worlds():
        pytest.skip(f"needs {world} GPUs")
    _widened_run(tmp_path, world, 0.05, "p2p")


@pytest.mark.parametrize("devices", [(0, 1), (0, 0)])
@pytest.mark.parametrize("mode", [1, 0])
def test_bucket_reduce_adam_bit_exact_against_the_oracle(oracle, native_lib, mode, devices):
    """ddpb_bucket_reduce_adam through the C ABI, two GPUs driven from ONE process (plain device
    buffers as the two windows, peer access enabled -- or both "ranks" on GPU 0, their kernels on two
    streams, which needs no second GPU): after three steps the parameters of both ranks,
    and each rank's shard of the Adam state, are BIT-identical to the CPU oracle's `sum in rank
    order -> x 1/W -> Adam` (oracle/ddp_oracle.c), nothing outside the bucket is touched, and the
    state of the shards a rank does not own stays zero."""
    import ctypes
    from distributed_training_pytorch_b200 import _lib
    if devices[0] != devices[1]:
        if torch.cuda.device_count() < 2:
            pytest.skip("needs 2 GPUs")
        _lib.check(native_lib.ddpb_enable_peer(0, 1), "ddpb_enable_peer")
    streams = [torch.cuda.Stream(device=d) for d in devices]
    W, P, start, n = 2, 10007, 16, 9987                    # n % 4 == 3: the scalar tail is rank 1's
    up = lambda b: (b + 255) // 256 * 256
    grads_off = up(_lib.BUCKET_FLAGS_BYTES)
    params_off = grads_off + up(4 * (P + 4))
    total = params_off + up(4 * P)
    rng = np.random.default_rng(3)
    p_ref = rng.standard_normal(P).astype(np.float32)
    m_ref, v_ref = np.zeros(P, np.float32), np.zeros(P, np.float32)
    win, grads, params, m, v, state, table, jobs = [], [], [], [], [], [], [], []
    for d in range(W):
        gpu = f"cuda:{devices[d]}"
        with torch.cuda.device(devices[d]):
            w = torch.zeros(total, dtype=torch.uint8, device=gpu)
            win.append(w)
            grads.append(w[grads_off:grads_off + 4 * P].view(torch.float32))
            params.append(w[params_off:params_off + 4 * P].view(torch.float32))
            params[d].copy_(torch.from_numpy(p_ref))
            m.append(torch.zeros(P, device=gpu)); v.append(torch.zeros(P, device=gpu))
            state.append(torch.zeros(4, dtype=torch.int64, device=gpu))
            table.append(torch.from_numpy(_lib.adam_bias_table(1e-3, 0.9, 0.999)).to(gpu))
    for d in range(W):
        j = _lib.BucketJob()
        for q in range(W):
            j.window[q] = win[q].data_ptr()
        j.grads_off, j.params_off = grads_off, params_off
        j.m, j.v, j.bias_table, j.state = m[d].data_ptr(), v[d].data_ptr(), table[d].data_ptr(), state[d].data_ptr()
        j.start, j.n, j.bias_table_n = start, n, table[d].shape[0]
        j.world, j.rank, j.bucket_index, j.mode, j.advance = W, d, 5, mode, 1
        j.beta1, j.beta2, j.eps, j.grad_scale = 0.9, 0.999, 1e-8, 1.0 / W
        jobs.append(j)
    for step in range(1, 4):
        g = [(rng.standard_normal(P) * 10 ** rng.uniform(-4, 1)).astype(np.float32) for _ in range(W)]
        total_g = g[0] + g[1]                                                # float32, rank order
        for d in range(W):
            grads[d].copy_(torch.from_numpy(g[d] if mode else total_g))    # mode 0: already summed in place
        for d in set(devices):
            torch.cuda.synchronize(d)
        for d in range(W):                                                   # both kernels in flight: they wait for each other
            with torch.cuda.device(devices[d]):
                jobs[d].ticket = step
                _lib.check(native_lib.ddpb_bucket_reduce_adam(ctypes.byref(jobs[d]), streams[d].cuda_stream),
                           "ddpb_bucket_reduce_adam")
        for d in set(devices):
            torch.cuda.synchronize(d)
        sl = slice(start, start + n)
        pb, mb, vb = p_ref[sl].copy(), m_ref[sl].copy(), v_ref[sl].copy()
        oracle.adam_step(pb, total_g[sl] * np.float32(1.0 / W), mb, vb, step)
        p_ref[sl], m_ref[sl], v_ref[sl] = pb, mb, vb
    n4 = n >> 2
    per = (n4 + W - 1) // W
    owned = [slice(start, start + 4 * per), slice(start + 4 * per, start + n)]
    for d in range(W):
        assert np.array_equal(params[d].cpu().numpy(), p_ref), d            # whole buffer: bucket updated, rest untouched
        got_m, got_v = m[d].cpu().numpy(), v[d].cpu().numpy()
        assert np.array_equal(got_m[owned[d]], m_ref[owned[d]]) and np.array_equal(got_v[owned[d]], v_ref[owned[d]])
        other = owned[1 - d]
        assert not got_m[other].any() and not got_v[other].any()
        assert not got_m[:start].any() and not got_m[start + n:].any()
        st = state[d].cpu().numpy()
        assert st[0] == 3 and st[2] == 0                                    # three steps, no peer time-out
