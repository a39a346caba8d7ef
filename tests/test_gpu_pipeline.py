"""The model-parallel configuration (BASELINE.json configs[3], demo_one_model_multi_gpu.py):
stage kernel through the C ABI against the oracle on one GPU, then the two-GPU pipeline through
the front end and through the drop-in entry point.  The split model is mathematically the full
toy model, so the oracle is oracle_mlp_loss_grad / oracle_segment_backward on the concatenated
parameters.  Tolerances: forward bit-exact, gradients 2e-6 of max|g|, trajectories 1e-5."""
import ctypes
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

from distributed_training_pytorch_b200 import _lib
from oracle import TOY_DIMS

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from gpu_util import bias_table, dev, new_state  # noqa: E402


def _seg_job(params, dims, act_last, has_loss, B):
    P = int(sum(dims[i + 1] * dims[i] + dims[i + 1] for i in range(len(dims) - 1)))
    bufs = dict(p=dev(params), g=torch.zeros(P + 1, device="cuda"), m=torch.zeros(P, device="cuda"),
                v=torch.zeros(P, device="cuda"), s=new_state(), ring=torch.zeros(8, device="cuda"), t=bias_table())
    j = _lib.SegJob()
    j.params, j.grads, j.m, j.v, j.state = (bufs[k].data_ptr() for k in ("p", "g", "m", "v", "s"))
    j.bias_table, j.bias_table_n = bufs["t"].data_ptr(), bufs["t"].shape[0]
    j.loss_ring, j.loss_ring_n = bufs["ring"].data_ptr(), 8
    j.B, j.nl, j.act_last, j.has_loss, j.opt = B, len(dims) - 1, int(act_last), int(has_loss), _lib.OPT_ADAM
    for i, d in enumerate(dims):
        j.dims[i] = d
    j.lr, j.beta1, j.beta2, j.eps = 1e-3, 0.9, 0.999, 1e-8
    return j, bufs, P


def _launch(native_lib, j):
    _lib.check(native_lib.ddpb_segment_step(ctypes.byref(j), torch.cuda.current_stream().cuda_stream), "ddpb_segment_step")


@pytest.mark.parametrize("B", [1, 77, 256, 300, 1000])
def test_stage_kernels_against_oracle(oracle, native_lib, B):
    rng = np.random.default_rng(B)
    dims0, dims1 = (2, 10, 10), (10, 10, 10, 1)
    p0 = (rng.standard_normal(140) * 0.5).astype(np.float32)
    p1 = (rng.standard_normal(231) * 0.5).astype(np.float32)
    x = rng.standard_normal((B, 2)).astype(np.float32)
    y = rng.standard_normal((B, 1)).astype(np.float32)
    # stage 0 forward
    j0, b0, _ = _seg_job(p0, dims0, True, False, B)
    dx, act = dev(x), torch.zeros(B, 10, device="cuda")
    j0.x, j0.out, j0.mode = dx.data_ptr(), act.data_ptr(), _lib.MODE_FWD
    _launch(native_lib, j0)
    act_ref = oracle.forward(p0, dims0, x, act_last=True)
    assert np.array_equal(act.cpu().numpy(), act_ref)
    # stage 1: loss + backward, gradient w.r.t. the activation
    j1, b1, P1 = _seg_job(p1, dims1, False, True, B)
    dy, dact = dev(y), torch.zeros(B, 10, device="cuda")
    j1.x, j1.y, j1.d_in, j1.mode = act.data_ptr(), dy.data_ptr(), dact.data_ptr(), _lib.MODE_GRAD
    _launch(native_lib, j1)
    full = np.concatenate([p0, p1])
    loss_ref, g_ref = oracle.loss_grad(full, TOY_DIMS, x, y)
    g1 = b1["g"].cpu().numpy()
    assert abs(float(b1["ring"][0]) - loss_ref) <= 2e-6 * abs(loss_ref)
    assert np.max(np.abs(g1[:P1] - g_ref[140:])) <= 2e-6 * np.max(np.abs(g_ref))
    # stage 0 backward from that gradient
    j0.d_out, j0.mode = dact.data_ptr(), _lib.MODE_GRAD
    _launch(native_lib, j0)
    g0 = b0["g"].cpu().numpy()
    assert np.max(np.abs(g0[:140] - g_ref[:140])) <= 2e-6 * np.max(np.abs(g_ref))
    g0_ref, _ = oracle.segment_backward(p0, dims0, x, dact.cpu().numpy(), act_last=True, want_din=False)
    assert np.max(np.abs(g0[:140] - g0_ref)) <= 2e-6 * max(np.max(np.abs(g0_ref)), 1e-30)


def test_stage_argument_errors(native_lib):
    j = _lib.SegJob()
    j.nl = 4
    assert native_lib.ddpb_segment_step(ctypes.byref(j), None) != 0
    assert b"layers" in native_lib.ddpb_last_error()


def _stage_devices(same_gpu):
    """(dev0, dev1): two GPUs, or both stages on GPU 0 -- which is what the reference's own launcher
    produces (interactive_job_cmds/salloc_one_model_multi_gpu_torchrun.sh:5 sets TASKS_PER_NODE=1, so
    demo_one_model_multi_gpu.py:90-91 computes dev0 = dev1 = 0) and needs no second GPU."""
    if same_gpu:
        return 0, 0
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    return 0, 1


@pytest.mark.parametrize("same_gpu", [False, True])
def test_two_gpu_pipeline_front_end(oracle, same_gpu):
    import distributed_training_pytorch_b200 as b200
    dev0, dev1 = _stage_devices(same_gpu)
    torch.manual_seed(7)
    model = b200.DistributedDataParallel(b200.MultiGPUModel(dev0, dev1), device_ids=None)
    opt = b200.optim.Adam(model.parameters(), lr=1e-3)
    loss_fn = b200.nn.MSELoss()
    p = model.flat_parameters().cpu().numpy().copy()
    m, v = np.zeros_like(p), np.zeros_like(p)
    rng = np.random.default_rng(0)
    for step in range(1, 31):
        B = (256, 100, 300)[step % 3]
        x = rng.standard_normal((B, 2)).astype(np.float32)
        y = rng.standard_normal((B, 1)).astype(np.float32)
        opt.zero_grad(set_to_none=True)
        target = torch.from_numpy(y).to(model.module.dev1)
        loss = loss_fn(model(torch.from_numpy(x)), target)
        loss.backward()
        opt.step()
        loss_ref, g = oracle.loss_grad(p, TOY_DIMS, x, y)
        oracle.adam_step(p, g, m, v, step)
        assert abs(float(loss.detach().cpu()) - loss_ref) <= 1e-5 * abs(loss_ref)
    assert np.max(np.abs(model.flat_parameters().cpu().numpy() - p)) <= 1e-5
    assert model.module.layers0[0].weight.device.index == dev0 and model.module.layers1[0].weight.device.index == dev1


@pytest.mark.parametrize("same_gpu", [False, True])
def test_model_parallel_demo_entry_point(tmp_path, oracle, same_gpu):
    _stage_devices(same_gpu)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    code = ("import sys, os, numpy as np; sys.path.insert(0, %r); import demo_one_model_multi_gpu as d; "
            "sys.argv = ['demo', '--backend=nccl', '--dry_run']; L = []; "
            "m = d.main(on_loss=lambda it, l: L.append(l)); "
            "np.savez(%r, losses=np.asarray(L, np.float32), p=m.flat_parameters().cpu().numpy())"
            % (ROOT, os.path.join(str(tmp_path), "out.npz")))
    env = dict(os.environ, WORLD_SIZE="1", TASKS_PER_NODE="2", SLURM_LOCALID="0", SLURM_PROCID="0",
               MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="0,1",
               DDPB_ITERATIONS="100", DDPB_SEED="1234", DDPB_NO_WANDB="1")
    if same_gpu:      # the reference launcher's mapping: TASKS_PER_NODE=1 puts both stages on GPU 0
        env.update(TASKS_PER_NODE="1", CUDA_VISIBLE_DEVICES="0", DDPB_SINGLE_GPU_STAGES="1")
    proc = subprocess.run([sys.executable, "-c", code], env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, (proc.stdout + proc.stderr)[-4000:]
    out = np.load(os.path.join(str(tmp_path), "out.npz"))
    # same RNG consumption as the entry point: one model, then the dataset
    import distributed_training_pytorch_b200 as b200
    torch.manual_seed(1234)
    init = torch.cat([q.detach().reshape(-1) for q in b200.ToyModel().parameters()]).numpy()
    data = b200.ToyData()
    p_ref, losses_ref = oracle.train_loop(init[None], TOY_DIMS, data.x_dense.numpy(), data.y_dense.numpy(), 1, 100)
    assert np.max(np.abs(out["losses"] - losses_ref[:, 0]) / losses_ref[:, 0]) <= 1e-5
    assert np.max(np.abs(out["p"] - p_ref[0])) <= 1e-5
    # and the reference's own run of demo_one_model_multi_gpu.py (tests/golden/make_golden.py modelparallel)
    g = np.load(os.path.join(ROOT, "tests", "golden", "ref_model_parallel_w1.npz"))
    assert np.array_equal(init, g["init"]) and np.array_equal(data.x_dense.numpy(), g["data_x"])
    assert np.max(np.abs(out["losses"] - g["loss"][:100]) / g["loss"][:100]) <= 2e-5
