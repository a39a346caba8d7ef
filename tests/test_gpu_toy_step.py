"""Parity of the fused toy-model kernel (ddpb_toy_step, through the C ABI) with the CPU
oracle and with the reference's own run (tests/golden).

Tolerances (fp32 path; the kernel uses the oracle's per-sample arithmetic order, only the
batch-sum order differs):
  forward values                     bit-exact
  mean loss of one batch             rel 1e-6
  gradient of one batch              2e-6 of max|grad| (absolute, per element)
  Adam/SGD update from that gradient 1e-6 abs on parameters (lr=1e-3 bounds the update)
  200-step trajectory vs the reference run: params abs 2e-5, loss rel 2e-5
  1000-step loss curve vs the reference run: rel 1e-2 (chaotic tail, see test_oracle_golden)
"""
import os

import numpy as np
import pytest
import torch

from distributed_training_pytorch_b200 import _lib
from oracle import TOY_DIMS

pytestmark = pytest.mark.gpu

from gpu_util import P, ToyHarness, dev  # noqa: E402


def _rand_params(rng, scale=0.5):
    return (rng.standard_normal(P) * scale).astype(np.float32)


@pytest.mark.parametrize("B", [1, 5, 31, 32, 33, 64, 100, 128, 255, 256, 257, 512, 1000, 4096])
def test_forward_bit_exact(oracle, native_lib, B):
    rng = np.random.default_rng(B)
    params = _rand_params(rng)
    x = rng.standard_normal((B, 2)).astype(np.float32)
    h = ToyHarness(params)
    out = torch.full((B, 1), float("nan"), device="cuda")
    h.bind(dev(x), None, _lib.MODE_FWD, out=out)
    h.launch()
    ref = oracle.forward(params, TOY_DIMS, x)
    assert np.array_equal(out.cpu().numpy(), ref)


@pytest.mark.parametrize("B", [1, 5, 31, 32, 33, 64, 100, 128, 255, 256, 257, 512, 1000, 4096])
def test_loss_and_gradient(oracle, native_lib, B):
    rng = np.random.default_rng(1000 + B)
    params = _rand_params(rng)
    x = rng.standard_normal((B, 2)).astype(np.float32)
    y = rng.standard_normal((B, 1)).astype(np.float32)
    h = ToyHarness(params)
    h.bind(dev(x), dev(y), _lib.MODE_GRAD)
    (slot,) = h.launch()
    loss_ref, grad_ref = oracle.loss_grad(params, TOY_DIMS, x, y)
    got = h.grads.cpu().numpy()
    assert abs(h.loss(slot) - loss_ref) <= 1e-6 * abs(loss_ref)
    assert abs(got[P] / B - loss_ref) <= 1e-6 * abs(loss_ref)          # bucket tail = sum (o-t)^2
    assert np.max(np.abs(got[:P] - grad_ref)) <= 2e-6 * np.max(np.abs(grad_ref))
    assert np.array_equal(h.params.cpu().numpy(), params)              # GRAD must not touch parameters
    assert h.state.cpu().numpy()[:2].tolist() == [0, 1]                # step unchanged, seq advanced


def test_gathered_batch_equals_direct_batch(oracle, native_lib, golden_dir):
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    rows = np.stack([oracle.sampler_indices(512, 2, 1, epoch=e) for e in range(3)]).astype(np.int32)
    X, Y, idx = dev(d["x"]), dev(d["y"]), dev(rows, torch.int32)
    a, b, c = ToyHarness(g["init_X"]), ToyHarness(g["init_X"]), ToyHarness(g["init_X"])
    for e in range(3):
        a.bind(X, Y, _lib.MODE_GRAD, idx=idx, idx_rows=3, idx_stride=256, B=256)     # row picked by seq % 3; dataset staged on chip
        a.launch()
        b.bind(dev(d["x"][rows[e]]), dev(d["y"][rows[e]]), _lib.MODE_GRAD)
        b.launch()
        c.bind(X, Y, _lib.MODE_GRAD, idx=idx, idx_rows=3, idx_stride=256, B=256, staged=False)   # dependent-load gather
        c.launch()
        assert torch.equal(a.grads, b.grads) and torch.equal(c.grads, b.grads), e


@pytest.mark.parametrize("opt,momentum", [(_lib.OPT_ADAM, 0.0), (_lib.OPT_SGD, 0.0), (_lib.OPT_SGD, 0.9)])
def test_fused_step_matches_oracle(oracle, native_lib, opt, momentum):
    rng = np.random.default_rng(7)
    params = _rand_params(rng)
    h = ToyHarness(params, opt=opt, lr=1e-3 if opt == _lib.OPT_ADAM else 1e-2, momentum=momentum)
    p, m, v = params.copy(), np.zeros(P, np.float32), np.zeros(P, np.float32)
    for step in range(1, 6):
        x = rng.standard_normal((256, 2)).astype(np.float32)
        y = rng.standard_normal((256, 1)).astype(np.float32)
        h.bind(dev(x), dev(y), _lib.MODE_STEP)
        (slot,) = h.launch()
        loss_ref, grad = oracle.loss_grad(p, TOY_DIMS, x, y)
        if opt == _lib.OPT_ADAM:
            oracle.adam_step(p, grad, m, v, step)
        else:
            oracle.sgd_step(p, grad, m, step, lr=1e-2, momentum=momentum)
        assert abs(h.loss(slot) - loss_ref) <= 2e-6 * abs(loss_ref)
        assert np.max(np.abs(h.params.cpu().numpy() - p)) <= 1e-6, step
        assert h.state.cpu().numpy()[:2].tolist() == [step, step]


def test_two_models_one_launch_equal_two_launches(native_lib, golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    rng = np.random.default_rng(3)
    x, y = dev(rng.standard_normal((256, 2))), dev(rng.standard_normal((256, 1)))
    ax, ay = ToyHarness(g["init_X"]), ToyHarness(g["init_Y"])
    bx, by = ToyHarness(g["init_X"]), ToyHarness(g["init_Y"])
    for _ in range(3):
        for h in (ax, ay, bx, by):
            h.bind(x, y, _lib.MODE_STEP)
        ax.launch(others=(ay,))
        bx.launch(); by.launch()
    assert torch.equal(ax.params, bx.params) and torch.equal(ay.params, by.params)     # deterministic reduction
    assert not torch.equal(ax.params, ay.params)


def _replay_reference_run(golden_dir, oracle, iters):
    """demo.py's loop at W=1 (two batches per epoch, demo.py:95-129) through the C ABI."""
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    X, Y = dev(d["x"]), dev(d["y"])
    hx, hy = ToyHarness(g["init_X"], ring=1024), ToyHarness(g["init_Y"], ring=1024)
    epochs = (iters + 1) // 2
    rows = np.stack([oracle.sampler_indices(512, 1, 0, epoch=e) for e in range(epochs)]).astype(np.int32)
    idx = dev(rows.reshape(-1, 256), torch.int32)               # one row per iteration
    for it in range(iters):
        for h in (hx, hy):
            h.bind(X, Y, _lib.MODE_STEP, idx=idx, idx_rows=idx.shape[0], idx_stride=256, B=256)
        hx.launch(others=(hy,))
    torch.cuda.synchronize()
    return g, hx, hy


def test_200_steps_against_reference_run(oracle, native_lib, golden_dir):
    g, hx, hy = _replay_reference_run(golden_dir, oracle, 200)
    assert int(g["mid_step"]) == 200
    assert np.max(np.abs(hx.params.cpu().numpy() - g["mid_X"])) <= 2e-5
    assert np.max(np.abs(hy.params.cpu().numpy() - g["mid_Y"])) <= 2e-5
    lx, ly = hx.ring.cpu().numpy()[:200], hy.ring.cpu().numpy()[:200]
    assert np.max(np.abs(lx - g["lossX"][:200]) / g["lossX"][:200]) <= 2e-5
    assert np.max(np.abs(ly - g["lossY"][:200]) / g["lossY"][:200]) <= 2e-5


def test_1000_steps_against_reference_run(oracle, native_lib, golden_dir):
    g, hx, hy = _replay_reference_run(golden_dir, oracle, 1000)
    lx, ly = hx.ring.cpu().numpy()[:1000], hy.ring.cpu().numpy()[:1000]
    assert np.max(np.abs(lx[:500] - g["lossX"][:500]) / g["lossX"][:500]) <= 2e-5
    assert np.max(np.abs(lx - g["lossX"]) / g["lossX"]) <= 1e-2
    assert np.max(np.abs(ly - g["lossY"]) / g["lossY"]) <= 1e-2
    assert abs(lx[-1] - 0.22993478178977966) < 1e-3            # BASELINE.md anchors
    assert abs(ly[-1] - 0.23505525290966034) < 1e-3


def test_argument_errors(native_lib):
    jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
    assert native_lib.ddpb_toy_step(jobs, 0, None) != 0
    assert native_lib.ddpb_toy_step(jobs, 5, None) != 0
    assert native_lib.ddpb_toy_step(jobs, 1, None) != 0          # null params
    assert b"params" in native_lib.ddpb_last_error()


def _launch_on(native_lib, stream, harnesses):
    jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
    for i, h in enumerate(harnesses):
        jobs[i] = h.job
    _lib.check(native_lib.ddpb_toy_step(jobs, len(harnesses), stream), "ddpb_toy_step")


def test_graph_replay_equals_eager_launches(oracle, native_lib, golden_dir):
    """K steps captured ONCE as a CUDA graph (ddpb_graph_*) and replayed twice must leave exactly what
    2K eager launches leave: parameters, both Adam moments, the gradient bucket, the loss ring and the
    device-resident counters (which is what lets a replay pick the next sampler row and the next
    bias-correction entry without new arguments) -- bit for bit, for both models of a launch."""
    import ctypes
    K = 12
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    X, Y = dev(d["x"]), dev(d["y"])
    rows = np.stack([oracle.sampler_indices(512, 2, 0, epoch=e) for e in range(16)]).astype(np.int32)
    idx = dev(rows, torch.int32)

    def pair():
        hs = (ToyHarness(g["init_X"], ring=32), ToyHarness(g["init_Y"], ring=32))
        for h in hs:
            h.bind(X, Y, _lib.MODE_STEP, idx=idx, idx_rows=16, idx_stride=256, B=256)      # row = seq % 16, chosen on the device
        return hs
    eager, replay = pair(), pair()
    side = torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(2 * K):
        _launch_on(native_lib, side.cuda_stream, eager)
    graph = ctypes.c_void_p()
    _lib.check(native_lib.ddpb_graph_begin(side.cuda_stream), "ddpb_graph_begin")
    for _ in range(K):
        _launch_on(native_lib, side.cuda_stream, replay)
    _lib.check(native_lib.ddpb_graph_end(side.cuda_stream, ctypes.byref(graph)), "ddpb_graph_end")
    side.synchronize()
    assert replay[0].state.cpu().tolist() == [0, 0, 0, 0]                # capture launched nothing
    for _ in range(2):
        _lib.check(native_lib.ddpb_graph_launch(graph, side.cuda_stream), "ddpb_graph_launch")
    side.synchronize()
    _lib.check(native_lib.ddpb_graph_destroy(graph), "ddpb_graph_destroy")
    for a, b in zip(eager, replay):
        assert a.state.cpu().tolist() == b.state.cpu().tolist() == [2 * K, 2 * K, 0, 0]
        for name in ("params", "m", "v", "grads", "ring"):
            assert torch.equal(getattr(a, name), getattr(b, name)), name
    assert not torch.equal(eager[0].params, dev(g["init_X"]))


def test_missing_peer_times_out_loudly_and_leaves_state_untouched(native_lib, golden_dir):
    """world = 2 but the other rank never launches: after ~2 s the kernel gives up -- it must say so
    (state.reserved[0] = 1, the loss slot holds DDPB_LOSS_TIMEOUT_BITS) and must NOT have applied a step
    computed from a window that never filled (parameters, moments and both counters unchanged)."""
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    rng = np.random.default_rng(11)
    x, y = dev(rng.standard_normal((256, 2))), dev(rng.standard_normal((256, 1)))
    windows = [torch.zeros(_lib.XCHG_WINDOW_BYTES // 4, dtype=torch.int32, device="cuda") for _ in range(2)]
    h = ToyHarness(g["init_X"], ring=4)
    h.bind(x, y, _lib.MODE_STEP)
    j = h.job
    j.world, j.rank, j.inv_world = 2, 0, 0.5
    for p, w in enumerate(windows):
        j.peer_data[p] = w.data_ptr()
    h.launch()
    torch.cuda.synchronize()
    assert h.state.cpu().tolist() == [0, 0, 1, 0]
    assert h.ring.cpu().numpy().view(np.uint32)[0] == _lib.LOSS_TIMEOUT_BITS
    assert np.array_equal(h.params.cpu().numpy(), g["init_X"])
    assert not h.m.any().item() and not h.v.any().item()


def test_one_call_iteration_gathers_launches_and_waits(oracle, native_lib, golden_dir):
    """ddpb_toy_iteration: host gather into pinned staging + launch (the kernel reads the pinned rows in
    place over PCIe) + wait for the pinned loss slots, in one call; same result as the separate calls
    on a device-resident copy of the same rows."""
    import ctypes
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    B = 200
    rows = oracle.sampler_indices(512, 2, 1, epoch=5)[:B].astype(np.int64)
    xs, ys = np.ascontiguousarray(d["x"]), np.ascontiguousarray(d["y"])
    stage = torch.zeros(256 * 3, dtype=torch.float32).pin_memory()
    ring = torch.zeros(8, dtype=torch.float32).pin_memory()
    ring.numpy().view(np.uint32)[:] = _lib.LOSS_SENTINEL
    a, b = ToyHarness(g["init_X"]), ToyHarness(g["init_X"])
    b.bind(dev(d["x"][rows]), dev(d["y"][rows]), _lib.MODE_STEP)
    b.launch()
    j = a.job
    j.x, j.y, j.idx, j.B, j.mode = stage.data_ptr(), stage.data_ptr() + 256 * 2 * 4, None, B, _lib.MODE_STEP
    j.loss_ring, j.loss_ring_n = ring.data_ptr(), 8
    jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
    jobs[0] = j
    it = _lib.ToyIter()
    it.jobs, it.njobs, it.stream = ctypes.addressof(jobs), 1, torch.cuda.current_stream().cuda_stream
    it.X, it.Y, it.idx, it.B, it.dx, it.dy = xs.ctypes.data, ys.ctypes.data, rows.ctypes.data, B, 2, 1
    it.out_x, it.out_y = j.x, j.y
    it.nwait, it.sentinel_bits, it.spin_ns = 1, _lib.LOSS_SENTINEL, 1_000_000
    it.wait_slot[0] = ring.data_ptr()
    _lib.check(native_lib.ddpb_toy_iteration(ctypes.addressof(it)), "ddpb_toy_iteration")
    assert np.array_equal(stage.numpy()[:2 * B].reshape(B, 2), d["x"][rows])
    assert it.wait_value[0] == float(ring[0]) == b.loss(0)          # already there when the call returns
    torch.cuda.synchronize()
    assert torch.equal(a.params, b.params) and torch.equal(a.grads, b.grads)
