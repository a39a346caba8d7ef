"""ddpb_toy_step_virtual_ranks: the fused multi-replica step whose gradient-exchange windows
are compiled into the library instead of being allocated by the caller.

It exists for callers that have no way to allocate device memory of their own (the GPU was
given to another memory pool, or the loop may not take the allocator's device-wide
synchronisation).  The check: 2, 3 and 4 replicas of one model, each with its own DistributedSampler
shard, take two steps in a row through the new entry point and through the existing one
(ddpb_toy_step with world = W and caller-allocated windows); parameters, Adam moments,
averaged gradients and losses must be IDENTICAL bit for bit, and both must agree with the CPU
oracle (divide-then-sum average + Adam) within the 1e-6 the other tests use.  Collected last
on purpose (file name), and self-contained so it also passes alone in a fresh process."""
import os

import numpy as np
import pytest
import torch

from distributed_training_pytorch_b200 import _lib
from oracle import TOY_DIMS

pytestmark = pytest.mark.gpu

from gpu_util import P, ToyHarness, dev  # noqa: E402


def _replicas(init, n, windows=None):
    reps = [ToyHarness(init) for _ in range(n)]
    for r, h in enumerate(reps):
        if windows is not None:                      # existing entry point: caller-allocated exchange windows
            j = h.job
            j.world, j.rank, j.inv_world = n, r, float(np.float32(1.0) / np.float32(n))
            for p, window in enumerate(windows):
                j.peer_data[p] = window.data_ptr()
    return reps


def _launch(fn, reps):
    jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
    for i, h in enumerate(reps):
        jobs[i] = h.job
    _lib.check(fn(jobs, len(reps), torch.cuda.current_stream().cuda_stream), "launch")
    for h in reps:
        h.seq += 1


@pytest.mark.parametrize("W", [2, 3, 4])
def test_two_steps_match_the_allocating_entry_point(oracle, native_lib, golden_dir, W):
    g = np.load(os.path.join(golden_dir, "ref_w2.npz"))
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    X, Y = dev(d["x"]), dev(d["y"])
    windows = [torch.zeros(_lib.XCHG_WINDOW_BYTES // 4, dtype=torch.int32, device="cuda") for _ in range(W)]
    old = _replicas(g["init_X"], W, windows)
    new = _replicas(g["init_X"], W)
    p, m, v = g["init_X"].copy(), np.zeros(P, np.float32), np.zeros(P, np.float32)
    B = -(-512 // W)                                                # 256 / 171 / 128 samples per replica
    for step in (1, 2):                                             # two calls in a row share the built-in array
        shards = [oracle.sampler_indices(512, W, r, epoch=step - 1) for r in range(W)]
        idx = [dev(s.astype(np.int32), torch.int32) for s in shards]
        for reps in (old, new):
            for r, h in enumerate(reps):
                h.bind(X, Y, _lib.MODE_STEP, idx=idx[r], idx_rows=1, idx_stride=B, B=B)
        _launch(native_lib.ddpb_toy_step, old)
        _launch(native_lib.ddpb_toy_step_virtual_ranks, new)
        torch.cuda.synchronize()
        for a, b in zip(old, new):
            assert torch.equal(a.params, b.params) and torch.equal(a.m, b.m) and torch.equal(a.v, b.v)
            assert torch.equal(a.grads, b.grads) and torch.equal(a.ring, b.ring)
            assert a.state.cpu().tolist() == b.state.cpu().tolist() == [step, step, 0, 0]
        assert torch.equal(new[0].params, new[1].params)            # replicas stay identical
        # against the oracle: per-rank gradients, DDP average (x 1/W then sum), torch's Adam
        per_rank = []
        for r in range(W):
            loss_r, grad_r = oracle.loss_grad(p, TOY_DIMS, d["x"][shards[r]], d["y"][shards[r]])
            per_rank.append(grad_r)
            assert abs(new[r].loss(step - 1) - loss_r) <= 2e-6 * abs(loss_r)
        avg = oracle.ddp_average(np.stack(per_rank))
        assert np.max(np.abs(new[0].grads.cpu().numpy()[:P] - avg)) <= 2e-6 * np.max(np.abs(avg))
        oracle.adam_step(p, avg, m, v, step)
        assert np.max(np.abs(new[0].params.cpu().numpy() - p)) <= 1e-6


def test_more_replicas_than_the_builtin_windows_hold_is_rejected(native_lib):
    jobs = (_lib.ToyJob * 8)()
    for n in (0, 1, 5):
        assert native_lib.ddpb_toy_step_virtual_ranks(jobs, n, None) != 0
        assert b"2..4" in native_lib.ddpb_last_error()
