"""Fused average+Adam / SGD pass and the gather kernel, through the C ABI, vs the oracle.
Same gradient in -> the update must be BIT-EXACT (explicit rn intrinsics mirror the
oracle's op order); the bias corrections of the device-counter variant come from a host-built
table computed with the same Python-double expressions as torch, so it is bit-exact too."""
import numpy as np
import pytest
import torch

from distributed_training_pytorch_b200 import _lib

pytestmark = pytest.mark.gpu

from gpu_util import bias_table, dev, new_state  # noqa: E402


@pytest.mark.parametrize("n", [1, 3, 371, 372, 1000, 4099, (1 << 20) + 3])
@pytest.mark.parametrize("scale", [1.0, 0.125])
def test_avg_adam_bit_exact(oracle, native_lib, n, scale):
    rng = np.random.default_rng(n)
    p = rng.standard_normal(n).astype(np.float32)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    dp, dm, dv = dev(p), dev(m), dev(v)
    stream = torch.cuda.current_stream().cuda_stream
    for step in range(1, 5):
        g = (rng.standard_normal(n) * 10 ** rng.uniform(-6, 1)).astype(np.float32)
        dg = dev(g)
        _lib.check(native_lib.ddpb_fused_avg_adam(dg.data_ptr(), dp.data_ptr(), dm.data_ptr(), dv.data_ptr(), n,
                                                  scale, 1e-3, 0.9, 0.999, 1e-8, step, stream))
        oracle.adam_step(p, g * np.float32(scale), m, v, step)
        assert np.array_equal(dp.cpu().numpy(), p) and np.array_equal(dm.cpu().numpy(), m)
        assert np.array_equal(dv.cpu().numpy(), v)


def test_avg_adam_unaligned_views(oracle, native_lib):
    n = 1001
    rng = np.random.default_rng(0)
    base = [torch.zeros(n + 3, device="cuda") for _ in range(4)]
    g, p, m, v = (b[1:1 + n] for b in base)                     # 4-byte aligned only
    hp = rng.standard_normal(n).astype(np.float32); hg = rng.standard_normal(n).astype(np.float32)
    p.copy_(torch.from_numpy(hp)); g.copy_(torch.from_numpy(hg))
    _lib.check(native_lib.ddpb_fused_avg_adam(g.data_ptr(), p.data_ptr(), m.data_ptr(), v.data_ptr(), n, 1.0,
                                              1e-3, 0.9, 0.999, 1e-8, 1, torch.cuda.current_stream().cuda_stream))
    hm, hv = np.zeros(n, np.float32), np.zeros(n, np.float32)
    oracle.adam_step(hp, hg, hm, hv, 1)
    assert np.array_equal(p.cpu().numpy(), hp)
    assert all(float(b[0]) == 0 and float(b[-1]) == 0 for b in base[1:])   # no out-of-bounds writes


def test_avg_adam_device_counters(oracle, native_lib):
    n = 5000
    rng = np.random.default_rng(5)
    p = rng.standard_normal(n).astype(np.float32)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    dp, dm, dv, state, table = dev(p), dev(m), dev(v), new_state(), bias_table()
    stream = torch.cuda.current_stream().cuda_stream
    for step in range(1, 30):
        g = rng.standard_normal(n).astype(np.float32)
        dg = dev(g)
        _lib.check(native_lib.ddpb_fused_avg_adam_dev(dg.data_ptr(), dp.data_ptr(), dm.data_ptr(), dv.data_ptr(),
                                                      n, 0.5, 0.9, 0.999, 1e-8, table.data_ptr(), table.shape[0],
                                                      state.data_ptr(), 1, stream))
        oracle.adam_step(p, g * np.float32(0.5), m, v, step)
    assert state.cpu().numpy()[0] == 29
    assert np.array_equal(dp.cpu().numpy(), p) and np.array_equal(dm.cpu().numpy(), m)   # table == torch's doubles


@pytest.mark.parametrize("momentum", [0.0, 0.9])
def test_avg_sgd_bit_exact(oracle, native_lib, momentum):
    n = 3001
    rng = np.random.default_rng(9)
    p, buf = rng.standard_normal(n).astype(np.float32), np.zeros(n, np.float32)
    dp, dbuf = dev(p), dev(buf)
    for step in range(1, 5):
        g = rng.standard_normal(n).astype(np.float32)
        dg = dev(g)
        _lib.check(native_lib.ddpb_fused_avg_sgd(dg.data_ptr(), dp.data_ptr(), dbuf.data_ptr(), n, 0.25, 1e-2,
                                                 momentum, step, torch.cuda.current_stream().cuda_stream))
        oracle.sgd_step(p, g * np.float32(0.25), buf, step, lr=1e-2, momentum=momentum)
        assert np.array_equal(dp.cpu().numpy(), p)


@pytest.mark.parametrize("B", [0, 1, 256, 5000])
def test_gather_batch(native_lib, B):
    rng = np.random.default_rng(B)
    X, Y = rng.standard_normal((512, 2)).astype(np.float32), rng.standard_normal((512, 1)).astype(np.float32)
    idx = rng.integers(0, 512, size=max(B, 1)).astype(np.int32)
    ox, oy = torch.zeros(max(B, 1), 2, device="cuda"), torch.zeros(max(B, 1), 1, device="cuda")
    dX, dY, didx = dev(X), dev(Y), dev(idx, torch.int32)          # keep the device copies alive across the launch
    _lib.check(native_lib.ddpb_gather_batch(dX.data_ptr(), dY.data_ptr(), didx.data_ptr(), B,
                                            2, 1, ox.data_ptr(), oy.data_ptr(), torch.cuda.current_stream().cuda_stream))
    if B:
        assert np.array_equal(ox.cpu().numpy(), X[idx]) and np.array_equal(oy.cpu().numpy(), Y[idx])
