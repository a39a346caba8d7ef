"""oracle/torch_ref.py (the CPU baseline that bench.py times) must BE the reference's
arithmetic: same seed -> bit-identical loss curve and parameters as the real reference's
run stored in tests/golden (single rank; the multi-rank runs are covered by the gloo test)."""
import os
import socket

import numpy as np

from oracle import torch_ref


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_torch_ref_is_bit_identical_to_the_reference_run(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    res = torch_ref.run(1, 0, _free_port(), iters=200, warmup=0, record=True)
    losses = np.asarray(res["losses"], np.float32)
    assert np.array_equal(losses[:, 0], g["lossX"][:200])
    assert np.array_equal(losses[:, 1], g["lossY"][:200])
    assert np.array_equal(res["params"][0], g["mid_X"]) and np.array_equal(res["params"][1], g["mid_Y"])


def test_torch_ref_standard_dataloader_single_rank_equals_the_two_rank_reference_run(golden_dir):
    """--dataloader=standard: every rank sees the same batches, so averaging over ranks changes
    nothing but rounding (x 1/2 then + is exact): a single-rank torch_ref run in standard mode
    reproduces the reference's W=2 standard run bit for bit."""
    g = np.load(os.path.join(golden_dir, "ref_w2_standard.npz"))
    res = torch_ref.run(1, 0, _free_port(), iters=200, warmup=0, mode="standard", record=True)
    losses = np.asarray(res["losses"], np.float32)
    assert np.array_equal(losses[:, 0], g["lossX"][:200]) and np.array_equal(losses[:, 1], g["lossY"][:200])
    assert np.array_equal(res["params"][0], g["mid_X"]) and np.array_equal(res["params"][1], g["mid_Y"])
