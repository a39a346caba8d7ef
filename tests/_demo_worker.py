"""Runs the drop-in demo.py entry point under torchrun and dumps what rank 0 logged plus the
final parameters of every rank (used by tests/test_gpu_multi.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import demo  # noqa: E402
import demo_assume_started_with_mpiexec as demo_mpi  # noqa: E402


def main():
    out_dir = sys.argv[1]
    sys.argv = [sys.argv[0]] + sys.argv[2:]
    losses = []
    entry = demo_mpi if os.environ.get("DDPB_TEST_ENTRY") == "mpiexec" else demo
    model_X, model_Y = entry.main(on_loss=lambda it, lx, ly: losses.append((lx, ly)))
    rank = int(os.environ["RANK"])
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), losses=np.asarray(losses, np.float32),
             px=model_X.flat_parameters().cpu().numpy(), py=model_Y.flat_parameters().cpu().numpy())


if __name__ == "__main__":
    main()
