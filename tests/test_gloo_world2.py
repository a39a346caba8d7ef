"""World-size-2 plumbing on the CPU (gloo): the three rendezvous styles of the reference
(torchrun env, srun-style env with tcp:// init, MPI-launcher env), sampler sharding from the
process group, the loss-reduction convention, and the loud failure of the compute path
without a GPU (config 1 of BASELINE.json is plumbing only).  Also pins the 2-rank CPU
baseline (oracle/torch_ref.py) to the reference's own 2-rank run."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "_gloo_worker.py")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(mode, tmp_path, per_rank_env):
    procs = []
    for env_extra in per_rank_env:
        env = dict(os.environ, CUDA_VISIBLE_DEVICES="", **env_extra)
        procs.append(subprocess.Popen([sys.executable, WORKER, mode, str(tmp_path)], env=env, cwd=ROOT,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    return [json.load(open(os.path.join(tmp_path, f"rank{r}.json"))) for r in range(len(per_rank_env))]


def _check(reports, golden_dir):
    gold = np.load(os.path.join(golden_dir, "sampler.npz"))
    for rep in reports:
        assert rep["world"] == 2
        for epoch, shards in enumerate(rep["shards"]):
            assert sorted(shards[0] + shards[1]) == list(range(512))            # disjoint cover
            for r in range(2):
                assert shards[r] == gold[f"w2_r{r}"][epoch].tolist()              # bit-exact vs torch's sampler
        assert rep["mean_loss"] == pytest.approx(1.5)
        assert rep["ddp_error"] and "CUDA" in rep["ddp_error"]


def test_srun_style_env(tmp_path, golden_dir):
    port = str(_free_port())
    base = dict(WORLD_SIZE="2", TASKS_PER_NODE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=port)
    reports = _spawn("slurm", tmp_path, [dict(base, SLURM_LOCALID=str(r), SLURM_PROCID=str(r)) for r in range(2)])
    _check(reports, golden_dir)
    assert [r["local_rank"] for r in reports] == [0, 1]


def test_node_rank_arithmetic_two_nodes_one_task(tmp_path, golden_dir):
    port = str(_free_port())
    base = dict(WORLD_SIZE="2", TASKS_PER_NODE="1", MASTER_ADDR="127.0.0.1", MASTER_PORT=port, USE_NODE_RANK="1",
                SLURM_LOCALID="0")
    reports = _spawn("slurm", tmp_path, [dict(base, NODE_RANK=str(r)) for r in range(2)])
    _check(reports, golden_dir)
    assert [r["local_rank"] for r in reports] == [0, 0]


def test_torchrun_style_env(tmp_path, golden_dir):
    port = str(_free_port())
    base = dict(WORLD_SIZE="2", LOCAL_WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=port)
    reports = _spawn("torchrun", tmp_path, [dict(base, RANK=str(r), LOCAL_RANK=str(r)) for r in range(2)])
    _check(reports, golden_dir)


def test_mpi_launcher_env_without_mpi4py(tmp_path, golden_dir):
    port = str(_free_port())
    base = dict(OMPI_COMM_WORLD_SIZE="2", TASKS_PER_NODE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=port)
    reports = _spawn("mpi", tmp_path, [dict(base, OMPI_COMM_WORLD_RANK=str(r)) for r in range(2)])
    _check(reports, golden_dir)


def test_two_rank_cpu_baseline_is_the_reference(tmp_path, golden_dir):
    """oracle/torch_ref.py at W=2 over gloo == the reference's own W=2 run, bit for bit."""
    port = _free_port()
    code = ("import sys, json, numpy as np; sys.path.insert(0, %r); from oracle import torch_ref; "
            "r = torch_ref.run(2, int(sys.argv[1]), %d, iters=200, warmup=0, record=True); "
            "np.savez(sys.argv[2], losses=np.asarray(r['losses'], np.float32), px=r['params'][0], py=r['params'][1])"
            % (ROOT, port))
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    procs = [subprocess.Popen([sys.executable, "-c", code, str(r), os.path.join(tmp_path, f"r{r}.npz")], env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    for p in procs:
        out = p.communicate(timeout=300)[0]
        assert p.returncode == 0, out[-3000:]
    g = np.load(os.path.join(golden_dir, "ref_w2.npz"))
    r0 = np.load(os.path.join(tmp_path, "r0.npz"))
    assert np.array_equal(r0["losses"][:, 0], g["lossX"][:200]) and np.array_equal(r0["losses"][:, 1], g["lossY"][:200])
    assert np.array_equal(r0["px"], g["mid_X"]) and np.array_equal(r0["py"], g["mid_Y"])
