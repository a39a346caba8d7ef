"""Process-group bootstrap for the three launch styles of the reference.

  torchrun   demo.py:25-34        env:// rendezvous, WORLD_SIZE / LOCAL_WORLD_SIZE / LOCAL_RANK
  srun/env   demo.py:35-49        TASKS_PER_NODE, SLURM_LOCALID, NODE_RANK or SLURM_PROCID,
                                  WORLD_SIZE, MASTER_ADDR, MASTER_PORT -> tcp:// rendezvous
  mpiexec    demo_assume_started_with_mpiexec.py:29-50
                                  rank/size from MPI.COMM_WORLD, master address and a free
                                  port broadcast from rank 0, then env:// rendezvous

Error behaviour follows the reference: a missing WORLD_SIZE raises RuntimeError with its
message (demo.py:31-33), missing MASTER_ADDR/PORT raises ValueError (demo.py:47-48).
The NCCL communicator the kernels use is created later, by DistributedDataParallel, from an
id carried over this process group (engine.py).
"""
import datetime
import os
import socket
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class Layout:
    rank: int
    world_size: int
    local_rank: int
    local_world_size: int

    @property
    def device(self):
        """CUDA device of this process: its local rank (demo.py:66).  DDPB_SHARE_GPUS=1 (additive knob)
        wraps the local ranks around the visible GPUs, so that W processes can run on fewer than W GPUs
        (gloo backend; used to exercise the W-rank gradient exchange on a one-GPU box)."""
        if os.environ.get("DDPB_SHARE_GPUS") == "1":
            return self.local_rank % max(1, torch.cuda.device_count())
        return self.local_rank


def _required_int(name):
    value = os.getenv(name, None)
    try:
        return int(value)
    except Exception:
        raise RuntimeError("WORLD_SIZE environment variable is required and " + "must be an interger.") from None


def init_torchrun(backend, timeout=datetime.timedelta(hours=1)):
    """demo.py:25-34."""
    if timeout is None:                        # demo_one_model_multi_gpu.py:49 passes no timeout
        dist.init_process_group(backend=backend)
    else:
        dist.init_process_group(backend=backend, timeout=timeout)
    world_size = _required_int("WORLD_SIZE")
    local_world_size = _required_int("LOCAL_WORLD_SIZE")
    local_rank = int(os.environ["LOCAL_RANK"])
    return Layout(dist.get_rank(), world_size, local_rank, local_world_size)


def layout_from_slurm_env(use_node_rank):
    """Rank arithmetic of demo.py:36-43 (no side effects; unit-testable)."""
    local_world_size = int(os.environ.get("TASKS_PER_NODE"))
    local_rank = int(os.environ.get("SLURM_LOCALID"))
    if use_node_rank:
        rank = int(os.environ.get("NODE_RANK")) * local_world_size + local_rank
    else:
        rank = int(os.environ.get("SLURM_PROCID"))
    world_size = int(os.getenv("WORLD_SIZE", None))
    return Layout(rank, world_size, local_rank, local_world_size)


def init_slurm_env(backend, use_node_rank):
    """demo.py:35-49."""
    layout = layout_from_slurm_env(use_node_rank)
    addr, port = os.getenv("MASTER_ADDR", None), os.getenv("MASTER_PORT", None)
    if addr is None or port is None:
        raise ValueError("MASTER_ADDR and MASTER_PORT must be propvided with the env:// init_method")
    dist.init_process_group(backend=backend, init_method=f"tcp://{addr}:{port}", rank=layout.rank,
                            world_size=layout.world_size)
    return layout


def find_free_port():
    """demo_assume_started_with_mpiexec.py:20-27."""
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("", 0))
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        return s.getsockname()[1]


def mpi_world():
    """(rank, size, bcast) from mpi4py when importable, else from the environment an MPI
    launcher (or tools/mpiexec_compat.py on a box without MPI) exports."""
    try:
        from mpi4py import MPI
        comm = MPI.COMM_WORLD
        return comm.rank, comm.size, lambda obj, root=0: comm.bcast(obj, root=root)
    except ImportError:
        for r, s in (("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE"), ("PMI_RANK", "PMI_SIZE"),
                     ("PMIX_RANK", "OMPI_COMM_WORLD_SIZE")):
            if r in os.environ and s in os.environ:
                return int(os.environ[r]), int(os.environ[s]), None
        raise RuntimeError("neither mpi4py nor an MPI launcher environment (OMPI_COMM_WORLD_RANK/SIZE, "
                           "PMI_RANK/SIZE) is available") from None


def init_mpi(backend):
    """demo_assume_started_with_mpiexec.py:35-50."""
    rank, world_size, bcast = mpi_world()
    tasks_per_node = int(os.getenv("TASKS_PER_NODE", None))
    local_rank = rank % tasks_per_node
    if bcast is not None:
        os.environ["MASTER_ADDR"] = bcast(socket.gethostbyname(socket.getfqdn()), root=0)
        os.environ["MASTER_PORT"] = str(bcast(find_free_port(), root=0))
    elif "MASTER_ADDR" not in os.environ or "MASTER_PORT" not in os.environ:
        raise ValueError("without mpi4py the launcher must export MASTER_ADDR and MASTER_PORT")
    os.environ["RANK"], os.environ["WORLD_SIZE"] = str(rank), str(world_size)
    dist.init_process_group(backend=backend)
    return Layout(rank, world_size, local_rank, tasks_per_node)


def announce(layout, config):
    """The rank-tagged prints of demo.py:51-63 (also bumps config.seed by the rank, :59-60)."""
    rank = dist.get_rank()
    if rank == 0:
        print(f"[Process {rank}] World_size: {dist.get_world_size()}")
    print(f"[Process {rank}] Hello from {socket.gethostname()}")
    if layout.local_rank == 0:
        print(f"[Process {rank}] Available devices on machine: {torch.cuda.device_count()}")
    base_seed = config.seed
    config.seed += rank
    print(f"[Process {rank}] Base seed: {base_seed} and  worker seed: {config.seed}")
    print(f"[Process {rank}] Using torchrun: {config.torchrun}")
    print(f"[Process {rank}] Using backend: {config.backend}")
