#!/usr/bin/env python3
"""Drop-in for the reference's demo_assume_started_with_mpiexec.py: the same two-model
data-parallel loop as demo.py, but rank/size come from the MPI launcher and rank 0's address
plus a free port are broadcast before the env:// rendezvous (reference :29-50).

    TASKS_PER_NODE=8 mpiexec -n 8 python demo_assume_started_with_mpiexec.py --backend=nccl

mpi4py is used when importable; without it (this image has neither mpi4py nor mpiexec) the
launcher's environment (OMPI_COMM_WORLD_RANK/SIZE or PMI_RANK/SIZE, MASTER_ADDR/PORT) is read
instead, which is what tools/mpiexec_compat.py exports:

    TASKS_PER_NODE=8 python tools/mpiexec_compat.py -n 8 python demo_assume_started_with_mpiexec.py --backend=nccl

Like the reference, this variant has no wandb, and its loss all-reduce operates on a temporary
(reference :113,115), i.e. the collective runs but nothing is logged.
"""
import os

import torch
import torch.distributed as dist
from torch.distributed.elastic.multiprocessing.errors import record
from tqdm import tqdm

from demo import get_dataloader
from distributed_training_pytorch_b200 import DistributedDataParallel as DDP
from distributed_training_pytorch_b200 import ToyData, ToyModel, bootstrap, nn, optim
from distributed_training_pytorch_b200.argument_parser import get_args


def setup_distributed(config):
    model_X, model_Y = ToyModel(), ToyModel()
    layout = bootstrap.init_mpi(config.backend)
    bootstrap.announce(layout, config)
    torch.cuda.set_device(layout.device)
    model_X = DDP(model_X.cuda(), device_ids=[layout.device])
    model_Y = DDP(model_Y.cuda(), device_ids=[layout.device])
    return model_X, model_Y


def training_demo(model_X, model_Y, dataloader, sampler, on_loss=None):
    rank = dist.get_rank()
    optimizer_X = optim.Adam(model_X.parameters(), lr=1e-3)
    optimizer_Y = optim.Adam(model_Y.parameters(), lr=1e-3)
    all_reduce_group = dist.new_group(backend="gloo")
    loss = nn.MSELoss()
    max_iterations = int(os.environ.get("DDPB_ITERATIONS", 1000))
    iteration, epoch, training = 0, 0, True
    iteration_pbar = tqdm(range(max_iterations), desc="Iteration", disable=rank != 0)
    while training:
        if sampler is not None:
            sampler.set_epoch(epoch)
        for data, target in dataloader:
            optimizer_X.zero_grad(set_to_none=True)
            optimizer_Y.zero_grad(set_to_none=True)
            data, target = data.cuda(), target.cuda()
            l_X = loss(model_X(data), target)
            l_X.backward()
            l_Y = loss(model_Y(data), target)
            l_Y.backward()
            optimizer_X.step()
            optimizer_Y.step()
            batch_size = target.size(0)
            local_X, local_Y = l_X.detach().cpu(), l_Y.detach().cpu()
            # the reference reduces temporaries here: the collectives synchronise, the sums are dropped
            dist.all_reduce(local_X * batch_size, group=all_reduce_group, op=dist.ReduceOp.SUM)
            dist.all_reduce(local_Y * batch_size, group=all_reduce_group, op=dist.ReduceOp.SUM)
            if on_loss is not None:
                on_loss(iteration, float(local_X), float(local_Y))
            iteration_pbar.update(1)
            iteration += 1
            if iteration == max_iterations:
                training = False
                break
        epoch += 1
    dist.destroy_process_group(all_reduce_group)
    if rank == 0:
        iteration_pbar.close()
    print(f"[Process {rank}] Finished")


@record
def main(on_loss=None):
    config = get_args()
    if config.dry_run:
        os.environ["WANDB_MODE"] = "dryrun"
    if "DDPB_SEED" in os.environ:
        torch.manual_seed(int(os.environ["DDPB_SEED"]))
    model_X, model_Y = setup_distributed(config)
    dataloader, sampler = get_dataloader(config, ToyData())
    training_demo(model_X, model_Y, dataloader, sampler, on_loss=on_loss)
    dist.barrier()
    dist.destroy_process_group()
    return model_X, model_Y


if __name__ == "__main__":
    main()
