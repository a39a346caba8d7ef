#!/usr/bin/env python3
"""Drop-in for the reference's demo.py: two toy models trained data-parallel, same flags,
same torchrun / srun-env rendezvous, same loop (reference demo.py:19-181) -- but every
model/loss/optimizer object comes from distributed_training_pytorch_b200, so the step runs
as sm_100a kernels.

    torchrun --standalone --nproc_per_node=8 demo.py --backend=nccl --torchrun [--dry_run]

Additive knobs (environment, so argument_parser stays verbatim): DDPB_ITERATIONS (default
1000, demo.py:88), DDPB_SEED (seed torch before building models/data; the reference never
applies --seed), DDPB_DEVICE_DATA=1 (HBM-resident dataset), DDPB_NO_WANDB=1, DDPB_FUSED_LOSS=1 (the logged
mean loss comes out of the gradient exchange instead of two gloo all-reduces per step, logged one
iteration late), DDPB_SHARE_GPUS=1 (more ranks than GPUs; gloo backend).
"""
import os

import torch
import torch.distributed as dist
from torch.distributed.elastic.multiprocessing.errors import record
from tqdm import tqdm

from distributed_training_pytorch_b200 import DataLoader, DistributedSampler, ToyData, ToyModel, bootstrap, nn, optim
from distributed_training_pytorch_b200 import DistributedDataParallel as DDP
from distributed_training_pytorch_b200.argument_parser import get_args


def _wandb():
    if os.environ.get("DDPB_NO_WANDB") == "1":
        return None
    import wandb
    return wandb


def setup_distributed(config):
    # models are built first, on the CPU, from the process RNG (reference demo.py:22-23)
    model_X, model_Y = ToyModel(), ToyModel()
    if config.torchrun:
        layout = bootstrap.init_torchrun(config.backend)
    else:
        layout = bootstrap.init_slurm_env(config.backend, config.use_node_rank)
    bootstrap.announce(layout, config)
    torch.cuda.set_device(layout.device)              # process -> GPU binding (demo.py:66)
    model_X = DDP(model_X.cuda(), device_ids=[layout.device])
    model_Y = DDP(model_Y.cuda(), device_ids=[layout.device])
    return model_X, model_Y, layout.world_size


def training_demo(model_X, model_Y, dataloader, sampler, world_size, on_loss=None):
    rank = dist.get_rank()
    wandb = _wandb() if rank == 0 else None
    if wandb is not None:
        wandb.init(project="distributed tester", group="base-demo")
    optimizer_X = optim.Adam(model_X.parameters(), lr=1e-3)
    optimizer_Y = optim.Adam(model_Y.parameters(), lr=1e-3)
    all_reduce_group = dist.new_group(backend="gloo")     # CPU side group for the logged losses (demo.py:84)
    loss = nn.MSELoss()
    max_iterations = int(os.environ.get("DDPB_ITERATIONS", 1000))
    iteration, epoch, training = 0, 0, True
    iteration_pbar = tqdm(range(max_iterations), desc="Iteration", disable=rank != 0)
    # DDPB_FUSED_LOSS=1 (additive knob): log from the loss that rides in the gradient bucket (see below)
    fused_loss, behind = os.environ.get("DDPB_FUSED_LOSS") == "1", None

    def log(it, mean_X, mean_Y):
        if rank != 0:
            return
        if wandb is not None:
            wandb.log({"loss/lossX": mean_X}, commit=False)
            wandb.log({"loss/lossY": mean_Y})
        if on_loss is not None:
            on_loss(it, float(mean_X), float(mean_Y))

    while training:
        if sampler is not None:
            sampler.set_epoch(epoch)
        for data, target in dataloader:
            optimizer_X.zero_grad(set_to_none=True)
            optimizer_Y.zero_grad(set_to_none=True)
            data = data.cuda()
            target = target.cuda()
            output_X = model_X(data)
            output_Y = model_Y(data)
            l_X = loss(output_X, target)
            l_X.backward()
            l_Y = loss(output_Y, target)
            l_Y.backward()
            optimizer_X.step()
            optimizer_Y.step()
            batch_size = target.size(0)
            if fused_loss:
                # The same logged value, sum_r(loss_r * B) / (B * W), without the two blocking host
                # collectives: the squared-error sum is element 371 of the gradient bucket, so the gradient
                # exchange has already reduced it.  It is read ONE ITERATION LATE (when it has long
                # arrived in pinned memory), so the host never waits for the GPU here.
                if behind is not None:
                    log(behind[0], behind[1].detach().global_mean(), behind[2].detach().global_mean())
                behind = (iteration, l_X, l_Y)
            else:
                # logged value: sum_r(loss_r * B) / (B * W); assumes equal batch on every rank
                all_reduce_loss_X = l_X.detach().cpu() * batch_size
                dist.all_reduce(all_reduce_loss_X, group=all_reduce_group, op=dist.ReduceOp.SUM)
                all_reduce_loss_Y = l_Y.detach().cpu() * batch_size
                dist.all_reduce(all_reduce_loss_Y, group=all_reduce_group, op=dist.ReduceOp.SUM)
                if rank == 0:
                    log(iteration, all_reduce_loss_X / (batch_size * world_size), all_reduce_loss_Y / (batch_size * world_size))
            iteration_pbar.update(1)
            iteration += 1
            if iteration == max_iterations:
                training = False
                break
        epoch += 1
    if behind is not None:
        log(behind[0], behind[1].detach().global_mean(), behind[2].detach().global_mean())
    dist.destroy_process_group(all_reduce_group)
    if rank == 0:
        iteration_pbar.close()
        if wandb is not None:
            wandb.finish()
    print(f"[Process {rank}] Finished")


def get_dataloader(config, dataset):
    if config.dataloader == "distributed":
        # each process works on its own partition of the (per-epoch shuffled) dataset
        sampler = DistributedSampler(dataset, shuffle=True)
        loader = DataLoader(dataset, batch_size=256, shuffle=False, pin_memory=True,
                            num_workers=config.num_workers, sampler=sampler)
        return loader, sampler
    # 'standard': every process walks the whole dataset in order, i.e. identical minibatches
    return DataLoader(dataset, pin_memory=True, batch_size=256, shuffle=False, num_workers=config.num_workers), None


@record
def main(on_loss=None):
    config = get_args()
    if config.dry_run:
        os.environ["WANDB_MODE"] = "dryrun"
    if "DDPB_SEED" in os.environ:
        torch.manual_seed(int(os.environ["DDPB_SEED"]))
    model_X, model_Y, world_size = setup_distributed(config)
    dataloader, sampler = get_dataloader(config, ToyData())
    training_demo(model_X, model_Y, dataloader, sampler, world_size, on_loss=on_loss)
    dist.barrier()
    dist.destroy_process_group()
    return model_X, model_Y


if __name__ == "__main__":
    main()
