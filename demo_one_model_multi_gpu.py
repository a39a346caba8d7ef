#!/usr/bin/env python3
"""Drop-in for the reference's demo_one_model_multi_gpu.py: ONE toy model split over two GPUs
of the process (stage 0: 2->10->10 on dev0, stage 1: 10->10->10->1 on dev1), wrapped with
DistributedDataParallel(device_ids=None), one Adam, loss logged by rank 0 (reference :17-181).

A launch that really uses two GPUs (the reference's own launcher sets TASKS_PER_NODE=1, which
maps both stages to GPU 0 -- SURVEY.md section 3.3):

    WORLD_SIZE=1 TASKS_PER_NODE=2 SLURM_LOCALID=0 SLURM_PROCID=0 MASTER_ADDR=127.0.0.1 \\
    MASTER_PORT=29500 CUDA_VISIBLE_DEVICES=0,1 python demo_one_model_multi_gpu.py --backend=nccl --dry_run
"""
import os

import torch
import torch.distributed as dist
from tqdm import tqdm

from demo import _wandb, get_dataloader
from distributed_training_pytorch_b200 import DistributedDataParallel as DDP
from distributed_training_pytorch_b200 import MultiGPUModel, ToyData, bootstrap, nn, optim
from distributed_training_pytorch_b200.argument_parser import get_args


def setup_distributed(config):
    if config.torchrun:
        layout = bootstrap.init_torchrun(config.backend, timeout=None)
    else:
        layout = bootstrap.init_slurm_env(config.backend, config.use_node_rank)
    bootstrap.announce(layout, config)
    # one task drives two GPUs (reference :86-91)
    # DDPB_SINGLE_GPU_STAGES=1 (additive knob) drops the reference's two-device assertion for the case its
    # own launcher produces (TASKS_PER_NODE=1 maps both stages to GPU 0) on a box that has one GPU
    assert torch.cuda.device_count() == 2 or os.environ.get("DDPB_SINGLE_GPU_STAGES") == "1"
    dev0 = (dist.get_rank() * 2) % layout.local_world_size
    dev1 = (dist.get_rank() * 2 + 1) % layout.local_world_size
    model = MultiGPUModel(dev0, dev1)
    return DDP(model, device_ids=None)      # device_ids must be None for a module that spans GPUs


def training_demo(model, dataloader, sampler, on_loss=None):
    rank = dist.get_rank()
    wandb = _wandb() if rank == 0 else None
    if wandb is not None:
        wandb.init(project="distributed tester", group="multi-gpu-per-node")
    optimizer = optim.Adam(model.parameters(), lr=1e-3)
    loss_fn = nn.MSELoss()
    max_iterations = int(os.environ.get("DDPB_ITERATIONS", 1000))
    iteration, epoch, training = 0, 0, True
    iteration_pbar = tqdm(range(max_iterations), desc="Iteration", disable=rank != 0)
    while training:
        if sampler is not None:
            sampler.set_epoch(epoch)
        for data, target in dataloader:
            optimizer.zero_grad(set_to_none=True)
            target = target.to(model.module.dev1)
            output = model(data)
            loss = loss_fn(output, target)
            loss.backward()
            optimizer.step()
            if rank == 0:
                value = loss.detach()
                if wandb is not None:
                    wandb.log({"loss/loss": float(value)})
                if on_loss is not None:
                    on_loss(iteration, float(value))
            iteration_pbar.update(1)
            iteration += 1
            if iteration == max_iterations:
                training = False
                break
        epoch += 1
    iteration_pbar.close()
    dist.barrier()
    if wandb is not None:
        wandb.finish()
    print(f"[Process {rank}] Finished")


def main(on_loss=None):
    config = get_args()
    if config.dry_run:
        os.environ["WANDB_MODE"] = "dryrun"
    if "DDPB_SEED" in os.environ:
        torch.manual_seed(int(os.environ["DDPB_SEED"]))
    model = setup_distributed(config)
    dataloader, sampler = get_dataloader(config, ToyData())
    training_demo(model, dataloader, sampler, on_loss=on_loss)
    dist.destroy_process_group()
    return model


if __name__ == "__main__":
    main()
