"""Worker for tests/test_gloo_world2.py (one process per rank, gloo, no GPU)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import distributed_training_pytorch_b200 as b200  # noqa: E402
from distributed_training_pytorch_b200 import bootstrap  # noqa: E402


def main():
    mode, out_dir = sys.argv[1], sys.argv[2]
    if mode == "slurm":
        layout = bootstrap.init_slurm_env("gloo", use_node_rank=os.environ.get("USE_NODE_RANK") == "1")
    elif mode == "torchrun":
        layout = bootstrap.init_torchrun("gloo")
    else:
        layout = bootstrap.init_mpi("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    assert (layout.rank, layout.world_size) == (rank, world)
    torch.manual_seed(1234)                                    # rank-independent, like the oracle shim
    mx, my, data = b200.ToyModel(), b200.ToyModel(), b200.ToyData()
    sampler = b200.DistributedSampler(data, shuffle=True)      # rank / world taken from the process group
    loader = b200.DataLoader(data, batch_size=256, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler)
    report = {"rank": rank, "world": world, "local_rank": layout.local_rank, "shards": []}
    for epoch in range(2):
        sampler.set_epoch(epoch)
        rows = []
        for x, y in loader:
            assert x.size(0) == y.size(0) == len(sampler)
            rows.append(x.tensor[:, 0].clone())
        mine = sampler.indices()
        assert torch.equal(torch.cat(rows), data.x_dense[mine, 0])
        gathered = [None] * world
        dist.all_gather_object(gathered, mine.tolist())
        report["shards"].append(gathered)
    # the logged-loss convention of demo.py:114-121: sum_r(loss_r * B) / (B * W)
    side = dist.new_group(backend="gloo")
    t = torch.tensor(float(rank + 1)) * 256
    dist.all_reduce(t, group=side, op=dist.ReduceOp.SUM)
    report["mean_loss"] = float(t / (256 * world))
    # the product path needs a GPU and says so
    try:
        b200.DistributedDataParallel(mx, device_ids=None)
        report["ddp_error"] = None
    except b200.DdpB200Error as e:
        report["ddp_error"] = str(e)
    dist.barrier()
    dist.destroy_process_group()
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump(report, f)


if __name__ == "__main__":
    main()
