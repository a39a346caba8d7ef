"""torchrun worker: the widened model trained data-parallel (per bucket: the fused peer-window
all-reduce + Adam kernel, or ncclAllReduce + fused avg+Adam with DDPB_ALLREDUCE=nccl, on the
communication stream); dumps the loss curve and final parameters of each rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import distributed_training_pytorch_b200 as b200  # noqa: E402


def main():
    out_dir, H, B, steps, cap = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(100 + rank)                 # different initial weights per rank: the wrapper must broadcast rank 0's
    model = b200.DistributedDataParallel(b200.ToyModel(hidden=H).cuda(), device_ids=[local], bucket_cap_mb=(cap or None))   # 0 = default plan
    init = model.flat_parameters().cpu().numpy().copy()
    opt = b200.optim.Adam(model.parameters(), lr=1e-3)
    loss_fn = b200.nn.MSELoss()
    g = torch.Generator().manual_seed(7)          # the same data stream on every rank; rank r takes its slice
    losses = []
    for _ in range(steps):
        z = torch.randn(B * dist.get_world_size(), generator=g)
        y = (0.5 * torch.randn(B * dist.get_world_size(), generator=g) + z * z)
        sl = slice(rank * B, (rank + 1) * B)
        x = torch.stack([z[sl], z[sl]], 1).contiguous().cuda()
        t = y[sl].view(B, 1).contiguous().cuda()
        opt.zero_grad(set_to_none=True)
        loss = loss_fn(model(x), t)
        loss.backward()
        opt.step()
        losses.append(float(loss.detach().cpu()))
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), losses=np.asarray(losses, np.float32), init=init,
             final=model.flat_parameters().cpu().numpy(), buckets=len(model._backend.buckets),
             fused=int(model._backend.fused_exchange), exp_avg=model._backend.optimizer_state_dict()["exp_avg"].cpu().numpy())
    model._backend.check_peers()
    model._backend.close()                        # collective teardown of the peer windows (no-op on the NCCL path)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
