"""The drop-in front end (DDP wrapper / MSELoss / Adam / DataLoader) on one GPU: the loop
of demo.py:95-129 written against this package must reproduce the reference run's loss
curve (tests/golden/ref_w1.npz, seed 1234) within the fp32 tolerances of
test_gpu_toy_step.py, in all three data paths (pinned host batches, plain CUDA tensors,
HBM-resident dataset) and in both lazy (fused step) and eager (separate optimizer) modes."""
import os

import numpy as np
import pytest
import torch

import distributed_training_pytorch_b200 as b200
from distributed_training_pytorch_b200 import DataLoader, DistributedSampler, ToyData, ToyModel, nn, optim
from distributed_training_pytorch_b200 import DistributedDataParallel as DDP

pytestmark = pytest.mark.gpu


def _loop(iters, device_resident=False, lazy=None, plain_tensors=False, summed=False):
    torch.manual_seed(1234)
    mx, my = ToyModel(), ToyModel()
    data = ToyData()
    model_X, model_Y = DDP(mx.cuda(), device_ids=[0], lazy_step=lazy), DDP(my.cuda(), device_ids=[0], lazy_step=lazy)
    sampler = DistributedSampler(data, num_replicas=1, rank=0, shuffle=True)
    loader = DataLoader(data, batch_size=256, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler,
                        device_resident=device_resident)
    opt_X, opt_Y = optim.Adam(model_X.parameters(), lr=1e-3), optim.Adam(model_Y.parameters(), lr=1e-3)
    loss = nn.MSELoss()
    losses, it, epoch = [], 0, 0
    while it < iters:
        sampler.set_epoch(epoch)
        for x, t in loader:
            opt_X.zero_grad(set_to_none=True); opt_Y.zero_grad(set_to_none=True)
            x, t = x.cuda(), t.cuda()
            if plain_tensors:
                x, t = x.clone(), t.clone()
            if summed:                       # demo_pytorch_lightning.py:27-33: one summed loss, one backward
                l_X, l_Y = loss(model_X(x), t), loss(model_Y(x), t)
                total = l_X + l_Y
                total.backward()
            else:
                l_X = loss(model_X(x), t); l_X.backward()
                l_Y = loss(model_Y(x), t); l_Y.backward()
            opt_X.step(); opt_Y.step()
            losses.append((float(l_X.detach().cpu()), float(l_Y.detach().cpu())))
            it += 1
            if it == iters:
                break
        epoch += 1
    return np.asarray(losses, np.float32), model_X, model_Y


@pytest.mark.parametrize("mode", ["host_batches", "plain_tensors", "device_resident", "eager"])
def test_frontend_reproduces_reference_curve(golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    losses, model_X, model_Y = _loop(200, device_resident=mode == "device_resident",
                                     lazy=False if mode == "eager" else None,
                                     plain_tensors=mode == "plain_tensors")
    ref = np.stack([g["lossX"], g["lossY"]], axis=1)[:200]
    assert np.max(np.abs(losses - ref) / ref) <= 2e-5
    assert np.max(np.abs(model_X.flat_parameters().cpu().numpy() - g["mid_X"])) <= 2e-5
    assert np.max(np.abs(model_Y.flat_parameters().cpu().numpy() - g["mid_Y"])) <= 2e-5
    # nn.Parameter views and .grad views track the flat buffers
    w0 = model_X.module.layers[0].weight
    assert w0.data_ptr() == model_X.flat_parameters().data_ptr() and w0.grad is not None
    assert list(model_X.state_dict())[0] == "module.layers.0.weight"


def test_summed_loss_single_backward():
    """(loss_X + loss_Y).backward() then both optimizers (the Lightning variant of the reference,
    demo_pytorch_lightning.py:27-38) is the same computation as two backward calls: identical
    parameters, still one launch per iteration, and the sum reads back as the sum."""
    from distributed_training_pytorch_b200.engine import engine_for
    a, ax, ay = _loop(20)
    before = engine_for(0).launches
    b, bx, by = _loop(20, summed=True)
    assert engine_for(0).launches - before == 20
    assert np.array_equal(a, b)
    assert torch.equal(ax.flat_parameters(), bx.flat_parameters()) and torch.equal(ay.flat_parameters(), by.flat_parameters())
    x = torch.randn(64, 1, device="cuda").repeat(1, 2).contiguous()
    t = torch.randn(64, 1, device="cuda")
    loss = nn.MSELoss()
    l1, l2 = loss(bx(x), t), loss(by(x), t)
    total = sum([l1, l2])
    total.backward()
    assert abs(float(total.detach().cpu()) - (float(l1.detach().cpu()) + float(l2.detach().cpu()))) <= 1e-6


def test_one_launch_per_iteration_when_lazy():
    from distributed_training_pytorch_b200.engine import engine_for
    eng = engine_for(0)
    before = eng.launches
    _loop(10)
    assert eng.launches - before == 10           # both models, forward+backward+Adam: one kernel per iteration
    before = eng.launches
    _loop(10, lazy=False)
    assert eng.launches - before == 10 * (1 + 2 * 2)   # eager: 1 gradient launch (both models) + 2 x (adam + counter) per iteration


def test_eager_forward_values(oracle):
    from oracle import TOY_DIMS
    torch.manual_seed(3)
    model = DDP(ToyModel().cuda(), device_ids=[0])
    x = torch.randn(77, 2, device="cuda")
    out = model(x).tensor()
    ref = oracle.forward(model.flat_parameters().cpu().numpy(), TOY_DIMS, x.cpu().numpy())
    assert np.array_equal(out.cpu().numpy(), ref)


def test_unsupported_models_fail_loudly():
    with pytest.raises(NotImplementedError):
        DDP(torch.nn.Linear(2, 2).cuda())
    with pytest.raises(b200.DdpB200Error):
        DDP(ToyModel())                            # still on the CPU
