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


def test_second_backward_without_step_is_rejected():
    """torch accumulates .grad over backward() calls; this package's bucket is overwritten by each one.
    The case is rejected loudly (NotImplementedError) instead of silently applying only the last
    gradient -- in both the lazy (fused) and the eager mode; zero_grad() or step() re-arms."""
    for lazy in (None, False):
        torch.manual_seed(5)
        model = DDP(ToyModel().cuda(), device_ids=[0], lazy_step=lazy)
        opt = optim.Adam(model.parameters(), lr=1e-3)
        loss = nn.MSELoss()
        x, t = torch.randn(64, 2, device="cuda"), torch.randn(64, 1, device="cuda")
        l1 = loss(model(x), t)
        l1.backward()
        with pytest.raises(NotImplementedError, match="ACCUMULATE"):
            loss(model(x), t).backward()
        opt.zero_grad(set_to_none=True)
        l2 = loss(model(x), t)
        l2.backward()
        opt.step()
        loss(model(x), t).backward()            # a step consumed the gradient: the next backward is fine
        opt.step()
        model.flat_parameters()                 # launches what is still only described
        assert int(model._backend.state[0].item()) == 2


def test_sgd_momentum_mixing_fused_and_eager_steps_matches_torch():
    """One device-resident step counter serves the fused step (backward(); step()) and the eager one
    (backward(); read the loss; step()): mixing them must initialise the momentum buffer exactly once.
    Checked against torch.optim.SGD(momentum=0.9) on the same fp32 model, 1e-6 abs on the parameters
    (lr 1e-2 bounds each update)."""
    torch.manual_seed(9)
    base = ToyModel()
    import copy
    twin = copy.deepcopy(base.layers).cuda()           # stock torch modules with the same initial values
    model = DDP(base.cuda(), device_ids=[0])
    opt = optim.SGD(model.parameters(), lr=1e-2, momentum=0.9)
    opt_t = torch.optim.SGD(twin.parameters(), lr=1e-2, momentum=0.9)
    loss = nn.MSELoss()
    g = torch.Generator(device="cuda").manual_seed(2)
    for it in range(8):
        x = torch.randn(128, 2, device="cuda", generator=g)
        t = torch.randn(128, 1, device="cuda", generator=g)
        opt.zero_grad(set_to_none=True); opt_t.zero_grad(set_to_none=True)
        l = loss(model(x), t)
        l.backward()
        if it in (0, 3, 4):                      # read before step(): gradient launch now, eager optimizer pass later
            float(l.detach().cpu())
        opt.step()
        lt = torch.nn.functional.mse_loss(twin(x), t)
        lt.backward()
        opt_t.step()
        assert abs(float(l.detach().cpu()) - lt.item()) <= 1e-5 * abs(lt.item()), it
    flat_t = torch.cat([p.detach().reshape(-1) for p in twin.parameters()])
    assert (model.flat_parameters() - flat_t).abs().max().item() <= 1e-6
    assert int(model._backend.state[0].item()) == 8


def test_loader_blocks_are_not_refilled_before_the_gpu_has_read_them(golden_dir):
    """The loader cycles through 8 pinned blocks that the toy kernel reads IN PLACE; a loop that never
    reads a loss lets the host run many iterations ahead of the GPU.  A block must not be refilled
    until the step that reads it has run: 300 iterations without a single loss read must end exactly
    where the loop that reads every loss ends."""
    read_every, mx, my = _loop(300)
    torch.manual_seed(1234)
    ax, ay = ToyModel(), ToyModel()
    data = ToyData()
    model_X, model_Y = DDP(ax.cuda(), device_ids=[0]), DDP(ay.cuda(), device_ids=[0])
    sampler = DistributedSampler(data, num_replicas=1, rank=0, shuffle=True)
    loader = DataLoader(data, batch_size=256, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler)
    opt_X, opt_Y = optim.Adam(model_X.parameters(), lr=1e-3), optim.Adam(model_Y.parameters(), lr=1e-3)
    loss = nn.MSELoss()
    slow = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    it, epoch = 0, 0
    while it < 300:
        sampler.set_epoch(epoch)
        for x, t in loader:
            if it % 16 == 0:
                for _ in range(20):
                    slow.fill_(it & 0xff)            # keep the GPU far behind the host
            opt_X.zero_grad(set_to_none=True); opt_Y.zero_grad(set_to_none=True)
            x, t = x.cuda(), t.cuda()
            loss(model_X(x), t).backward()
            loss(model_Y(x), t).backward()
            opt_X.step(); opt_Y.step()
            it += 1
            if it == 300:
                break
        epoch += 1
    assert torch.equal(model_X.flat_parameters(), mx.flat_parameters())
    assert torch.equal(model_Y.flat_parameters(), my.flat_parameters())


def test_global_mean_on_one_rank_is_the_local_loss():
    torch.manual_seed(6)
    model = DDP(ToyModel().cuda(), device_ids=[0])
    opt = optim.Adam(model.parameters(), lr=1e-3)
    x, t = torch.randn(64, 2, device="cuda"), torch.randn(64, 1, device="cuda")
    l = nn.MSELoss()(model(x), t)
    l.backward()
    opt.step()
    assert float(l.detach().global_mean()) == float(l.detach().cpu())


def test_device_resident_index_rows_across_block_boundaries():
    """The HBM-resident loader uploads the sampler's index rows 256 epochs at a time and computes the next
    block on a helper thread from the middle of the current one: every batch handle of 700 consecutive epochs
    must point at exactly DistributedSampler's indices of that epoch (bit-exact, as everywhere)."""
    data = ToyData()
    sampler = DistributedSampler(data, num_replicas=2, rank=1, shuffle=True)
    loader = DataLoader(data, batch_size=256, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler,
                        device_resident=True)
    for epoch in range(700):
        sampler.set_epoch(epoch)
        batches = list(loader)
        assert len(batches) == 1
        x, _ = batches[0]
        rows = x.keep
        first = (x.idx_ptr - rows.data_ptr()) // 4
        got = rows.reshape(-1)[first:first + x.B].cpu().numpy()
        assert np.array_equal(got, sampler.indices(epoch).astype(np.int32)), epoch
