"""CPU-side checks: the C-ABI library loads and exports what include/ddp_b200.h declares,
host entry points (sampler, gather) are bit-exact, the front end fails loudly without CUDA."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import distributed_training_pytorch_b200 as b200
from distributed_training_pytorch_b200 import _lib, bootstrap
from distributed_training_pytorch_b200.argument_parser import get_args

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(native_lib):
    header = open(os.path.join(ROOT, "include", "ddp_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(ddpb_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(native_lib, name), f"{name} declared in include/ddp_b200.h but not exported"
    assert set(_lib.exported_symbols()) <= set(declared)
    assert native_lib.ddpb_version() == 100


def test_struct_layout_matches_header():
    # 12 pointers + 8 int32 + 5 doubles + 16 peer pointers + 4 words, naturally aligned
    assert ctypes.sizeof(_lib.ToyJob) == 12 * 8 + 8 * 4 + 5 * 8 + 16 * 8 + 16
    assert ctypes.sizeof(_lib.OptState) == 32


def test_ctypes_structs_agree_with_the_c_compiler(tmp_path):
    """Every job struct of include/ddp_b200.h, compiled as plain C: size and the offset of each field
    must be what the ctypes mirrors in _lib.py say (the header is the contract, gcc is the judge)."""
    import subprocess
    structs = {"ddpb_toy_job": _lib.ToyJob, "ddpb_seg_job": _lib.SegJob, "ddpb_bucket_job": _lib.BucketJob,
               "ddpb_opt_state": _lib.OptState, "ddpb_toy_iter": _lib.ToyIter}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "ddp_b200.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for field, _ in cls._fields_:
            lines.append(f'  printf("{cname}.{field} %zu\\n", offsetof({cname}, {field}));')
    lines += ['  return 0;', '}']
    src, exe = tmp_path / "layout.c", tmp_path / "layout"
    src.write_text("\n".join(lines))
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    out = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(out[cname]) == ctypes.sizeof(cls), cname
        for field, _ in cls._fields_:
            assert int(out[f"{cname}.{field}"]) == getattr(cls, field).offset, (cname, field)


class _Range:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def test_native_sampler_bit_exact(native_lib, golden_dir):
    gold = np.load(os.path.join(golden_dir, "sampler.npz"))
    for world in (1, 2, 4, 8):
        for rank in range(world):
            s = b200.DistributedSampler(_Range(512), num_replicas=world, rank=rank, shuffle=True)
            ref = gold[f"w{world}_r{rank}"]
            for epoch in range(ref.shape[0]):
                s.set_epoch(epoch)
                assert np.array_equal(s.indices(), ref[epoch])
                assert list(iter(s)) == ref[epoch].tolist()
            assert np.array_equal(s.epochs_i32(0, ref.shape[0]), ref.astype(np.int32))
    for key in gold.files:
        if key.startswith("perm_"):
            _, n, seed = key.split("_")
            out = np.empty(int(n[1:]), np.int64)
            assert native_lib.ddpb_randperm(int(seed[1:]), out.size, out.ctypes.data) == 0
            assert np.array_equal(out, gold[key]), key
        elif key.startswith("n"):
            parts = key.split("_")
            n, world, rank = int(parts[0][1:]), int(parts[1][1:]), int(parts[2][1:])
            shuffle = not key.endswith("noshuffle")
            s = b200.DistributedSampler(_Range(n), num_replicas=world, rank=rank, shuffle=shuffle)
            s.set_epoch(3 if shuffle else 0)
            assert np.array_equal(s.indices(), gold[key]), key


def test_sampler_matches_torch_live():
    from torch.utils.data import DistributedSampler as TorchSampler
    for n, world, drop_last in ((512, 8, False), (100, 3, False), (100, 3, True), (5, 8, False), (9, 4, True)):
        for rank in range(world):
            ours = b200.DistributedSampler(_Range(n), num_replicas=world, rank=rank, shuffle=True, seed=5, drop_last=drop_last)
            ref = TorchSampler(range(n), num_replicas=world, rank=rank, shuffle=True, seed=5, drop_last=drop_last)
            ours.set_epoch(7); ref.set_epoch(7)
            assert list(iter(ours)) == list(iter(ref)) and len(ours) == len(ref)


def test_sampler_argument_errors(native_lib):
    with pytest.raises(ValueError):
        b200.DistributedSampler(_Range(10), num_replicas=2, rank=2)
    out = np.empty(4, np.int64)
    assert native_lib.ddpb_sampler_indices(512, 2, 0, 1, 0, 0, 0, out.ctypes.data, out.size) == -1
    assert b"room" in native_lib.ddpb_last_error()


def test_toy_model_and_data_bit_exact(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_w1.npz"))
    d = np.load(os.path.join(golden_dir, "dataset.npz"))
    torch.manual_seed(1234)                      # RNG order of demo.py:22-23 then :174
    mx, my, data = b200.ToyModel(), b200.ToyModel(), b200.ToyData()
    flat = lambda m: torch.cat([p.detach().reshape(-1) for p in m.parameters()]).numpy()
    assert np.array_equal(flat(mx), g["init_X"]) and np.array_equal(flat(my), g["init_Y"])
    assert np.array_equal(data.x_dense.numpy(), d["x"]) and np.array_equal(data.y_dense.numpy(), d["y"])
    assert list(mx.state_dict()) == [f"layers.{i}.{k}" for i in (0, 2, 4, 6, 8) for k in ("weight", "bias")]
    x0, y0 = data[3]
    assert x0.shape == (2,) and y0.shape == (1,) and len(data) == 512


def test_host_loader_batches_equal_gathered_rows():
    torch.manual_seed(0)
    data = b200.ToyData()
    sampler = b200.DistributedSampler(data, num_replicas=4, rank=1, shuffle=True)
    loader = b200.DataLoader(data, batch_size=50, shuffle=False, pin_memory=True, num_workers=0, sampler=sampler)
    sampler.set_epoch(2)
    idx = sampler.indices()
    seen = 0
    for x, y in loader:
        rows = idx[seen:seen + x.size(0)]
        assert torch.equal(x.tensor, data.x_dense[rows]) and torch.equal(y.tensor, data.y_dense[rows])
        seen += x.size(0)
    assert seen == 128 and len(loader) == 3            # ragged tail kept (drop_last=False)
    plain = b200.DataLoader(data, pin_memory=True, batch_size=256, shuffle=False)
    sizes = [t.size(0) for _, t in plain]
    assert sizes == [256, 256]


def test_flags_are_the_reference_flags():
    a = get_args([])
    assert (a.dataloader, a.backend, a.torchrun, a.use_node_rank, a.num_workers, a.dry_run) == \
        ("distributed", "nccl", False, False, 0, False)
    assert 0 <= a.seed < 2 ** 32
    b = get_args(["--dataloader=standard", "--backend", "gloo", "--torchrun", "--use_node_rank", "--seed", "7",
                  "--num_workers", "2", "--dry_run"])
    assert (b.dataloader, b.backend, b.torchrun, b.use_node_rank, b.seed, b.num_workers, b.dry_run) == \
        ("standard", "gloo", True, True, 7, 2, True)
    with pytest.raises(SystemExit):
        get_args(["--backend", "ucc"])


def test_slurm_rank_arithmetic(monkeypatch):
    env = dict(TASKS_PER_NODE="4", SLURM_LOCALID="3", NODE_RANK="2", SLURM_PROCID="9", WORLD_SIZE="16")
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    assert bootstrap.layout_from_slurm_env(True).rank == 2 * 4 + 3
    lay = bootstrap.layout_from_slurm_env(False)
    assert (lay.rank, lay.world_size, lay.local_rank, lay.local_world_size) == (9, 16, 3, 4)
    monkeypatch.delenv("MASTER_ADDR", raising=False)
    with pytest.raises(ValueError, match="MASTER_ADDR and MASTER_PORT"):
        bootstrap.init_slurm_env("gloo", False)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(b200.DdpB200Error, match="no CPU path|CUDA"):
        b200.DistributedDataParallel(b200.ToyModel(), device_ids=None)
    with pytest.raises(RuntimeError):
        b200.ToyModel()(torch.zeros(4, 2))


def test_adam_bias_table_is_torchs_python_arithmetic():
    table = _lib.adam_bias_table(1e-3, 0.9, 0.999)
    for t in (1, 2, 10, 999, 5000, table.shape[0]):
        bc1, bc2 = 1 - 0.9 ** t, 1 - 0.999 ** t                   # torch:optim/adam.py:531-536
        assert table[t - 1, 0] == np.float32(-(1e-3 / bc1)) and table[t - 1, 1] == np.float32(bc2 ** 0.5)
    assert table[-1].tolist() == [np.float32(-1e-3), 1.0]        # converged: later steps reuse the last row


def test_cached_epoch_blocks_equal_single_epoch_calls():
    s = b200.DistributedSampler(_Range(512), num_replicas=4, rank=3, shuffle=True)
    for epoch in (0, 1, 63, 64, 65, 200, 5):
        s.set_epoch(epoch)
        assert np.array_equal(s.cached_indices(), s.indices())


def test_bucket_plan_covers_the_flat_gradient_in_reverse_layer_order():
    from distributed_training_pytorch_b200.wide import plan_buckets
    H = 1024
    sizes = [2 * H, H, H * H, H, H * H, H, H * H, H, H, 1]
    offsets = [int(x) for x in np.cumsum([0] + sizes[:-1])]
    total = sum(sizes)
    for cap in (0.5, 1, 5, 25, 1e9):
        plan = plan_buckets(offsets, total, cap)
        assert [l for l, _, _ in plan] == sorted((l for l, _, _ in plan), reverse=True)
        covered = sorted((s, s + n) for _, s, n in plan)
        assert covered[0][0] == 0 and covered[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(covered, covered[1:]))           # no gaps, no overlap
        assert all(n * 4 <= max(cap * (1 << 20), 16) for _, _, n in plan)
    assert len(plan_buckets(offsets, total, 1e9)) == 5 and len(plan_buckets(offsets, total, 1)) == 3 * 5 + 2


def test_sampler_blocks_prefetched_on_the_helper_thread_are_the_same_permutations():
    """Walking the epochs in order (what every training loop does) makes the sampler compute the next block
    of 64 epochs on a helper thread from the middle of the current one; across three block boundaries every
    epoch must still be DistributedSampler's permutation, bit for bit, and jumping around must not be
    served from a stale prefetch."""
    s = b200.DistributedSampler(_Range(512), num_replicas=2, rank=1, shuffle=True)
    for epoch in range(0, 200):
        s.set_epoch(epoch)
        assert np.array_equal(s.cached_indices(), s.indices(epoch)), epoch
    assert getattr(s, "_next", None) is None or s._next[0] == 256      # a block ahead at most
    for epoch in (500, 33, 97, 96, 1000):
        s.set_epoch(epoch)
        assert np.array_equal(s.cached_indices(), s.indices(epoch)), epoch


def test_host_loader_gathers_lazily_and_exactly_once():
    """The loader records WHICH rows a pinned block is to hold; the gather itself runs when the host tensor is
    first touched (or, on the GPU path, inside the call that launches the step).  Touching a batch twice must
    not gather twice into a block that has meanwhile been handed out again."""
    torch.manual_seed(1)
    data = b200.ToyData()
    sampler = b200.DistributedSampler(data, num_replicas=1, rank=0, shuffle=True)
    loader = b200.DataLoader(data, batch_size=128, shuffle=False, pin_memory=False, num_workers=0, sampler=sampler)
    sampler.set_epoch(3)
    idx = sampler.indices()
    it = iter(loader)
    x0, y0 = next(it)
    assert x0.stage.pending_idx is not None                 # nothing gathered yet
    assert x0.shape == (128, 2) and len(y0) == 128          # shapes need no gather
    assert x0.stage.pending_idx is not None
    first = x0.tensor.clone()
    assert x0.stage.pending_idx is None and torch.equal(first, data.x_dense[idx[:128]])
    assert torch.equal(y0.tensor, data.y_dense[idx[:128]])
    x1, _ = next(it)
    assert x1.stage is not x0.stage                         # the ring hands out another block
    assert torch.equal(x1.tensor, data.x_dense[idx[128:256]]) and torch.equal(x0.tensor, first)


def test_layout_device_wraps_only_when_asked(monkeypatch):
    from distributed_training_pytorch_b200.bootstrap import Layout
    lay = Layout(rank=5, world_size=8, local_rank=5, local_world_size=8)
    monkeypatch.delenv("DDPB_SHARE_GPUS", raising=False)
    assert lay.device == 5
    monkeypatch.setenv("DDPB_SHARE_GPUS", "1")
    monkeypatch.setattr(torch.cuda, "device_count", lambda: 2)
    assert lay.device == 1


def test_local_only_wrappers_ignore_the_process_group():
    from distributed_training_pytorch_b200.ddp import LOCAL_ONLY, _world_of
    assert _world_of(LOCAL_ONLY) == (1, 0, None)
    assert _world_of(None) == (1, 0, None)                  # no process group initialised in this test process
