#!/usr/bin/env python3
"""Generate the golden fixtures in this directory by running the REFERENCE ITSELF.

This script only runs in the build container (it needs /root/reference, which does
not exist on the GPU box).  It drives the reference's own ``demo.main()``
(/root/reference/demo.py:156-181) unmodified, on CPU with the gloo backend, under a
small shim that

  * turns the hard-coded CUDA calls (demo.py:66,69,71,102,103) into identities,
  * forces DistributedDataParallel(device_ids=None) and pin_memory=False,
  * seeds torch rank-independently (the reference parses --seed but never applies
    it: argument_parser.py:18, demo.py:59-61),
  * records what wandb.log receives (demo.py:120-121) and the final parameters.

Usage:  python tests/golden/make_golden.py            # spawns torchrun for W=1,2,4,8
        python tests/golden/make_golden.py standard   # only ref_w2_standard.npz (--dataloader=standard, W=2)
        python tests/golden/make_golden.py modelparallel   # only ref_model_parallel_w1.npz (demo_one_model_multi_gpu.py)
Outputs: ref_w{W}.npz (loss curves, initial / step-200 / final parameters), dataset.npz,
         sampler.npz, meta.json; ref_w2_standard.npz.
"""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
SEED = 1234
MID = 200     # parameters are also snapshotted after this many steps (tight pin before chaos sets in)


def _flat(model):
    import torch
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()]).numpy().copy()


def worker():
    import torch
    sys.path.insert(0, REF)
    torch.cuda.set_device = lambda d: None
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    import torch.nn.parallel as tnp
    import torch.utils.data as tud

    stock_ddp, stock_loader = tnp.DistributedDataParallel, tud.DataLoader

    class HostDDP(stock_ddp):
        def __init__(self, module, device_ids=None, **kw):
            super().__init__(module, device_ids=None, **kw)

    tnp.DistributedDataParallel = HostDDP
    tud.DataLoader = lambda *a, **k: stock_loader(*a, **{**k, "pin_memory": False})

    import demo
    import wandb

    logged = {"loss/lossX": [], "loss/lossY": []}
    wandb.init = lambda *a, **k: None
    wandb.finish = lambda *a, **k: None

    def _log(d, **k):
        for key, v in d.items():
            logged[key].append(float(v))
        # both optimizers have stepped when lossY is logged (demo.py:110-121): snapshot after MID steps
        if len(logged["loss/lossY"]) == MID and "mid_X" not in captured and "models" in captured:
            captured["mid_X"], captured["mid_Y"] = (_flat(m.module) for m in captured["models"])

    wandb.log = _log
    captured = {}
    stock_setup, stock_loop, stock_loader_fn = demo.setup_distributed, demo.training_demo, demo.get_dataloader

    def setup(config):
        mx, my, ws = stock_setup(config)
        captured["init_X"], captured["init_Y"] = _flat(mx.module), _flat(my.module)
        captured["models"] = (mx, my)
        return mx, my, ws

    def loader(config, dataset):
        captured["data_x"] = torch.stack(dataset.data).numpy().copy()
        captured["data_y"] = torch.stack(dataset.y).numpy().copy()
        return stock_loader_fn(config, dataset)

    def loop(mx, my, dl, sampler, ws):
        stock_loop(mx, my, dl, sampler, ws)
        captured["final_X"], captured["final_Y"] = _flat(mx.module), _flat(my.module)

    demo.setup_distributed, demo.training_demo, demo.get_dataloader = setup, loop, loader
    torch.manual_seed(SEED)
    sys.argv = [sys.argv[0], "--backend=gloo", "--torchrun", "--dry_run"] + os.environ.get("GOLDEN_ARGS", "").split()
    demo.main()
    captured.pop("models")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    np.savez(os.path.join(os.environ["GOLDEN_TMP"], f"w{world}_r{rank}.npz"),
             lossX=np.asarray(logged["loss/lossX"], np.float64),
             lossY=np.asarray(logged["loss/lossY"], np.float64), **captured)


def worker_model_parallel():
    """demo_one_model_multi_gpu.py (one model split over two devices, one Adam) executed as a script
    on CPU: Module.to / Tensor.to with an integer device become identities, device_count() says 2
    (its assert at :89), pin_memory is off, wandb is captured."""
    import runpy
    import torch
    sys.path.insert(0, REF)
    torch.cuda.device_count = lambda: 2
    stock_module_to, stock_tensor_to = torch.nn.Module.to, torch.Tensor.to
    torch.nn.Module.to = lambda self, *a, **k: self if (a and isinstance(a[0], int)) else stock_module_to(self, *a, **k)
    torch.Tensor.to = lambda self, *a, **k: self if (a and isinstance(a[0], int)) else stock_tensor_to(self, *a, **k)
    import torch.nn.parallel as tnp
    import torch.utils.data as tud
    import toy_model_and_data
    import wandb
    stock_ddp, stock_loader, stock_data = tnp.DistributedDataParallel, tud.DataLoader, toy_model_and_data.ToyData
    captured, logged = {}, []

    class HostDDP(stock_ddp):
        def __init__(self, module, device_ids=None, **kw):
            super().__init__(module, device_ids=None, **kw)
            captured["model"] = self
            captured["init"] = _flat(self.module)

    class RecordedData(stock_data):
        def __init__(self):
            super().__init__()
            captured["data_x"] = torch.stack(self.data).numpy().copy()
            captured["data_y"] = torch.stack(self.y).numpy().copy()

    tnp.DistributedDataParallel = HostDDP
    tud.DataLoader = lambda *a, **k: stock_loader(*a, **{**k, "pin_memory": False})
    toy_model_and_data.ToyData = RecordedData

    def _log(d, **k):
        logged.append(float(d["loss/loss"]))
        if len(logged) == MID:
            captured["mid"] = _flat(captured["model"].module)

    wandb.init = lambda *a, **k: None
    wandb.log = _log
    wandb.Settings = lambda *a, **k: None
    torch.manual_seed(SEED)
    sys.argv = [os.path.join(REF, "demo_one_model_multi_gpu.py"), "--backend=gloo", "--torchrun", "--dry_run"]
    runpy.run_path(sys.argv[0], run_name="__main__")
    model = captured.pop("model")
    np.savez(os.path.join(HERE, "ref_model_parallel_w1.npz"), loss=np.asarray(logged, np.float32),
             final=_flat(model.module), mid_step=np.int64(MID), **captured)
    print("ref_model_parallel_w1.npz: loss0", logged[0], "loss999", logged[-1], "final sha", sha(_flat(model.module)))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def _run_reference(world, extra_args=""):
    """One run of the reference under torchrun; returns the per-rank captures."""
    import shutil
    tmp = os.path.join(HERE, "..", "..", "gpurun_out", f"golden_tmp_w{world}")      # scratch inside the repo (git-ignored)
    shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(tmp)
    env = dict(os.environ, GOLDEN_TMP=tmp, OMP_NUM_THREADS="1", GOLDEN_WORKER="1", GOLDEN_ARGS=extra_args)
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1",
                    f"--nproc_per_node={world}", os.path.abspath(__file__)], check=True, env=env, cwd=tmp)
    ranks = [dict(np.load(os.path.join(tmp, f"w{world}_r{r}.npz"))) for r in range(world)]
    shutil.rmtree(tmp, ignore_errors=True)
    return ranks


def driver_standard():
    """--dataloader=standard at W=2 (demo.py:149-154): no sampler, every rank walks all 512 samples in
    order (2 steps per epoch, identical batches on both ranks)."""
    ranks = _run_reference(2, "--dataloader=standard")
    r0 = ranks[0]
    assert np.array_equal(ranks[1]["final_X"], r0["final_X"]) and np.array_equal(ranks[1]["final_Y"], r0["final_Y"])
    np.savez(os.path.join(HERE, "ref_w2_standard.npz"),
             lossX=r0["lossX"].astype(np.float32), lossY=r0["lossY"].astype(np.float32),
             init_X=r0["init_X"], init_Y=r0["init_Y"], final_X=r0["final_X"], final_Y=r0["final_Y"],
             mid_X=r0["mid_X"], mid_Y=r0["mid_Y"], mid_step=np.int64(MID))
    print("ref_w2_standard.npz: loss0", float(r0["lossX"][0]), float(r0["lossY"][0]), "final_X sha", sha(r0["final_X"]))


def driver():
    import tempfile
    import torch
    from torch.utils.data import DistributedSampler
    meta = {"torch": torch.__version__, "seed": SEED, "reference": "ammunk/distributed-training-pytorch",
            "runs": {}}
    for world in (1, 2, 4, 8):
        with tempfile.TemporaryDirectory() as tmp:
            env = dict(os.environ, GOLDEN_TMP=tmp, OMP_NUM_THREADS="1", GOLDEN_WORKER="1")
            subprocess.run([sys.executable, "-m", "torch.distributed.run", "--standalone",
                            "--local-addr", "127.0.0.1", f"--nproc_per_node={world}",
                            os.path.abspath(__file__)], check=True, env=env, cwd=tmp)
            ranks = [dict(np.load(os.path.join(tmp, f"w{world}_r{r}.npz"))) for r in range(world)]
        r0 = ranks[0]
        for r in ranks[1:]:   # DDP keeps replicas identical (demo.py:70,72 broadcast + allreduce)
            assert np.array_equal(r["final_X"], r0["final_X"]) and np.array_equal(r["final_Y"], r0["final_Y"])
        np.savez(os.path.join(HERE, f"ref_w{world}.npz"),
                 lossX=r0["lossX"].astype(np.float32), lossY=r0["lossY"].astype(np.float32),
                 init_X=r0["init_X"], init_Y=r0["init_Y"], final_X=r0["final_X"], final_Y=r0["final_Y"],
                 mid_X=r0["mid_X"], mid_Y=r0["mid_Y"], mid_step=np.int64(MID))
        meta["runs"][f"w{world}"] = {"final_X_sha": sha(r0["final_X"]), "final_Y_sha": sha(r0["final_Y"]),
                                     "loss0": [float(r0["lossX"][0]), float(r0["lossY"][0])],
                                     "loss999": [float(r0["lossX"][-1]), float(r0["lossY"][-1])]}
        if world == 1:
            np.savez(os.path.join(HERE, "dataset.npz"), x=r0["data_x"], y=r0["data_y"])
            meta["dataset"] = {"x_sha": sha(r0["data_x"]), "y_sha": sha(r0["data_y"])}
        else:
            # every rank builds the same dataset because the shim seeds rank-independently
            assert all(np.array_equal(r["data_x"], r0["data_x"]) for r in ranks)

    # sampler index streams: DistributedSampler(dataset, shuffle=True) as demo.py:144 builds it
    samp = {}
    for world in (1, 2, 4, 8):
        for rank in range(world):
            s = DistributedSampler(range(512), num_replicas=world, rank=rank, shuffle=True)
            rows = []
            for epoch in range(4):
                s.set_epoch(epoch)
                rows.append(np.fromiter(iter(s), dtype=np.int64))
            samp[f"w{world}_r{rank}"] = np.stack(rows)
    # ragged sizes exercise the padding branch (torch utils/data/distributed.py:118-128)
    for n, world in ((10, 4), (7, 8), (513, 8), (1, 2)):
        for rank in range(world):
            s = DistributedSampler(range(n), num_replicas=world, rank=rank, shuffle=True)
            s.set_epoch(3)
            samp[f"n{n}_w{world}_r{rank}"] = np.fromiter(iter(s), dtype=np.int64)
            s2 = DistributedSampler(range(n), num_replicas=world, rank=rank, shuffle=False)
            samp[f"n{n}_w{world}_r{rank}_noshuffle"] = np.fromiter(iter(s2), dtype=np.int64)
    for n in (1, 2, 5, 512, 1000, 4097):
        for seed in (0, 1, 7):
            g = torch.Generator(); g.manual_seed(seed)
            samp[f"perm_n{n}_s{seed}"] = torch.randperm(n, generator=g).numpy()
    np.savez_compressed(os.path.join(HERE, "sampler.npz"), **samp)
    meta["sampler_perm512_sha"] = {str(e): sha(samp[f"perm_n512_s{e}"]) for e in (0, 1)}
    with open(os.path.join(HERE, "meta.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps(meta, indent=1, sort_keys=True))


if __name__ == "__main__":
    if os.environ.get("GOLDEN_WORKER"):
        worker()
    elif os.environ.get("GOLDEN_MP_WORKER"):
        worker_model_parallel()
    elif sys.argv[1:] == ["standard"]:
        driver_standard()
    elif sys.argv[1:] == ["modelparallel"]:
        scratch = os.path.join(HERE, "..", "..", "gpurun_out", "golden_tmp_mp")     # wandb dry-run files land in the cwd
        os.makedirs(scratch, exist_ok=True)
        subprocess.run([sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1",
                        "--nproc_per_node=1", os.path.abspath(__file__)], check=True, cwd=scratch,
                       env=dict(os.environ, OMP_NUM_THREADS="1", GOLDEN_MP_WORKER="1"))
    else:
        driver()
