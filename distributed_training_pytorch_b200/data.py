"""Sampler and loaders with the call shapes of demo.py:139-154.

DistributedSampler reproduces torch.utils.data.DistributedSampler's index stream
bit-exactly (native host code, csrc/sampler.cu).  DataLoader replaces the reference's
per-sample __getitem__ + default_collate (0.29 ms per batch on the host, more than a whole
GPU step) by one native gather from the dense dataset into pinned staging rows; batches
come out as HostBatch objects whose .cuda() is one async H2D copy of the whole (x, y)
block.  With device_resident=True (or DDPB_DEVICE_DATA=1) the dataset and the sampler's
index rows live in HBM and a batch is just an index-row handle (no per-step copies).
"""
import math
import os

import numpy as np
import torch

from . import _lib


class DistributedSampler:
    """Same constructor and semantics as torch.utils.data.DistributedSampler
    (torch:utils/data/distributed.py:66-157); built at demo.py:144."""

    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False):
        if num_replicas is None or rank is None:
            import torch.distributed as dist
            if not dist.is_available() or not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available")
            num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
            rank = dist.get_rank() if rank is None else rank
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset, self.num_replicas, self.rank = dataset, num_replicas, rank
        self.epoch, self.drop_last, self.shuffle, self.seed = 0, drop_last, shuffle, seed
        n = len(dataset)
        if drop_last and n % num_replicas != 0:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    CACHE_EPOCHS = 64

    def _epoch_block(self, epoch0):
        rows = np.empty((self.CACHE_EPOCHS, self.num_samples), np.int64)
        _lib.check(_lib.lib().ddpb_sampler_epochs_i64(len(self.dataset), self.num_replicas, self.rank,
                                                      int(self.shuffle), self.seed, epoch0, self.CACHE_EPOCHS,
                                                      int(self.drop_last), rows.ctypes.data, self.num_samples),
                   "ddpb_sampler_epochs_i64")
        return rows

    def cached_indices(self):
        """Indices of the current epoch from a block of CACHE_EPOCHS epochs computed in one native
        call (bit-identical to indices()).  One epoch costs ~8 us of MT19937 + Fisher-Yates on the
        host -- as much as a whole GPU step -- so while a block is being consumed the NEXT one is
        computed on a helper thread (the native call releases the GIL): in a loop that walks the epochs
        in order the permutations are never on the critical path."""
        cache = getattr(self, "_cache", None)
        e, key = self.epoch, (self.seed, self.shuffle, self.drop_last)
        if cache is None or not (cache[0] <= e < cache[0] + self.CACHE_EPOCHS) or cache[2] != key:
            nxt = getattr(self, "_next", None)
            self._next = None
            rows = None
            if nxt is not None and nxt[0] <= e < nxt[0] + self.CACHE_EPOCHS and nxt[2] == key:
                rows = nxt[1].result()
                e0 = nxt[0]
            if rows is None:
                e0 = e
                rows = self._epoch_block(e0)
            cache = self._cache = (e0, rows, key)
        elif e - cache[0] == self.CACHE_EPOCHS // 2 and getattr(self, "_next", None) is None:
            pool = DistributedSampler._pool
            if pool is None:
                from concurrent.futures import ThreadPoolExecutor
                pool = DistributedSampler._pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="ddpb-sampler")
            e0 = cache[0] + self.CACHE_EPOCHS
            self._next = (e0, pool.submit(self._epoch_block, e0), key)
        return cache[1][e - cache[0]]

    _pool = None

    def indices(self, epoch=None):
        """int64 numpy array of this rank's indices for `epoch` (default: the current one)."""
        out = np.empty(self.num_samples, np.int64)
        got = _lib.lib().ddpb_sampler_indices(len(self.dataset), self.num_replicas, self.rank, int(self.shuffle),
                                              self.seed, self.epoch if epoch is None else epoch,
                                              int(self.drop_last), out.ctypes.data, out.size)
        if got != self.num_samples:
            _lib.check(2, "ddpb_sampler_indices")
        return out

    def epochs_i32(self, epoch0, nepochs):
        """[nepochs, num_samples] int32 index rows for consecutive epochs (device-resident path)."""
        out = np.empty((nepochs, self.num_samples), np.int32)
        _lib.check(_lib.lib().ddpb_sampler_epochs_i32(len(self.dataset), self.num_replicas, self.rank,
                                                      int(self.shuffle), self.seed, epoch0, nepochs,
                                                      int(self.drop_last), out.ctypes.data, self.num_samples),
                   "ddpb_sampler_epochs_i32")
        return out

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self):
        return self.num_samples

    def set_epoch(self, epoch):
        self.epoch = epoch


class HostBatch:
    """One half (data or target) of a staged batch in pinned host memory.

    Quacks like the pinned tensor the reference's DataLoader yields: .cuda()/.to(device)
    hand back the device-side handle (the toy kernel reads the pinned rows itself over PCIe; other
    consumers get one async H2D copy of the whole (x, y) block), .size()/.shape describe it; any
    other attribute is looked up on the pinned tensor itself."""

    __slots__ = ("stage", "which")

    def __init__(self, stage, which):
        self.stage, self.which = stage, which

    @property
    def tensor(self):
        self.stage.materialize()
        return self.stage.host[self.which]

    def cuda(self, device=None, non_blocking=False):
        return self.stage.staged(device)[self.which]

    def to(self, device=None, *args, **kwargs):
        dev = torch.device(device) if device is not None and not isinstance(device, torch.dtype) else None
        if dev is None or dev.type == "cpu":
            return self.tensor
        return self.stage.staged(dev)[self.which]

    def size(self, dim=None):
        return self.stage.host[self.which].size() if dim is None else self.stage.host[self.which].size(dim)

    @property
    def shape(self):
        return self.stage.host[self.which].shape

    def __len__(self):
        return self.stage.B

    def __getattr__(self, name):
        return getattr(self.tensor, name)


class StagedBatch:
    """Device-side handle of a HostBatch.  Nothing moves when it is made: the fused toy step reads the
    pinned rows in place, and a consumer that needs a real device tensor (`.tensor()` or any tensor
    attribute) triggers the H2D copy of the block then."""

    __slots__ = ("stage", "which")

    def __init__(self, stage, which):
        self.stage, self.which = stage, which

    def tensor(self):
        self.stage.copy_now()
        return self.stage.dev[self.which]

    def cuda(self, device=None, non_blocking=False):
        return self

    def to(self, *args, **kwargs):
        return self

    def size(self, dim=None):
        shape = self.stage.host[self.which].shape
        return shape if dim is None else shape[dim]

    @property
    def shape(self):
        return self.stage.host[self.which].shape

    def __len__(self):
        return self.stage.B

    def __getattr__(self, name):
        return getattr(self.tensor(), name)


class _EventGuard:
    __slots__ = ("events",)

    def __init__(self, events):
        self.events = events

    def synchronize(self):
        for ev in self.events:
            ev.synchronize()


class _Stage:
    """A pinned (x | y) block, the description of the rows that belong in it, and (on demand) its
    device twin; a DataLoader cycles through a few.

    The rows are gathered LAZILY: the loader only records which indices the block is to hold; the
    gather runs inside the native call that launches the consuming step (engine.flush), or on the
    first access to the host tensor.  A block is handed out again only after its last consumer has
    run on the GPU (`wait_free`): the toy kernel reads the block in place, and the copies issued for
    the other back ends are asynchronous, so without that wait a host that runs ahead of the GPU
    would overwrite rows that are still to be read."""

    def __init__(self, loader, batch, dx, dy):
        self.block = torch.empty(batch * (dx + dy), dtype=torch.float32, pin_memory=torch.cuda.is_available())
        self.loader = loader
        self.batch, self.dx, self.dy = batch, dx, dy
        self.nbytes = self.block.numel() * 4
        self.host_ptr = self.block.data_ptr()
        self.host_x_ptr, self.host_y_ptr = self.host_ptr, self.host_ptr + batch * dx * 4
        self._host_views, self._dev_views = {}, {}
        self.device = None          # device index of the twin
        self.dblock = None
        self.dev_x_ptr = self.dev_y_ptr = 0
        self.copied = True          # nothing to copy yet
        self.pending_idx = None     # (index array kept alive, address of the first index) until gathered
        self.guard = None           # who still reads the block: (engine, launch number) or a CUDA event
        self.halves = (HostBatch(self, 0), HostBatch(self, 1))
        self.twins = (StagedBatch(self, 0), StagedBatch(self, 1))
        self.B = -1
        self.set_rows(batch)

    def set_rows(self, B):
        if B != self.B:
            self.B = B
            views = self._host_views.get(B)
            if views is None:
                views = (self.block[:B * self.dx].view(B, self.dx),
                         self.block[self.batch * self.dx:self.batch * self.dx + B * self.dy].view(B, self.dy))
                self._host_views[B] = views
            self.host = views
            if self.dblock is not None:
                self.dev = self._device_views(B)
        self.copied = False

    def _device_views(self, B):
        views = self._dev_views.get(B)
        if views is None:
            views = (self.dblock[:B * self.dx].view(B, self.dx),
                     self.dblock[self.batch * self.dx:self.batch * self.dx + B * self.dy].view(B, self.dy))
            self._dev_views[B] = views
        return views

    # -- life cycle of one batch --------------------------------------------------------
    def wait_free(self):
        """Block until the GPU work that reads this block (in place or through its async copy) has run."""
        g = self.guard
        if g is not None:
            self.guard = None
            if type(g) is tuple:
                engine, launch_no = g
                if engine.completed < launch_no:
                    engine.sync()
            else:
                g.synchronize()

    def begin(self, B, idx, idx_addr):
        """The block is to hold rows idx[...] (B of them, first index at address idx_addr)."""
        if self.guard is not None:
            self.wait_free()
        self.set_rows(B)
        self.pending_idx = (idx, idx_addr)

    def materialize(self):
        """Gather the rows now (host code, csrc/sampler.cu) unless a launch already did."""
        p = self.pending_idx
        if p is not None:
            self.pending_idx = None
            ld = self.loader
            if _lib.lib().ddpb_host_gather(ld.x_ptr, ld.y_ptr, p[1], self.B, self.dx, self.dy, self.host_x_ptr, self.host_y_ptr):
                _lib.check(1, "ddpb_host_gather")

    def staged(self, device=None):
        if self.device is None:
            idx = torch.cuda.current_device() if device is None else torch.device(device).index
            self.device = torch.cuda.current_device() if idx is None else idx
        elif device is not None:
            idx = torch.device(device).index
            if idx is not None and idx != self.device:
                self.device, self.dblock = idx, None
        return self.twins

    def _twin(self):
        if self.dblock is None:
            self.dblock = torch.empty_like(self.block, device=f"cuda:{self.device}")
            self.dev_x_ptr = self.dblock.data_ptr()
            self.dev_y_ptr = self.dev_x_ptr + self.batch * self.dx * 4
            self._dev_views = {}
            self.dev = self._device_views(self.B)

    def copy_now(self):
        """Gather (if still pending) and issue the H2D copy of the block on the engine's stream, once."""
        if not self.copied:
            from .engine import engine_for
            if self.device is None:
                self.device = torch.cuda.current_device()
            self._twin()
            self.materialize()
            eng = engine_for(self.device)
            _lib.check(eng.lib.ddpb_memcpy_h2d_async(self.dev_x_ptr, self.host_ptr, self.nbytes, eng.stream()), "h2d")
            self.copied = True
            self.guard_with_event(eng.torch_stream())

    def guard_with_event(self, *torch_streams):
        """The block must not be refilled before what has been enqueued on these streams so far has run."""
        events = getattr(self, "_events", None)
        if events is None or len(events) != len(torch_streams):
            events = self._events = [torch.cuda.Event() for _ in torch_streams]
        for ev, st in zip(events, torch_streams):
            ev.record(st)
        self.guard = _EventGuard(events)


class DeviceBatch:
    """Handle of a batch of the HBM-resident dataset: `B` rows selected by an index row."""

    __slots__ = ("x", "y", "idx_ptr", "B", "which", "keep")

    def __init__(self, x, y, idx_ptr, B, which, keep):
        self.x, self.y, self.idx_ptr, self.B, self.which, self.keep = x, y, idx_ptr, B, which, keep

    def cuda(self, device=None, non_blocking=False):
        return self

    def to(self, *args, **kwargs):
        return self

    def size(self, dim=None):
        width = self.x.shape[1] if self.which == 0 else self.y.shape[1]
        shape = torch.Size((self.B, width))
        return shape if dim is None else shape[dim]

    @property
    def shape(self):
        return self.size()


def _dense(dataset):
    """Dense [n,dx], [n,dy] float32 CPU tensors of a map-style dataset of (x, y) pairs."""
    if hasattr(dataset, "x_dense") and hasattr(dataset, "y_dense"):
        return dataset.x_dense, dataset.y_dense
    xs, ys = zip(*(dataset[i] for i in range(len(dataset))))
    x = torch.stack([torch.as_tensor(v, dtype=torch.float32).reshape(-1) for v in xs]).contiguous()
    y = torch.stack([torch.as_tensor(v, dtype=torch.float32).reshape(-1) for v in ys]).contiguous()
    return x, y


class DataLoader:
    """DataLoader(dataset, batch_size, shuffle=False, sampler=None, pin_memory, num_workers, drop_last)
    with the argument meaning of torch's (call sites: demo.py:144-154).

    num_workers is accepted and ignored: batch assembly is one native gather (microseconds),
    there is nothing for worker processes to do.  shuffle=True without a sampler is not
    used by the reference and is rejected rather than approximated."""

    STAGES = 8
    EPOCH_ROWS = 256        # index rows uploaded at a time on the device-resident path

    def __init__(self, dataset, batch_size=1, shuffle=False, sampler=None, pin_memory=False, num_workers=0,
                 drop_last=False, device_resident=None, **unused):
        if shuffle and sampler is None:
            raise NotImplementedError("DataLoader(shuffle=True) without a sampler is not part of the reference path")
        self.dataset, self.batch_size, self.sampler, self.drop_last = dataset, int(batch_size), sampler, drop_last
        self.num_workers, self.pin_memory = num_workers, pin_memory
        self.x, self.y = _dense(dataset)
        self.dx, self.dy = self.x.shape[1], self.y.shape[1]
        self.x_ptr, self.y_ptr = self.x.data_ptr(), self.y.data_ptr()
        if device_resident is None:
            device_resident = os.environ.get("DDPB_DEVICE_DATA", "0") == "1"
        self.device_resident = bool(device_resident)
        self._stages = None
        self._next_stage = 0
        self._dev = None          # (x_dev, y_dev) on the device-resident path
        self._rows = None         # (epoch0, host int32 rows, device rows)
        self._rows_next = None    # (epoch0, future of the next block's host rows)
        self._row_bufs, self._row_turn = None, 0

    def __len__(self):
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def _indices(self):
        if self.sampler is None:
            return np.arange(len(self.dataset), dtype=np.int64)
        if hasattr(self.sampler, "cached_indices"):
            return self.sampler.cached_indices()
        if hasattr(self.sampler, "indices"):
            return self.sampler.indices()
        return np.fromiter(iter(self.sampler), dtype=np.int64)

    def __iter__(self):
        return self._iter_device() if self.device_resident else self._iter_host()

    def _iter_host(self):
        if self._stages is None:
            self._stages = [_Stage(self, self.batch_size, self.dx, self.dy) for _ in range(self.STAGES)]
        idx = self._indices()
        n, bs = len(idx), self.batch_size
        idx_addr = idx.ctypes.data
        stages, nstages = self._stages, self.STAGES
        for start in range(0, n, bs):
            B = min(bs, n - start)
            if B < bs and self.drop_last:
                return
            k = self._next_stage
            stage = stages[k]
            self._next_stage = k + 1 if k + 1 < nstages else 0
            stage.begin(B, idx, idx_addr + start * 8)       # rows are gathered by the call that launches the step
            yield stage.halves

    def _iter_device(self):
        if self._dev is None:
            self._dev = (self.x.cuda(), self.y.cuda())
        epoch = self.sampler.epoch if self.sampler is not None else 0
        per_epoch = len(self.sampler) if self.sampler is not None else len(self.dataset)
        if self._rows is None or not (self._rows[0] <= epoch < self._rows[0] + self._rows[1].shape[0]):
            rows = None
            nxt, self._rows_next = self._rows_next, None
            if nxt is not None and nxt[0] == epoch:
                rows = nxt[1].result()                      # computed on the helper thread while the last block was in use
            if rows is None:
                if self.sampler is not None and hasattr(self.sampler, "epochs_i32"):
                    rows = self.sampler.epochs_i32(epoch, self.EPOCH_ROWS)
                else:
                    rows = np.tile(self._indices().astype(np.int32), (1, 1))
            # two device buffers used in turn and never reallocated: a fresh device allocation here can cost tens of
            # ms once peer mappings exist (the driver maps every new segment into all peers), which the other
            # ranks would spend waiting inside their step; the block that was current is still readable by the
            # steps in flight while the new one is filled
            bufs = self._row_bufs
            if bufs is None or bufs[0][0].shape != rows.shape:
                bufs = self._row_bufs = [(torch.empty(rows.shape, dtype=torch.int32, device="cuda"),
                                          torch.empty(rows.shape, dtype=torch.int32, pin_memory=True)) for _ in range(2)]
                self._row_turn = 0
            drows, pinned = bufs[self._row_turn]
            self._row_turn ^= 1
            pinned.numpy()[...] = rows                      # host copy into pinned staging, then an ASYNC upload on the
            drows.copy_(pinned, non_blocking=True)          # launch stream (ordered before the steps that read the rows)
            self._rows = (epoch, rows, drows)
        elif (self._rows_next is None and self._rows[1].shape[0] > 1 and epoch - self._rows[0] == self._rows[1].shape[0] // 2
              and self.sampler is not None and hasattr(self.sampler, "epochs_i32")):
            # half-way through the block: the next block's permutations (2 ms of MT19937 + Fisher-Yates for 256
            # epochs) are computed on a helper thread, so the refresh at the block boundary is only the upload --
            # with several ranks a long refresh on one of them is time the others spend waiting inside their step
            pool = DistributedSampler._pool
            if pool is None:
                from concurrent.futures import ThreadPoolExecutor
                pool = DistributedSampler._pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="ddpb-sampler")
            e0 = self._rows[0] + self._rows[1].shape[0]
            self._rows_next = (e0, pool.submit(self.sampler.epochs_i32, e0, self.EPOCH_ROWS))
        epoch0, rows, drows = self._rows
        row = epoch - epoch0 if rows.shape[0] > 1 else 0
        base = drows.data_ptr() + row * rows.shape[1] * 4
        for start in range(0, per_epoch, self.batch_size):
            B = min(self.batch_size, per_epoch - start)
            if B < self.batch_size and self.drop_last:
                return
            ptr = base + start * 4
            yield (DeviceBatch(self._dev[0], self._dev[1], ptr, B, 0, drows),
                   DeviceBatch(self._dev[0], self._dev[1], ptr, B, 1, drows))
