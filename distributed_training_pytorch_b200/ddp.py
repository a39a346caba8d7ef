"""DistributedDataParallel with the call shapes of demo.py:70,72 and
demo_one_model_multi_gpu.py:98.

    model = DistributedDataParallel(ToyModel().cuda(), device_ids=[local_rank])
    out = model(data); loss = MSELoss()(out, target); loss.backward(); optimizer.step()

What the reference's wrapper does with ~30 tensors, hooks and a C++ Reducer
(torch:nn/parallel/distributed.py:655-900; reducer.hpp) becomes, for the toy model:

  * construction (collective): parameters are flattened into ONE fp32 HBM vector in
    module.parameters() order, every nn.Parameter is re-pointed at its slice, `.grad`
    at the matching slice of ONE flat gradient bucket (gradient_as_bucket_view), shapes are
    verified across ranks and rank 0's values are broadcast (demo.py:70; torch:...:879-890);
  * forward/loss/backward/step: described lazily, launched by engine.py as one fused
    kernel per iteration (single rank) or fused-gradient kernel -> ncclAllReduce of the
    bucket -> fused average+Adam pass (several ranks).

There is no fallback to PyTorch modules: a model this package has no kernels for, or a
CPU-resident module, raises.
"""
import torch
import torch.distributed as dist

from . import _lib
import ctypes
import os
import weakref

import numpy as np

from .data import DeviceBatch, HostBatch, StagedBatch
from .engine import engine_for
from .toy_model_and_data import MultiGPUModel, ToyModel

_OWNERS = weakref.WeakValueDictionary()          # id(nn.Parameter) -> backend (alive as long as its wrapper)
LOSS_RING = 1024      # loss slots per model; a LossValue stays readable for this many steps
LOSS_SENTINEL = _lib.LOSS_SENTINEL
LOSS_TIMEOUT = _lib.LOSS_TIMEOUT_BITS
LOSS_SPIN_NS = 200_000       # poll this long before falling back to a stream synchronisation
ZERO_COPY_ROWS = 1024        # loader batches up to this many rows are read by the kernel straight from pinned memory
PEER_CHECK_EVERY = 1024      # steps between looks at the device-side time-out flag when nobody reads losses
_SECOND_BACKWARD = ("a second backward() reached this model before optimizer.step() or zero_grad(): torch would "
                    "ACCUMULATE the two gradients in .grad, this package's gradient bucket is overwritten by every "
                    "backward. Gradient accumulation is not part of the reference loop (demo.py:100-111) and is "
                    "rejected rather than approximated")


LOCAL_ONLY = object()   # process_group=LOCAL_ONLY: this wrapper is not replicated across processes (no collectives),
                        # whatever torch.distributed says -- e.g. one rank of a job driving a model of its own


def _world_of(process_group):
    """(world, rank, backend name) the wrapper replicates over."""
    if process_group is LOCAL_ONLY or not (dist.is_available() and dist.is_initialized()):
        return 1, 0, None
    return dist.get_world_size(process_group), dist.get_rank(process_group), dist.get_backend(process_group)


def owner_of(param):
    try:
        return _OWNERS[id(param)]
    except KeyError:
        raise _lib.DdpB200Error("this parameter does not belong to a distributed_training_pytorch_b200."
                                "DistributedDataParallel model; wrap the model before building its optimizer") from None


class Output:
    """Lazy result of wrapper(data).  Holds the batch; values exist after `.tensor()`."""

    __slots__ = ("backend", "data")

    def __init__(self, backend, data):
        self.backend, self.data = backend, data

    def tensor(self):
        return self.backend.forward_now(self.data)

    def detach(self):
        return self.tensor()

    def cpu(self):
        return self.tensor().cpu()


class DistributedDataParallel:
    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 process_group=None, bucket_cap_mb=None, find_unused_parameters=False,
                 gradient_as_bucket_view=True, static_graph=False, lazy_step=None, **unused):
        self.module = module
        self.process_group = process_group
        self.bucket_cap_mb = bucket_cap_mb
        self.device_ids = device_ids
        if isinstance(module, ToyModel) and module.hidden == _lib.TOY_HIDDEN:
            self._backend = ToyBackend(module, process_group, lazy_step)
        elif isinstance(module, ToyModel):
            from .wide import WideBackend
            self._backend = WideBackend(module, process_group, bucket_cap_mb, _OWNERS)
        elif isinstance(module, MultiGPUModel) and module.hidden == _lib.TOY_HIDDEN:
            if device_ids is not None:
                raise ValueError("device_ids MUST be None when the module spans several GPUs "
                                 "(demo_one_model_multi_gpu.py:96-98)")
            from .pipeline import PipelineBackend
            self._backend = PipelineBackend(module, process_group, _OWNERS)
        else:
            raise NotImplementedError(
                f"no sm_100a kernels for {type(module).__name__}"
                f"{'(hidden=%d)' % module.hidden if hasattr(module, 'hidden') else ''}; "
                "there is no PyTorch fallback in this package")
        self.require_backward_grad_sync = True

    # -- nn.Module-like surface ---------------------------------------------------
    def __call__(self, data):
        return Output(self._backend, data)

    forward = __call__

    def parameters(self, recurse=True):
        return self.module.parameters(recurse)

    def named_parameters(self, *a, **k):
        return self.module.named_parameters(*a, **k)

    def state_dict(self, *a, **k):
        self._backend.engine.flush()
        return {"module." + key: v for key, v in self.module.state_dict(*a, **k).items()}

    def train(self, mode=True):
        self.module.train(mode)
        return self

    def eval(self):
        return self.train(False)

    def flat_parameters(self):
        """The flat fp32 parameter vector (after everything described so far has been launched)."""
        if hasattr(self._backend, "flat_parameters"):
            return self._backend.flat_parameters()
        self._backend.engine.flush()
        self._backend.check_peers()
        return self._backend.params


class ToyBackend:
    """Flat state + job description of one H=10 ToyModel on one device."""

    def __init__(self, module, process_group, lazy_step):
        tensors = list(module.parameters())
        dev = tensors[0].device
        if dev.type != "cuda" or any(p.device != dev for p in tensors):
            raise _lib.DdpB200Error("DistributedDataParallel needs the module on ONE CUDA device (call .cuda() first, "
                                    "demo.py:69-70); there is no CPU path")
        if any(p.dtype != torch.float32 for p in tensors):
            raise _lib.DdpB200Error("parameters must be float32, like the reference's")
        self.device = dev
        self.engine = engine_for(dev.index if dev.index is not None else torch.cuda.current_device())
        self.num_tensors = len(tensors)
        P = sum(p.numel() for p in tensors)
        assert P == _lib.TOY_PARAMS
        self.P = P
        with torch.cuda.device(dev):
            self.params = torch.empty(P, dtype=torch.float32, device=dev)
            self.grads = torch.zeros(P + 1, dtype=torch.float32, device=dev)   # [P] = sum (o-t)^2 of the batch
            off = 0
            for p in tensors:
                n = p.numel()
                self.params[off:off + n].copy_(p.data.reshape(-1))
                p.data = self.params[off:off + n].view(p.shape)
                p.grad = self.grads[off:off + n].view(p.shape)
                _OWNERS[id(p)] = self
                off += n
            self.state = torch.zeros(4, dtype=torch.int64, device=dev)       # ddpb_opt_state {step, seq, ...}
            # The loss ring lives in pinned host memory that the GPU writes directly (unified
            # addressing): reading a loss is a poll of one host word, not a copy + stream sync.
            # A loop that never reads its losses keeps them in HBM instead (a kernel whose last
            # store crosses PCIe retires ~2 us later, which a graph-replayed step would feel).
            self.loss_dev = torch.zeros(LOSS_RING, dtype=torch.float32, device=dev)
            self.loss_host = torch.zeros(LOSS_RING, dtype=torch.float32, pin_memory=True)
            self.loss_bits = self.loss_host.numpy().view(np.uint32)
            self.loss_ptr = self.loss_host.data_ptr()
            self._loss_out = ctypes.c_float()
            self._loss_out_ref = ctypes.byref(self._loss_out)
            self.m = self.v = None
        self.process_group = process_group
        self.world, self.rank, self.backend_name = _world_of(process_group)
        self._sync_replicas()
        self.fused_exchange = False     # W > 1: gradient all-reduce fused into the step kernel (peer windows)
        self.peer = None
        self.lazy = (self.world == 1) if lazy_step is None else bool(lazy_step) and self.world == 1
        self.opt_kind, self.hyper = None, None
        self.seq = 0                    # host mirror of state.seq
        self.unread = 0                 # losses described since the last one was read
        self.ring_on_host = True
        self.waiting = None             # Loss whose backward() was called and not yet described
        self.last_loss = None           # Loss of the most recently described job
        self.grads_fresh = False        # a GRAD job was described; an optimizer step may consume it
        self.keep = []                  # tensors referenced by the described job
        self.bound = None               # the loader block / batch handle the job currently points at
        self.slot, self.job = self.engine.job_slot(self)
        j = self.job
        j.params, j.grads = self.params.data_ptr(), self.grads.data_ptr()
        j.state, j.loss_ring, j.loss_ring_n = self.state.data_ptr(), self.loss_ptr, LOSS_RING
        j.idx_rows, j.idx_stride = 1, 0
        j.world, j.rank, j.inv_world = 1, 0, 1.0
        j.mode = -1
        self._mode = -1
        self._out = None
        # the mean loss over all ranks (demo.py:112-121), a by-product of the fused exchange or of the
        # bucket all-reduce: pinned ring the GPU writes / a 4-byte async copy lands in
        self.gloss_host = self.gloss_bits = None
        self.gloss_wait, self.gloss_ptr = False, 0    # engine.flush(wait) also waits for the all-rank mean of a fused STEP
        if self.world > 1:
            self.gloss_host = torch.zeros(LOSS_RING, dtype=torch.float32, pin_memory=True)
            self.gloss_bits = self.gloss_host.numpy().view(np.uint32)
        if self.world > 1 and os.environ.get("DDPB_ALLREDUCE", "p2p") == "p2p" and lazy_step is not False:
            self._map_peer_windows()

    # -- construction-time collective (demo.py:70,72) --------------------------------
    def _sync_replicas(self):
        if self.world == 1:
            return
        shapes = [None] * self.world
        dist.all_gather_object(shapes, int(self.P), group=self.process_group)
        if any(s != self.P for s in shapes):
            raise _lib.DdpB200Error(f"parameter counts differ across ranks: {shapes}")
        if self.backend_name == "nccl":
            comm = self.engine.ensure_comm(self.process_group)
            _lib.check(self.engine.lib.ddpb_broadcast_f32(comm, self.params.data_ptr(), self.P, 0, self.engine.stream()),
                       "ddpb_broadcast_f32")
        else:                                   # gloo / mpi: stage through the host group
            host = self.params.cpu()
            dist.broadcast(host, src=0, group=self.process_group)
            self.params.copy_(host)

    def _map_peer_windows(self):
        """Allocate this rank's exchange window, trade IPC handles through the process group and
        map every peer's window (all ranks of the group must share one NVLink/NVSwitch box)."""
        if self.world > _lib.MAX_PEERS:
            return
        lib = self.engine.lib
        window_bytes = _lib.XCHG_WINDOW_BYTES
        mine, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            _lib.check(lib.ddpb_peer_alloc(window_bytes, ctypes.byref(mine), handle), "ddpb_peer_alloc")
            import socket
            handles = [None] * self.world
            dist.all_gather_object(handles, (socket.gethostname(), handle.raw), group=self.process_group)
            ok = all(h[0] == handles[0][0] for h in handles)
            mapped = []
            if ok:
                for r, (_, raw) in enumerate(handles):
                    if r == self.rank:
                        mapped.append(mine.value)
                        continue
                    ptr = ctypes.c_void_p()
                    if lib.ddpb_peer_open(raw, ctypes.byref(ptr)) != 0:
                        ok = False
                        break
                    mapped.append(ptr.value)
            agree = [None] * self.world
            dist.all_gather_object(agree, bool(ok), group=self.process_group)
        if not all(agree):                       # no peer access somewhere: every rank falls back to NCCL together
            for r, ptr in enumerate(mapped):
                if r != self.rank:
                    lib.ddpb_peer_close(ptr)
            lib.ddpb_peer_free(mine)
            return
        self.peer = dict(mine=mine, mapped=mapped)
        j = self.job
        for r, ptr in enumerate(mapped):
            j.peer_data[r] = ptr
        j.world, j.rank, j.inv_world = self.world, self.rank, 1.0 / self.world
        j.global_loss_ring = self.gloss_ptr = self.gloss_host.data_ptr()
        self.fused_exchange = True
        self.lazy = True

    def check_peers(self):
        """Raise if a fused exchange timed out waiting for a peer (state.reserved[0] set by the kernel)."""
        if self.fused_exchange and int(self.state[2].item()) != 0:
            self._peer_timeout()

    def _peer_timeout(self):
        raise _lib.DdpB200Error("a peer GPU did not deliver its gradient slice within 2 s: the fused all-reduce gave up; "
                                "parameters and optimizer state are those of the last completed step")

    # -- optimizer binding -----------------------------------------------------------
    def attach_optimizer(self, kind, hyper):
        self.opt_kind, self.hyper = kind, hyper
        with torch.cuda.device(self.device):
            self.m = torch.zeros(self.P, dtype=torch.float32, device=self.device)
            self.v = torch.zeros(self.P, dtype=torch.float32, device=self.device)
        j = self.job
        j.m, j.v, j.opt = self.m.data_ptr(), self.v.data_ptr(), kind
        j.lr = hyper["lr"]
        if kind == _lib.OPT_ADAM:
            j.beta1, j.beta2 = hyper["betas"]
            j.eps = hyper["eps"]
            table = _lib.adam_bias_table(hyper["lr"], *hyper["betas"])
            self.bias_table = torch.from_numpy(table).to(self.device)
            j.bias_table, j.bias_table_n = self.bias_table.data_ptr(), table.shape[0]
        else:
            j.beta1 = j.beta2 = 1.0
            j.momentum = hyper["momentum"]

    def optimizer_state_dict(self):
        self.engine.flush()
        self.check_peers()
        return {"step": int(self.state[0].item()), "exp_avg": self.m, "exp_avg_sq": self.v}

    # -- describing work ---------------------------------------------------------------
    def _bind_batch(self, data, target):
        j = self.job
        if type(data) is HostBatch:
            data = data.cuda(self.device)
        if type(target) is HostBatch:
            target = target.cuda(self.device)
        if type(data) is StagedBatch:                       # the loader's own batch: trusted layout
            stage = data.stage
            if target is not None and (type(target) is not StagedBatch or target.stage is not stage):
                raise _lib.DdpB200Error("data and target must come from the same loader batch")
            if stage.B <= ZERO_COPY_ROWS:
                # the kernel reads the pinned rows in place (mapped host memory, one PCIe round trip
                # inside the launch) -- this IS the `.cuda()` of demo.py:102-103; the rows themselves
                # are gathered by the native call that launches the step
                if self.bound is not stage or j.B != stage.B:
                    j.x, j.y, j.idx, j.B, j.idx_stride, j.n_data = stage.host_x_ptr, stage.host_y_ptr, None, stage.B, 0, 0
                    self.bound = stage
                self.engine.attach_stage(stage)
            else:
                stage.copy_now()                           # large batches: one async H2D copy of the block
                j.x, j.y, j.idx, j.B, j.idx_stride, j.n_data = stage.dev_x_ptr, stage.dev_y_ptr, None, stage.B, 0, 0
                self.bound = None
            self.keep = stage
            return
        self.bound = None
        if type(data) is DeviceBatch:
            j.x, j.y, j.idx, j.B = data.x.data_ptr(), data.y.data_ptr(), data.idx_ptr, data.B
            j.idx_stride, j.n_data = data.B, data.x.shape[0]
            self.keep = data.keep
            return
        if type(target) is StagedBatch:
            target = target.tensor()
        if not (isinstance(data, torch.Tensor) and data.is_cuda and data.dtype == torch.float32
                and data.dim() == 2 and data.shape[1] == 2 and data.is_contiguous()):
            raise _lib.DdpB200Error("model input must be a contiguous float32 CUDA tensor [B,2] (or a loader batch)")
        j.x, j.idx, j.B, j.idx_stride, j.n_data = data.data_ptr(), None, data.shape[0], 0, 0
        self.keep = [data]
        if target is not None:
            if isinstance(target, DeviceBatch):
                raise _lib.DdpB200Error("target is a device-resident batch handle but data is not")
            if not (isinstance(target, torch.Tensor) and target.is_cuda and target.dtype == torch.float32
                    and target.numel() == data.shape[0] and target.is_contiguous()):
                raise _lib.DdpB200Error("target must be a contiguous float32 CUDA tensor [B,1]")
            j.y = target.data_ptr()
            self.keep.append(target)

    def _describe(self, loss, mode):
        eng = self.engine
        if eng.queue and self in eng.queue:      # the job struct still describes a step that has not gone out
            eng.flush()
        self._bind_batch(loss.output.data, loss.target)
        if mode != self._mode:
            self.job.mode = self._mode = mode
        seq = self.seq
        slot = loss.slot = seq % LOSS_RING
        loss.batch = self.job.B
        if self.unread > 8:                      # nobody has been reading losses: keep them on the device
            if self.ring_on_host:
                self.ring_on_host, self.job.loss_ring = False, self.loss_dev.data_ptr()
                if self.fused_exchange:
                    self.job.global_loss_ring = None
            if self.fused_exchange and (seq % PEER_CHECK_EVERY) == PEER_CHECK_EVERY - 1:
                self.engine.flush()
                self.check_peers()               # nobody polls the loss ring, so look at the device flag now and then
        else:
            self.unread += 1
            self.loss_bits[slot] = LOSS_SENTINEL
            if self.gloss_bits is not None:
                self.gloss_bits[slot] = LOSS_SENTINEL
        loss.on_host = self.ring_on_host
        self.gloss_wait = self.fused_exchange and self.ring_on_host and mode == _lib.MODE_STEP
        self.last_loss = loss
        self.seq = seq + 1
        self.grads_fresh = mode == _lib.MODE_GRAD
        self.engine.enqueue(self)
        if self.world > 1 and mode == _lib.MODE_GRAD:
            self.engine.flush()           # start the bucket all-reduce now: it overlaps whatever the host does next

    def after_launch(self, stream):
        """Called by the engine right after the fused kernel of this model went out."""
        self.keep = None
        if self.world > 1 and self._mode == _lib.MODE_GRAD and self._sync_grads:
            self._allreduce_bucket(stream)

    _sync_grads = True

    def _allreduce_bucket(self, stream):
        n = self.P + 1                    # the loss sum rides in the bucket's last element
        lib = self.engine.lib
        if self.backend_name == "nccl":
            _lib.check(lib.ddpb_allreduce_sum_f32(self.engine.comm, self.grads.data_ptr(), n, stream), "ddpb_allreduce_sum_f32")
        else:
            host = self.grads.cpu()
            dist.all_reduce(host, group=self.process_group)
            self.grads.copy_(host)
        loss = self.last_loss
        if loss is not None and loss.on_host:    # sum_r sum_b (o-t)^2 for the logged mean: 4 bytes, asynchronously
            _lib.check(lib.ddpb_memcpy_d2h_async(self.gloss_host.data_ptr() + 4 * loss.slot,
                                                 self.grads.data_ptr() + 4 * self.P, 4, stream), "d2h")
            loss.global_is_sum = True

    def request_backward(self, loss):
        if (self.waiting is not None and self.waiting is not loss) or self.grads_fresh:
            raise NotImplementedError(_SECOND_BACKWARD)
        if self.lazy and self.opt_kind is not None:
            self.waiting = loss                               # decide GRAD vs STEP when step() or a read arrives
        else:
            self._describe(loss, _lib.MODE_GRAD)

    def discard_pending_backward(self):
        self.waiting = None
        self.grads_fresh = False

    def request_step(self):
        loss = self.waiting
        if loss is not None:
            self.waiting = None
            self._describe(loss, _lib.MODE_STEP)
            return
        if self.opt_kind is None:
            raise _lib.DdpB200Error("no optimizer attached")
        if not self.grads_fresh:
            return                                            # torch skips parameters whose grad is None
        self.grads_fresh = False
        self.engine.flush()
        lib, stream, h = self.engine.lib, self.engine.stream(), self.hyper
        scale = 1.0 / self.world
        if self.opt_kind == _lib.OPT_ADAM:
            _lib.check(lib.ddpb_fused_avg_adam_dev(self.grads.data_ptr(), self.params.data_ptr(), self.m.data_ptr(),
                                                   self.v.data_ptr(), self.P, scale, h["betas"][0], h["betas"][1],
                                                   h["eps"], self.job.bias_table, self.job.bias_table_n,
                                                   self.state.data_ptr(), 1, stream),
                       "ddpb_fused_avg_adam_dev")
        else:
            # the device-resident step counter decides "first step", as in the fused kernel, so the
            # two paths can be mixed freely (momentum buffer initialised exactly once)
            _lib.check(lib.ddpb_fused_avg_sgd_dev(self.grads.data_ptr(), self.params.data_ptr(), self.m.data_ptr(), self.P,
                                                  scale, h["lr"], h["momentum"], self.state.data_ptr(), 1, stream),
                       "ddpb_fused_avg_sgd_dev")

    # -- reading results ----------------------------------------------------------------
    def _ensure_described(self, loss):
        if loss.slot is None:
            if self.waiting is loss:
                self.waiting = None
            self._describe(loss, _lib.MODE_GRAD)

    def _value(self, loss):
        """The local mean loss of `loss` as a Python float (launches / waits as needed)."""
        v = loss.value
        if v is None:
            if loss.slot is None:
                self._ensure_described(loss)
            eng = self.engine
            if eng.queue:
                # launch (+ wait, in the same native call, when the newest step is the one being read; an
                # older loss has long arrived and must not make the host wait for the step just launched)
                eng.flush(loss is self.last_loss)
                v = loss.value
            self.unread = 0
            if not self.ring_on_host:            # from the next step on, losses go straight to host memory again
                self.ring_on_host, self.job.loss_ring = True, self.loss_ptr
                if self.fused_exchange:
                    self.job.global_loss_ring = self.gloss_host.data_ptr()
            if v is None:
                if not loss.on_host:             # this one was written to HBM: copy it out the classic way
                    off = 4 * loss.slot
                    _lib.check(eng.lib.ddpb_memcpy_d2h_async(self.loss_ptr + off, self.loss_dev.data_ptr() + off, 4, eng.last_stream), "d2h")
                    eng.sync()
                    v = float(self.loss_host[loss.slot])
                else:
                    rc = eng.lib.ddpb_wait_f32(self.loss_ptr + 4 * loss.slot, LOSS_SENTINEL, LOSS_SPIN_NS, eng.last_stream,
                                               self._loss_out_ref)
                    if rc:
                        _lib.check(rc, "ddpb_wait_f32")
                    v = self._loss_out.value
                    if loss.launch_no > eng.completed:
                        eng.completed = loss.launch_no       # its kernel has run, hence every earlier launch too
                loss.value = v
        else:
            self.unread = 0
        if v != v and self.fused_exchange and int(self.loss_bits[loss.slot]) == LOSS_TIMEOUT:
            self._peer_timeout()                 # the kernel wrote the time-out sentinel instead of a loss
        return v

    def read_loss(self, loss):
        """Local mean loss as a 0-dim CPU tensor; blocks like the reference's .cpu() (demo.py:115)."""
        a = np.empty((), np.float32)
        a[()] = self._value(loss)
        return torch.from_numpy(a)

    def read_global_loss(self, loss):
        """Mean loss over ALL ranks as a 0-dim CPU tensor -- the value demo.py:112-121 logs -- without a
        host collective: the squared-error sum rides in the gradient bucket (element 371), so the
        gradient exchange has already reduced it."""
        if self.world == 1:
            return self.read_loss(loss)
        v = self._value(loss)                    # launches; the global value lands right after the local one
        if not loss.on_host:
            raise _lib.DdpB200Error("this loss was produced while nobody was reading losses (kept in HBM); read "
                                    "losses regularly or call .cpu() for the local value")
        if not (self.fused_exchange or loss.global_is_sum):
            raise _lib.DdpB200Error("no reduced loss for this step: it took no gradient exchange (backward only)")
        a = np.empty((), np.float32)
        if loss.gvalue is not None:              # the call that launched the step waited for it as well
            a[()] = loss.gvalue
        else:
            eng = self.engine
            rc = eng.lib.ddpb_wait_f32(self.gloss_host.data_ptr() + 4 * loss.slot, LOSS_SENTINEL, LOSS_SPIN_NS, eng.last_stream,
                                       self._loss_out_ref)
            if rc:
                _lib.check(rc, "ddpb_wait_f32")
            a[()] = self._loss_out.value
        if loss.global_is_sum:                   # NCCL / gloo route: the all-reduced sum; the mean is formed here
            a /= np.float32(loss.batch * self.world)
        elif int(self.gloss_bits[loss.slot]) == LOSS_TIMEOUT:
            self._peer_timeout()
        return torch.from_numpy(a)

    def loss_device_view(self, loss):
        return self.read_loss(loss).to(self.device)

    def forward_now(self, data):
        """Eager forward (MODE_FWD) for callers that want the output values."""
        self.engine.flush()
        self._bind_batch(data, None)
        B = self.job.B
        out = torch.empty(B, 1, dtype=torch.float32, device=self.device)
        self.job.out = out.data_ptr()
        self.job.mode = self._mode = _lib.MODE_FWD
        self.last_loss = None
        self.engine.enqueue(self)
        self.engine.flush()
        return out
