"""Model-parallel backend: MultiGPUModel (two stages on two GPUs of ONE process) behind
DistributedDataParallel(model, device_ids=None) -- demo_one_model_multi_gpu.py:17-42,98.

Per iteration (reference :118-128: model(data) -> loss -> backward -> optimizer.step()):

    dev0  stage-0 forward   (2->10->10)            writes the [B,10] activation INTO dev1's memory
    dev1  stage-1 step      (10->10->10->1, loss)  backward + Adam; writes d(activation) INTO dev0's memory
    dev0  stage-0 step      recompute forward, backward from d(activation), Adam

Three launches of ddpb_segment_step ordered by two events; the transfers the reference does with
.to(dev1) (cudaMemcpyPeer) and its autograd reverse are the producing kernels' own NVLink stores.
Across processes (W > 1) each device's flat bucket is all-reduced with its own NCCL
communicator and updated by the fused average+Adam pass.
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .data import HostBatch, StagedBatch
from .engine import engine_for

LOSS_RING = 1024
LOSS_SENTINEL = 0x7FC0DEAD


class _Stage:
    """Flat state of one pipeline stage on its device."""

    def __init__(self, seq_module, dims, act_last, has_loss, owners, backend):
        tensors = list(seq_module.parameters())
        dev = tensors[0].device
        if dev.type != "cuda":
            raise _lib.DdpB200Error("MultiGPUModel stages must live on CUDA devices; there is no CPU path")
        self.device, self.dims, self.tensors = dev, dims, tensors
        self.engine = engine_for(dev.index)
        self.P = sum(p.numel() for p in tensors)
        with torch.cuda.device(dev):
            self.params = torch.empty(self.P, dtype=torch.float32, device=dev)
            self.grads = torch.zeros(self.P + 1, dtype=torch.float32, device=dev)
            off = 0
            for p in tensors:
                n = p.numel()
                self.params[off:off + n].copy_(p.data.reshape(-1))
                p.data = self.params[off:off + n].view(p.shape)
                p.grad = self.grads[off:off + n].view(p.shape)
                owners[id(p)] = backend
                off += n
            self.state = torch.zeros(4, dtype=torch.int64, device=dev)
        self.m = self.v = self.bias_table = None
        j = self.job = _lib.SegJob()
        j.params, j.grads, j.state = self.params.data_ptr(), self.grads.data_ptr(), self.state.data_ptr()
        j.nl, j.act_last, j.has_loss = len(dims) - 1, int(act_last), int(has_loss)
        for i, d in enumerate(dims):
            j.dims[i] = d

    def attach(self, kind, hyper):
        with torch.cuda.device(self.device):
            self.m = torch.zeros(self.P, dtype=torch.float32, device=self.device)
            self.v = torch.zeros(self.P, dtype=torch.float32, device=self.device)
        j = self.job
        j.m, j.v, j.opt, j.lr = self.m.data_ptr(), self.v.data_ptr(), kind, hyper["lr"]
        if kind == _lib.OPT_ADAM:
            j.beta1, j.beta2 = hyper["betas"]
            j.eps = hyper["eps"]
            table = _lib.adam_bias_table(hyper["lr"], *hyper["betas"])
            self.bias_table = torch.from_numpy(table).to(self.device)
            j.bias_table, j.bias_table_n = self.bias_table.data_ptr(), table.shape[0]
        else:
            j.beta1 = j.beta2 = 1.0
            j.momentum = hyper["momentum"]

    def launch(self, mode):
        self.job.mode = mode
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(self.engine.lib.ddpb_segment_step(ctypes.byref(self.job), stream), "ddpb_segment_step")


class PipelineBackend:
    def __init__(self, module, process_group, owners):
        self.module = module
        self.stage0 = _Stage(module.layers0, module.dims0, True, False, owners, self)
        self.stage1 = _Stage(module.layers1, module.dims1, False, True, owners, self)
        self.dev0, self.dev1 = self.stage0.device, self.stage1.device
        self.num_tensors = len(self.stage0.tensors) + len(self.stage1.tensors)
        self.engine = self.stage1.engine
        if self.dev0 != self.dev1:
            _lib.check(self.engine.lib.ddpb_enable_peer(self.dev0.index, self.dev1.index), "ddpb_enable_peer")
        self.process_group = process_group
        from .ddp import _world_of
        self.world, self.rank, self.backend_name = _world_of(process_group)
        self._sync_replicas()
        self.loss_host = torch.zeros(LOSS_RING, dtype=torch.float32, pin_memory=True)
        self.loss_bits = self.loss_host.numpy().view(np.uint32)
        self.stage1.job.loss_ring, self.stage1.job.loss_ring_n = self.loss_host.data_ptr(), LOSS_RING
        self._loss_out = ctypes.c_float()
        self.opt_kind = self.hyper = None
        self.seq = 0
        self.waiting = None
        self.grads_fresh = False
        self.cap = 0                    # rows the transfer buffers hold
        self.ev_act = torch.cuda.Event()
        self.ev_dact = torch.cuda.Event()

    # -- construction-time collective (reference :98) ------------------------------------
    def _sync_replicas(self):
        if self.world == 1:
            return
        for st in (self.stage0, self.stage1):
            if self.backend_name == "nccl":
                with torch.cuda.device(st.device):
                    comm = st.engine.ensure_comm(self.process_group)
                    _lib.check(st.engine.lib.ddpb_broadcast_f32(comm, st.params.data_ptr(), st.P, 0, st.engine.stream()),
                               "ddpb_broadcast_f32")
            else:
                host = st.params.cpu()
                dist.broadcast(host, src=0, group=self.process_group)
                st.params.copy_(host)

    def _buffers(self, B):
        if B > self.cap:
            h = self.stage0.dims[-1]
            self.cap = max(B, 256)
            self.x0 = torch.empty(self.cap, self.stage0.dims[0], dtype=torch.float32, device=self.dev0)
            self.y1 = torch.empty(self.cap, self.stage1.dims[-1], dtype=torch.float32, device=self.dev1)
            self.act = torch.empty(self.cap, h, dtype=torch.float32, device=self.dev1)      # written by dev0's kernel
            self.dact = torch.empty(self.cap, h, dtype=torch.float32, device=self.dev0)     # written by dev1's kernel

    # -- optimizer binding ---------------------------------------------------------------
    def attach_optimizer(self, kind, hyper):
        self.opt_kind, self.hyper = kind, hyper
        self.stage0.attach(kind, hyper)
        self.stage1.attach(kind, hyper)

    def optimizer_state_dict(self):
        return {"step": int(self.stage1.state[0].item())}

    def discard_pending_backward(self):
        self.waiting = None
        self.grads_fresh = False

    # -- the step ------------------------------------------------------------------------
    def _place(self, data, target):
        """x on dev0, y on dev1 (reference :41 x.to(dev0), :124 target.to(dev1)); returns B."""
        lib = self.engine.lib
        if isinstance(data, (HostBatch, StagedBatch)):
            stage = data.stage
            B = stage.B
            self._buffers(B)
            stage.materialize()                       # the loader gathers lazily
            with torch.cuda.device(self.dev0):
                _lib.check(lib.ddpb_memcpy_h2d_async(self.x0.data_ptr(), stage.host_x_ptr, B * stage.dx * 4,
                                                     torch.cuda.current_stream(self.dev0).cuda_stream), "h2d x")
            with torch.cuda.device(self.dev1):
                _lib.check(lib.ddpb_memcpy_h2d_async(self.y1.data_ptr(), stage.host_y_ptr, B * stage.dy * 4,
                                                     torch.cuda.current_stream(self.dev1).cuda_stream), "h2d y")
            # the pinned block may be refilled only after both asynchronous copies have run
            stage.guard_with_event(torch.cuda.current_stream(self.dev0), torch.cuda.current_stream(self.dev1))
            return B, self.x0, self.y1
        if isinstance(target, (HostBatch, StagedBatch)):
            target = target.tensor if isinstance(target, HostBatch) else target.tensor()
        x = data.to(self.dev0, dtype=torch.float32).contiguous()
        y = target.to(self.dev1, dtype=torch.float32).contiguous()
        self._buffers(x.shape[0])
        return x.shape[0], x, y

    def _run(self, loss, mode):
        B, x, y = self._place(loss.output.data, loss.target)
        j0, j1 = self.stage0.job, self.stage1.job
        slot = loss.slot = self.seq % LOSS_RING
        self.loss_bits[slot] = LOSS_SENTINEL
        self.seq += 1
        s0, s1 = torch.cuda.current_stream(self.dev0), torch.cuda.current_stream(self.dev1)
        # stage 0 forward: activation goes straight into dev1's buffer
        j0.B, j0.x, j0.out, j0.d_out, j0.d_in = B, x.data_ptr(), self.act.data_ptr(), None, None
        self.stage0.launch(_lib.MODE_FWD)
        self.ev_act.record(s0)
        s1.wait_event(self.ev_act)
        # stage 1: forward, loss, backward (+ update); d(activation) goes straight into dev0's buffer
        j1.B, j1.x, j1.y, j1.d_in = B, self.act.data_ptr(), y.data_ptr(), self.dact.data_ptr()
        self.stage1.launch(mode)
        self.ev_dact.record(s1)
        s0.wait_event(self.ev_dact)
        # stage 0: recompute forward, backward from d(activation) (+ update)
        j0.d_out = self.dact.data_ptr()
        self.stage0.launch(mode)
        self.keep = (x, y)
        self.grads_fresh = mode == _lib.MODE_GRAD
        if self.world > 1 and mode == _lib.MODE_GRAD:
            self._allreduce()

    def _allreduce(self):
        for st in (self.stage0, self.stage1):
            with torch.cuda.device(st.device):
                if self.backend_name == "nccl":
                    _lib.check(st.engine.lib.ddpb_allreduce_sum_f32(st.engine.comm, st.grads.data_ptr(), st.P,
                                                                    st.engine.stream()), "ddpb_allreduce_sum_f32")
                else:
                    host = st.grads.cpu()
                    dist.all_reduce(host, group=self.process_group)
                    st.grads.copy_(host)

    def timed_split(self, x, y, reps):
        """Median host-side latency (us) of the three launches of one iteration when each is waited for
        before the next is issued: stage-0 forward (its stores carry the activation to dev1), stage-1
        forward+backward+Adam (its stores carry the gradient back), stage-0 backward+Adam.  Launch and
        stream-synchronisation overhead are inside each figure, so the sum exceeds a pipelined step."""
        import time
        j0, j1 = self.stage0.job, self.stage1.job
        B = x.shape[0]
        self._buffers(B)
        out = []
        for _ in range(reps):
            t = [time.perf_counter()]
            j0.B, j0.x, j0.out, j0.d_out, j0.d_in = B, x.data_ptr(), self.act.data_ptr(), None, None
            self.stage0.launch(_lib.MODE_FWD)
            torch.cuda.synchronize(self.dev0); t.append(time.perf_counter())
            j1.B, j1.x, j1.y, j1.d_in = B, self.act.data_ptr(), y.data_ptr(), self.dact.data_ptr()
            self.stage1.launch(_lib.MODE_STEP)
            torch.cuda.synchronize(self.dev1); t.append(time.perf_counter())
            j0.d_out = self.dact.data_ptr()
            self.stage0.launch(_lib.MODE_STEP)
            torch.cuda.synchronize(self.dev0); t.append(time.perf_counter())
            out.append([(b - a) * 1e6 for a, b in zip(t, t[1:])])
            self.seq += 1
        med = np.median(np.asarray(out), axis=0)
        return {"stage0_forward": float(med[0]), "stage1_forward_backward_adam": float(med[1]),
                "stage0_backward_adam": float(med[2]), "note": "each phase = launch + stream synchronisation, host clock"}

    def request_backward(self, loss):
        if (self.waiting is not None and self.waiting is not loss) or self.grads_fresh:
            from .ddp import _SECOND_BACKWARD
            raise NotImplementedError(_SECOND_BACKWARD)
        if self.world == 1 and self.opt_kind is not None:
            self.waiting = loss
        else:
            self.waiting = None
            self._run(loss, _lib.MODE_GRAD)

    def request_step(self):
        if self.opt_kind is None:
            raise _lib.DdpB200Error("no optimizer attached")
        if self.waiting is not None:
            loss, self.waiting = self.waiting, None
            self._run(loss, _lib.MODE_STEP)
            return
        if not self.grads_fresh:
            return
        self.grads_fresh = False
        h, scale = self.hyper, 1.0 / self.world
        for st in (self.stage0, self.stage1):
            with torch.cuda.device(st.device):
                lib, stream = st.engine.lib, st.engine.stream()
                if self.opt_kind == _lib.OPT_ADAM:
                    _lib.check(lib.ddpb_fused_avg_adam_dev(st.grads.data_ptr(), st.params.data_ptr(), st.m.data_ptr(),
                                                           st.v.data_ptr(), st.P, scale, h["betas"][0], h["betas"][1],
                                                           h["eps"], st.job.bias_table, st.job.bias_table_n,
                                                           st.state.data_ptr(), 1, stream), "ddpb_fused_avg_adam_dev")
                else:
                    st.sgd_steps = getattr(st, "sgd_steps", 0) + 1
                    _lib.check(lib.ddpb_fused_avg_sgd(st.grads.data_ptr(), st.params.data_ptr(), st.m.data_ptr(), st.P,
                                                      scale, h["lr"], h["momentum"], st.sgd_steps, stream),
                               "ddpb_fused_avg_sgd")

    # -- results ---------------------------------------------------------------------------
    def read_loss(self, loss):
        if loss.slot is None:
            if self.waiting is loss:
                self.waiting = None
            self._run(loss, _lib.MODE_GRAD)
        s1 = torch.cuda.current_stream(self.dev1).cuda_stream
        _lib.check(self.engine.lib.ddpb_wait_f32(self.loss_host.data_ptr() + 4 * loss.slot, LOSS_SENTINEL, 200_000, s1,
                                                 ctypes.byref(self._loss_out)), "ddpb_wait_f32")
        return torch.tensor(self._loss_out.value)

    def loss_device_view(self, loss):
        return self.read_loss(loss).to(self.dev1)

    def forward_now(self, data):
        raise NotImplementedError("eager forward of the split model is not part of the reference loop")

    def check_peers(self):
        pass

    def flat_parameters(self):
        """Concatenated [371] parameter vector in module.parameters() order, on dev1."""
        torch.cuda.synchronize(self.dev0)
        torch.cuda.synchronize(self.dev1)
        return torch.cat([self.stage0.params.to(self.dev1), self.stage1.params])
