"""Widened toy model (hidden width H, BASELINE.json configs[4]) behind the same wrapper.

The reference has no width knob (toy_model_and_data.py:13-21 are literals); ToyModel(hidden=H) /
DDPB_HIDDEN is the additive one.  With H in the thousands the step is no longer latency-bound:
the three H x H layers are tcgen05/TMA tf32 GEMMs (csrc/gemm_tf32.cu), the thin ends are
streaming kernels (csrc/wide_edges.cu), weight gradients are written straight into the flat
bucket, and the optimizer is the fused average+Adam pass (28 B/param).  With several ranks the
flat gradient is cut into buckets (bucket_cap_mb, reverse parameter order like the c10d reducer,
torch:nn/parallel/distributed.py:1255-1283); each bucket's all-reduce and its Adam pass run on
a communication stream as soon as the kernels that produce the bucket have been issued, i.e.
overlapped with the rest of backward (demo.py:107-111 -> reducer hooks + optimizer.step()).
On one NVSwitch box that is ONE kernel per bucket (ddpb_bucket_reduce_adam: reduce-scatter by
pull over peer windows, Adam on the owned shard, all-gather of the new parameters by push);
DDPB_ALLREDUCE=nccl selects ncclAllReduce + the fused average+Adam pass instead.
"""
import ctypes
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .data import DeviceBatch, HostBatch, StagedBatch
from .engine import engine_for

LOSS_RING = 1024
LOSS_SENTINEL = 0x7FC0DEAD
CAP_B = 256          # local batch capacity of the work buffers (the reference's batch_size, demo.py:145)


def plan_buckets(offsets, total, cap_mb):
    """Cut the flat gradient into all-reduce buckets: [(layer, start, count)] in REVERSE layer order
    (gradients become ready last layer first, torch:nn/parallel/distributed.py:1255-1283), each
    layer's [weight | bias] range cut into pieces of at most cap_mb MiB (counts stay multiples of 4
    floats so every piece is 16-byte aligned).  `offsets[2l]` is the start of layer l's weight."""
    cap = max(4, int(cap_mb * (1 << 20) / 4) // 4 * 4)
    n_layers = len(offsets) // 2
    out = []
    for l in range(n_layers - 1, -1, -1):
        start = offsets[2 * l]
        end = offsets[2 * l + 2] if l + 1 < n_layers else total
        while start < end:
            n = min(cap, end - start)
            out.append((l, start, n))
            start += n
    return out


class WideBackend:
    def __init__(self, module, process_group, bucket_cap_mb, owners):
        tensors = list(module.parameters())
        dev = tensors[0].device
        if dev.type != "cuda" or any(p.device != dev for p in tensors):
            raise _lib.DdpB200Error("DistributedDataParallel needs the module on ONE CUDA device; there is no CPU path")
        H = module.hidden
        if H < 32 or H % 32:
            raise NotImplementedError(f"hidden width {H}: the tensor-core path needs a multiple of 32 (>= 32); "
                                      "width 10 runs on the fused toy kernel")
        self.H, self.device, self.module = H, dev, module
        self.engine = engine_for(dev.index if dev.index is not None else torch.cuda.current_device())
        self.lib = self.engine.lib
        self.num_tensors = len(tensors)
        self.P = P = sum(p.numel() for p in tensors)
        assert P == 3 * H * H + 7 * H + 1
        self.process_group = process_group
        from .ddp import _world_of
        self.world, self.rank, self.backend_name = _world_of(process_group)
        self.peer = None
        with torch.cuda.device(dev):
            if not self._map_peer_windows():
                self.params = torch.empty(P, dtype=torch.float32, device=dev)
                self.grads = torch.zeros(P + 1, dtype=torch.float32, device=dev)
            self.off = []
            off = 0
            for p in tensors:
                n = p.numel()
                self.params[off:off + n].copy_(p.data.reshape(-1))
                p.data = self.params[off:off + n].view(p.shape)
                p.grad = self.grads[off:off + n].view(p.shape)
                owners[id(p)] = self
                self.off.append(off)
                off += n
            self.state = torch.zeros(4, dtype=torch.int64, device=dev)
            ld = self.ld = CAP_B
            self.act = [torch.empty(CAP_B, H, device=dev) for _ in range(4)]
            self.act_t = [torch.zeros(H + 1, ld, device=dev) for _ in range(4)]
            for t in self.act_t:
                t[H].fill_(1.0)                               # the operand row that turns wgrad's last column into db
            # one pre-activation-gradient buffer (+ transpose) per layer 3..0: wgrad of a layer runs on a
            # side stream while dgrad produces the next layer's buffer, so nothing is reused in flight
            self.dz = [torch.empty(CAP_B, H, device=dev) for _ in range(4)]
            self.dz_t = [torch.zeros(H, ld, device=dev) for _ in range(4)]
            self.side_stream = torch.cuda.Stream(device=dev)
            self.opt_stream = torch.cuda.Stream(device=dev)
            self.out = torch.empty(CAP_B, device=dev)
            self.dz4 = torch.empty(CAP_B, device=dev)
            self.sq = torch.empty(CAP_B, device=dev)
        # one rank: "end" = one pass over all parameters after backward (default); "layer" = each hidden layer's
        # Adam pass right after its wgrad, on its own stream beside the rest of backward, with a capped grid.
        # Measured on B200 (profiles/r02_wide_adam_modes.txt): "layer" is SLOWER at every width (0.563 vs
        # 0.548 ms at H = 4096 with 4 CTAs per SM, 0.73 ms with 1) -- the GEMMs are latency-bound and lose
        # more to the contention than the L2-hot gradient saves -- so it stays an experiment knob.  "fused" = Adam
        # inside each hidden layer's wgrad kernel (ddpb_linear_wgrad_adam: 28 instead of 32 B/param, bit-identical
        # results); its epilogue keeps too few bytes in flight (0.572 vs 0.544 ms at H = 4096), also a knob.
        self.adam_mode = os.environ.get("DDPB_WIDE_ADAM", "end")
        self.adam_ctas = int(os.environ.get("DDPB_WIDE_ADAM_CTAS", "4"))
        self.loss_dev = torch.zeros(LOSS_RING, dtype=torch.float32, device=dev)
        self.loss_host = torch.zeros(LOSS_RING, dtype=torch.float32, pin_memory=True)
        self.loss_bits = self.loss_host.numpy().view(np.uint32)
        self._loss_out = ctypes.c_float()
        self._sync_replicas()
        # torch's default of 25 MiB costs 15 % here (each bucket pays ~20 us of launch + NCCL latency,
        # DESIGN.md section 7): when the caller names no cap, one bucket per layer
        cap = os.environ.get("DDPB_BUCKET_CAP_MB", bucket_cap_mb if bucket_cap_mb else None)
        if cap is None:
            # no cap named: one bucket per HIDDEN layer, the thin ends riding with their neighbours (layer 4 with
            # layer 3, layer 0 with layer 1).  Every bucket costs ~50 us of launch skew + flag handshakes at 8
            # ranks whatever its size (profiles/r02_bucket_diag_8gpu.txt), so two 16 KB buckets of their own were
            # 10 % of the step.  The tag is the layer that becomes ready LAST.
            o = self.off
            self.buckets = [(3, o[6], self.P - o[6]), (2, o[4], o[6] - o[4]), (0, 0, o[4])]
        else:
            self.buckets = self._plan_buckets(float(cap))
        self.comm_stream = torch.cuda.Stream(device=dev) if self.world > 1 else None
        # DDPB_COMM_STREAMS=2: the fused bucket kernels alternate between two communication streams, so one
        # bucket's handshakes and tail run under the next bucket's transfers
        self.comm_streams = [self.comm_stream]
        if self.world > 1 and int(os.environ.get("DDPB_COMM_STREAMS", "1")) > 1:
            self.comm_streams.append(torch.cuda.Stream(device=dev))
        self.opt_kind = self.hyper = self.m = self.v = self.bias_table = None
        self.seq = 0
        self.waiting = None
        self.grads_fresh = False
        self.ticket = 0
        if self.peer is not None and len(self.buckets) > _lib.BUCKET_MAX:
            raise _lib.DdpB200Error(f"{len(self.buckets)} gradient buckets > {_lib.BUCKET_MAX}: raise bucket_cap_mb "
                                    "or set DDPB_ALLREDUCE=nccl")
        self.fused_exchange = self.peer is not None

    # -- layout ---------------------------------------------------------------------------
    def _w(self, l):          # data_ptr of W_l / b_l in params, of their gradients in grads
        return self.params.data_ptr() + 4 * self.off[2 * l], self.params.data_ptr() + 4 * self.off[2 * l + 1]

    def _g(self, l):
        return self.grads.data_ptr() + 4 * self.off[2 * l], self.grads.data_ptr() + 4 * self.off[2 * l + 1]

    def _plan_buckets(self, cap_mb):
        return plan_buckets(self.off, self.P, cap_mb)

    def _map_peer_windows(self):
        """Several ranks on one NVSwitch box (DDPB_ALLREDUCE=p2p, the default): the flat gradient and
        parameter buffers live in one IPC-shareable window per rank, [flags | gradients | parameters],
        mapped into every peer, so that a bucket's all-reduce and its Adam pass are ONE kernel
        (ddpb_bucket_reduce_adam).  Returns False (nothing allocated) when that is not possible; all
        ranks decide together."""
        if (self.world == 1 or self.backend_name != "nccl" or self.world > _lib.MAX_PEERS
                or os.environ.get("DDPB_ALLREDUCE", "p2p") != "p2p"):
            return False
        import socket
        lib, P = self.lib, self.P
        up = lambda b: (b + 255) // 256 * 256
        grads_off = up(_lib.BUCKET_FLAGS_BYTES)
        params_off = grads_off + up(4 * (P + 4))
        total = params_off + up(4 * P)
        mine, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
        _lib.check(lib.ddpb_peer_alloc(total, ctypes.byref(mine), handle), "ddpb_peer_alloc")
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
        if not all(agree):
            for r, ptr in enumerate(mapped):
                if r != self.rank:
                    lib.ddpb_peer_close(ptr)
            lib.ddpb_peer_free(mine)
            return False

        class _Span:                               # a device range as __cuda_array_interface__ (the window outlives it)
            def __init__(self, ptr, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}
        self.grads = torch.as_tensor(_Span(mine.value + grads_off, P + 1), device=self.device)
        self.params = torch.as_tensor(_Span(mine.value + params_off, P), device=self.device)
        if self.grads.data_ptr() != mine.value + grads_off or self.params.data_ptr() != mine.value + params_off:
            raise _lib.DdpB200Error("could not wrap the peer window as tensors without a copy")
        self.peer = dict(mine=mine, mapped=mapped, grads_off=grads_off, params_off=params_off)
        return True

    def _bucket_job(self, index, start, n):
        """The ddpb_bucket_job of one bucket (built once; ticket / mode / advance change per call)."""
        h, j = self.hyper, _lib.BucketJob()
        for r, ptr in enumerate(self.peer["mapped"]):
            j.window[r] = ptr
        j.grads_off, j.params_off = self.peer["grads_off"], self.peer["params_off"]
        j.m, j.v, j.bias_table, j.state = self.m.data_ptr(), self.v.data_ptr(), self.bias_table.data_ptr(), self.state.data_ptr()
        j.start, j.n, j.bias_table_n = start, n, self.bias_table.shape[0]
        j.world, j.rank, j.bucket_index = self.world, self.rank, index
        j.beta1, j.beta2, j.eps, j.grad_scale = h["betas"][0], h["betas"][1], h["eps"], 1.0 / self.world
        return j

    def _fused_update(self, index, stream, reduce, advance):
        """One kernel: all-reduce of bucket `index` over the peer windows + Adam on the owned shard +
        push of the new parameters to every rank."""
        j = self.bucket_jobs[index]
        j.ticket, j.mode, j.advance = self.ticket, int(reduce), int(advance)
        _lib.check(self.lib.ddpb_bucket_reduce_adam(ctypes.byref(j), stream), "ddpb_bucket_reduce_adam")

    def _sync_replicas(self):
        if self.world == 1:
            return
        if self.backend_name == "nccl":
            comm = self.engine.ensure_comm(self.process_group)
            _lib.check(self.lib.ddpb_broadcast_f32(comm, self.params.data_ptr(), self.P, 0, self.engine.stream()),
                       "ddpb_broadcast_f32")
        else:
            host = self.params.cpu()
            dist.broadcast(host, src=0, group=self.process_group)
            self.params.copy_(host)

    # -- optimizer ------------------------------------------------------------------------
    def attach_optimizer(self, kind, hyper):
        self.opt_kind, self.hyper = kind, hyper
        with torch.cuda.device(self.device):
            self.m = torch.zeros(self.P, dtype=torch.float32, device=self.device)
            self.v = torch.zeros(self.P, dtype=torch.float32, device=self.device) if kind == _lib.OPT_ADAM else self.m
            if kind == _lib.OPT_ADAM:
                table = _lib.adam_bias_table(hyper["lr"], *hyper["betas"])
                self.bias_table = torch.from_numpy(table).to(self.device)
        self.sgd_steps = 0
        self.fused_exchange = self.peer is not None and kind == _lib.OPT_ADAM     # SGD: ncclAllReduce + fused avg+SGD
        if self.fused_exchange:
            self.bucket_jobs = [self._bucket_job(i, start, n) for i, (_, start, n) in enumerate(self.buckets)]

    def optimizer_state_dict(self):
        m, v = self.m, self.v
        self.check_peers()
        if self.fused_exchange:
            # each rank holds Adam state for the shards it owns and zeros elsewhere: the sum is the full state
            m, v = m.clone(), v.clone()
            dist.all_reduce(m, group=self.process_group)
            dist.all_reduce(v, group=self.process_group)
        return {"step": int(self.state[0].item()), "exp_avg": m, "exp_avg_sq": v}

    def discard_pending_backward(self):
        self.waiting = None
        self.grads_fresh = False

    def _update(self, start, n, stream, advance, ctas_per_sm=8):
        """Fused average + optimizer over grads[start:start+n]."""
        h, scale = self.hyper, 1.0 / self.world
        g, p, m = self.grads.data_ptr() + 4 * start, self.params.data_ptr() + 4 * start, self.m.data_ptr() + 4 * start
        if self.opt_kind == _lib.OPT_ADAM:
            _lib.check(self.lib.ddpb_fused_avg_adam_dev_grid(g, p, m, self.v.data_ptr() + 4 * start, n, scale, h["betas"][0],
                                                             h["betas"][1], h["eps"], self.bias_table.data_ptr(),
                                                             self.bias_table.shape[0], self.state.data_ptr(), int(advance),
                                                             ctas_per_sm, stream), "ddpb_fused_avg_adam_dev")
        else:
            _lib.check(self.lib.ddpb_fused_avg_sgd(g, p, m, n, scale, h["lr"], h["momentum"], self.sgd_steps, stream),
                       "ddpb_fused_avg_sgd")

    # -- the step ---------------------------------------------------------------------------
    def _bind(self, data, target):
        if type(data) is HostBatch:
            data = data.cuda(self.device)
        if type(target) is HostBatch:
            target = target.cuda(self.device)
        if type(data) is StagedBatch:
            stage = data.stage
            stage.copy_now()
            return stage.dev_x_ptr, stage.dev_y_ptr, None, stage.B, stage
        if type(data) is DeviceBatch:
            return data.x.data_ptr(), data.y.data_ptr(), data.idx_ptr, data.B, data.keep
        if type(target) is StagedBatch:
            target = target.tensor()
        if not (isinstance(data, torch.Tensor) and data.is_cuda and data.dtype == torch.float32 and data.dim() == 2
                and data.shape[1] == 2 and data.is_contiguous()):
            raise _lib.DdpB200Error("model input must be a contiguous float32 CUDA tensor [B,2] (or a loader batch)")
        if target is not None and not (isinstance(target, torch.Tensor) and target.is_cuda and target.dtype == torch.float32
                                       and target.numel() == data.shape[0] and target.is_contiguous()):
            raise _lib.DdpB200Error("target must be a contiguous float32 CUDA tensor [B,1]")
        return data.data_ptr(), (target.data_ptr() if target is not None else None), None, data.shape[0], (data, target)

    def _run(self, loss, with_update):
        x, y, idx, B, keep = self._bind(loss.output.data, loss.target)
        if B > CAP_B:
            raise NotImplementedError(f"local batch {B} > {CAP_B} on the widened path")
        lib, H, ld, eng = self.lib, self.H, self.ld, self.engine
        stream = eng.stream()
        slot = loss.slot = self.seq % LOSS_RING
        self.seq += 1
        if self.fused_exchange and self.seq % 256 == 0:
            self.check_peers()                  # a timed-out bucket exchange must not go unnoticed for long
        self.loss_bits[slot] = LOSS_SENTINEL
        loss_slot = self.loss_host.data_ptr() + 4 * slot
        act, act_t, dz, dz_t = self.act, self.act_t, self.dz, self.dz_t
        ck = _lib.check
        # ---- forward
        w, b = self._w(0)
        ck(lib.ddpb_wide_first_fwd(x, idx, w, b, act[0].data_ptr(), act_t[0].data_ptr(), B, H, ld, stream), "first_fwd")
        for l in (1, 2, 3):
            w, b = self._w(l)
            ck(lib.ddpb_linear_fwd(act[l - 1].data_ptr(), w, b, act[l].data_ptr(), act_t[l].data_ptr(), B, H, H, ld, 1,
                                   stream), "linear_fwd")
        w, b = self._w(4)
        ck(lib.ddpb_wide_last_fwd(act[3].data_ptr(), w, b, y, idx, self.out.data_ptr(), self.dz4.data_ptr(),
                                  self.sq.data_ptr(), B, H, stream), "last_fwd")
        # ---- backward; a layer's bucket(s) go to the communication stream as soon as they are produced
        gw, gb = self._g(4)
        ck(lib.ddpb_wide_last_bwd(act[3].data_ptr(), w, self.dz4.data_ptr(), self.sq.data_ptr(), dz[0].data_ptr(),
                                  dz_t[0].data_ptr(), gw, gb, self.grads.data_ptr() + 4 * self.P, loss_slot, B, H, ld,
                                  stream), "last_bwd")
        self._layer_done(4, with_update, None)
        main, side = torch.cuda.current_stream(self.device), self.side_stream
        per_layer = with_update and self.world == 1 and self.opt_kind == _lib.OPT_ADAM and self.adam_mode == "layer"
        cur = 0
        fused_adam = with_update and self.world == 1 and self.opt_kind == _lib.OPT_ADAM and self.adam_mode == "fused"
        for l in (3, 2, 1):
            w, bias = self._w(l)
            gw, gb = self._g(l)
            if fused_adam:
                # one rank: Adam rides in the wgrad epilogue (28 instead of 32 B/param, no optimizer pass for the
                # layer).  It overwrites W_l, so it follows dgrad(l), and runs beside dgrad(l-1) on the side stream.
                ck(lib.ddpb_linear_dgrad(dz[cur].data_ptr(), w, act[l - 1].data_ptr(), dz[cur + 1].data_ptr(),
                                         dz_t[cur + 1].data_ptr(), B, H, H, ld, stream), "dgrad")
                side.wait_stream(main)
                h, ow, ob = self.hyper, 4 * self.off[2 * l], 4 * self.off[2 * l + 1]
                ck(lib.ddpb_linear_wgrad_adam(dz_t[cur].data_ptr(), act_t[l - 1].data_ptr(), gw, gb, w, bias,
                                              self.m.data_ptr() + ow, self.m.data_ptr() + ob, self.v.data_ptr() + ow,
                                              self.v.data_ptr() + ob, B, H, H, ld, h["betas"][0], h["betas"][1], h["eps"],
                                              self.bias_table.data_ptr(), self.bias_table.shape[0], self.state.data_ptr(),
                                              side.cuda_stream), "wgrad_adam")
                cur += 1
                continue
            side.wait_stream(main)                                # dz_t[cur] is complete on the main stream
            ck(lib.ddpb_linear_wgrad(dz_t[cur].data_ptr(), act_t[l - 1].data_ptr(), gw, gb, B, H, H, ld, side.cuda_stream), "wgrad")
            ck(lib.ddpb_linear_dgrad(dz[cur].data_ptr(), w, act[l - 1].data_ptr(), dz[cur + 1].data_ptr(),
                                     dz_t[cur + 1].data_ptr(), B, H, H, ld, stream), "dgrad")
            # the bucket's update overwrites W_l: it must follow wgrad (side) AND dgrad, which reads W_l (main)
            self._layer_done(l, with_update, side)
            if per_layer:
                # one rank: Adam on this layer NOW, on its own stream, while its gradient is still in L2 (24
                # instead of 32 B/param of HBM traffic) and beside the next layer's GEMMs, which leave most
                # of the HBM bandwidth unused; a capped grid leaves thread slots for their CTAs
                opt = self.opt_stream
                opt.wait_stream(side)
                opt.wait_stream(main)
                self._update(self.off[2 * l], self.off[2 * l + 2] - self.off[2 * l], opt.cuda_stream, False, self.adam_ctas)
            cur += 1
        gw, gb = self._g(0)
        ck(lib.ddpb_wide_first_bwd(dz[cur].data_ptr(), x, idx, gw, gb, B, H, stream), "first_bwd")
        self._layer_done(0, with_update, None)
        main.wait_stream(side)
        self.keep = keep
        if self.world > 1:
            for comm in self.comm_streams:
                main.wait_stream(comm)                          # parameters are final after this
            if with_update and self.fused_exchange and len(self.comm_streams) > 1:
                self._update(0, 0, stream, True)                # several communication streams: advance the counter here
        elif with_update:
            self.sgd_steps += 1
            if per_layer or fused_adam:                       # the thin ends (layer 0 and layer 4), then the counter
                main.wait_stream(self.opt_stream)
                self._update(0, self.off[2], stream, False)
                self._update(self.off[8], self.P - self.off[8], stream, True)
            else:
                self._update(0, self.P, stream, True)
        self.grads_fresh = not with_update

    def _layer_done(self, layer, with_update, produced_on):
        """Several ranks: all-reduce (and update) the buckets of `layer` on the communication stream,
        after the stream that produced them (`produced_on`, default the main stream)."""
        if self.world == 1:
            return
        if layer == 4:
            self.sgd_steps += int(with_update)
            self.ticket += 1
        for index, (bucket_layer, start, n) in enumerate(self.buckets):
            if bucket_layer != layer:
                continue
            fused = self.fused_exchange and with_update
            comm = self.comm_streams[index % len(self.comm_streams)] if fused else self.comm_stream
            comm.wait_stream(torch.cuda.current_stream(self.device))
            if produced_on is not None:
                comm.wait_stream(produced_on)
            if layer < 4:
                comm.wait_stream(self.side_stream)       # a merged bucket also holds an earlier layer's wgrad output
            cs = comm.cuda_stream
            last = index == len(self.buckets) - 1
            if fused:
                # the step counter is advanced once, after every bucket of the step (see _run)
                self._fused_update(index, cs, True, last and len(self.comm_streams) == 1)
                continue
            ptr = self.grads.data_ptr() + 4 * start
            if self.backend_name == "nccl":
                _lib.check(self.lib.ddpb_allreduce_sum_f32(self.engine.comm, ptr, n, cs), "ddpb_allreduce_sum_f32")
            else:
                with torch.cuda.stream(self.comm_stream):
                    host = self.grads[start:start + n].cpu()
                    dist.all_reduce(host, group=self.process_group)
                    self.grads[start:start + n].copy_(host)
            if with_update:
                self._update(start, n, cs, last)

    def request_backward(self, loss):
        if (self.waiting is not None and self.waiting is not loss) or self.grads_fresh:
            from .ddp import _SECOND_BACKWARD
            raise NotImplementedError(_SECOND_BACKWARD)
        if self.opt_kind is not None:
            self.waiting = loss
        else:
            self.waiting = None
            self._run(loss, False)

    def request_step(self):
        if self.opt_kind is None:
            raise _lib.DdpB200Error("no optimizer attached")
        if self.waiting is not None:
            loss, self.waiting = self.waiting, None
            self._run(loss, True)
            return
        if not self.grads_fresh:
            return
        self.grads_fresh = False
        self.sgd_steps += 1
        if self.fused_exchange:
            # the buckets were summed in place by ncclAllReduce during backward: owned shards are
            # updated and pushed (mode 0), so the sharded Adam state stays the only one
            self.ticket += 1
            for index in range(len(self.buckets)):
                self._fused_update(index, self.engine.stream(), False, index == len(self.buckets) - 1)
            return
        self._update(0, self.P, self.engine.stream(), True)      # the 1/W scale is applied inside the pass

    # -- results ------------------------------------------------------------------------------
    def read_loss(self, loss):
        if loss.slot is None:
            if self.waiting is loss:
                self.waiting = None
            self._run(loss, False)
        _lib.check(self.lib.ddpb_wait_f32(self.loss_host.data_ptr() + 4 * loss.slot, LOSS_SENTINEL, 2_000_000,
                                          self.engine.stream(), ctypes.byref(self._loss_out)), "ddpb_wait_f32")
        return torch.tensor(self._loss_out.value)

    def loss_device_view(self, loss):
        return self.read_loss(loss).to(self.device)

    def forward_now(self, data):
        x, _, idx, B, keep = self._bind(data, None)
        lib, H, ld, stream = self.lib, self.H, self.ld, self.engine.stream()
        w, b = self._w(0)
        _lib.check(lib.ddpb_wide_first_fwd(x, idx, w, b, self.act[0].data_ptr(), self.act_t[0].data_ptr(), B, H, ld, stream))
        for l in (1, 2, 3):
            w, b = self._w(l)
            _lib.check(lib.ddpb_linear_fwd(self.act[l - 1].data_ptr(), w, b, self.act[l].data_ptr(), None, B, H, H, ld, 1, stream))
        w, b = self._w(4)
        _lib.check(lib.ddpb_wide_last_fwd(self.act[3].data_ptr(), w, b, None, None, self.out.data_ptr(), None, None, B, H, stream))
        return self.out[:B].clone().view(B, 1)

    def close(self):
        """Collective: unmap the peers' windows and free this rank's (every rank calls it, after which
        the module's parameters and gradients must not be used).  Without it the windows live until
        the process exits."""
        if self.peer is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.process_group)               # nobody is still inside a bucket kernel
        for r, ptr in enumerate(self.peer["mapped"]):
            if r != self.rank:
                self.lib.ddpb_peer_close(ptr)
        dist.barrier(group=self.process_group)               # every mapping of my window is closed
        self.lib.ddpb_peer_free(self.peer["mine"])
        self.peer, self.fused_exchange = None, False
        self.params = self.grads = None

    def check_peers(self):
        """Raise if a fused bucket exchange timed out waiting for a peer (state.reserved[0] set by the kernel)."""
        if self.fused_exchange and int(self.state[2].item()) != 0:
            raise _lib.DdpB200Error("a peer GPU did not arrive at a gradient bucket in time (fused all-reduce timed out)")
