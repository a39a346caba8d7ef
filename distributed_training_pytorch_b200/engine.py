"""Per-device launch engine: the lazy step queue, the NCCL communicator, launch counters.

The reference issues ~140 library kernels per iteration from Python (demo.py:104-111).
Here the front-end objects (wrapper, loss, optimizer) only *describe* work; the engine
launches it when a result is needed (loss.cpu(), a second step of the same model, an
explicit flush).  All models of the process that are pending at that moment go out in ONE
ddpb_toy_step launch (blockIdx.y = model), so demo.py's two models cost one launch per
iteration on a single rank.
"""
import ctypes
import weakref

import torch

from . import _lib

_ENGINES = {}


def engine_for(device_index):
    eng = _ENGINES.get(device_index)
    if eng is None:
        eng = _ENGINES[device_index] = Engine(device_index)
    return eng


def all_engines():
    return list(_ENGINES.values())


class Engine:
    def __init__(self, device_index):
        if not torch.cuda.is_available():
            raise _lib.DdpB200Error("no CUDA device: distributed_training_pytorch_b200 has no CPU fallback")
        self.device = device_index
        self.lib = _lib.lib()
        major, minor, sms = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.ddpb_device_info(device_index, ctypes.byref(major), ctypes.byref(minor), ctypes.byref(sms)),
                   "ddpb_device_info")
        if major.value != 10:
            raise _lib.DdpB200Error(f"device {device_index} is sm_{major.value}{minor.value}; this library is built "
                                    "for sm_100a (B200) only")
        self.sm_count = sms.value
        self.queue = []                                    # backends with a described, unlaunched job
        # Job slots: every toy model of the device owns one element of this array and describes its
        # work in place, so a launch of models in slot order passes the array as it is (no copies).
        self.jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
        self._jobs_addr = ctypes.addressof(self.jobs)
        self._job_size = ctypes.sizeof(_lib.ToyJob)
        self._free_slots = list(range(_lib.MAX_JOBS))
        self._scratch = (_lib.ToyJob * _lib.MAX_JOBS)()    # launches that are not a run of consecutive slots
        self._scratch_addr = ctypes.addressof(self._scratch)
        # the descriptor of ddpb_toy_iteration: gather + launch + loss wait in ONE host call
        self.it = it = _lib.ToyIter()
        self._it_addr = ctypes.addressof(it)
        it.sentinel_bits, it.spin_ns = _lib.LOSS_SENTINEL, 200_000
        self._iteration = self.lib.ddpb_toy_iteration
        self._it_jobs = self._it_n = self._it_loader = self._it_B = None    # what the descriptor currently holds
        self._it_gather, self._it_nwait = False, 0
        self.pending_stage = None                          # a data._Stage whose gather rides with the next launch
        self.last_stream = None                            # stream of the most recent launch
        self._stream = None
        self._torch_stream = None
        self.launch_no = 0                                 # toy launches issued / known to have run on the GPU
        self.completed = 0
        self.comm = None                                   # NCCL communicator handle (void*)
        self.comm_rank, self.comm_world = 0, 1

    @property
    def launches(self):
        """Kernels the library has launched in this process (counted at the launch sites, csrc/util.cu)."""
        return int(self.lib.ddpb_launch_count())

    def job_slot(self, owner):
        """A ddpb_toy_job for a new model: an element of the engine's array while there are free ones
        (given back when `owner` is garbage-collected), else a private struct that is packed at launch."""
        free = self._free_slots
        if not free:
            return -1, _lib.ToyJob()
        k = free.pop(0)
        ctypes.memset(self._jobs_addr + k * self._job_size, 0, self._job_size)
        weakref.finalize(owner, free.append, k)
        return k, self.jobs[k]

    # -- streams -------------------------------------------------------------
    def stream(self):
        """cudaStream_t the engine launches on: torch's current stream of the device when the
        engine was created (normally the default stream), cached because querying torch costs
        more than describing a whole step.  Call use_stream() to move the engine elsewhere."""
        s = self._stream
        if s is None:
            s = self._stream = torch.cuda.current_stream(self.device).cuda_stream
            self.it.stream = s
        return s

    def torch_stream(self):
        """The same stream as a torch object (for event records)."""
        ts = self._torch_stream
        if ts is None:
            h = self.stream()
            ts = self._torch_stream = (torch.cuda.ExternalStream(h, device=self.device) if h
                                       else torch.cuda.default_stream(self.device))
        return ts

    def use_stream(self, stream=None):
        """Launch on `stream` (a torch.cuda.Stream) from now on; None = re-read torch's current stream."""
        self.flush()
        self._stream = None if stream is None else stream.cuda_stream
        self._torch_stream = None
        if self._stream is not None:
            self.it.stream = self._stream

    def sync(self):
        _lib.check(self.lib.ddpb_stream_sync(self.stream()), "ddpb_stream_sync")
        self.completed = self.launch_no

    # -- lazy queue ------------------------------------------------------------
    def enqueue(self, model):
        if model in self.queue:
            self.flush()
        self.queue.append(model)
        if len(self.queue) == _lib.MAX_JOBS:
            self.flush()

    def attach_stage(self, stage):
        """The next launch reads `stage` (a pinned (x|y) block) in place; its rows are gathered by the
        same native call."""
        prev = self.pending_stage
        if prev is not stage:
            if prev is not None:
                prev.materialize()
                prev.guard = (self, self.launch_no + 1)
            self.pending_stage = stage

    def flush(self, wait=False):
        """Launch everything described so far (one kernel for all pending toy jobs).  wait=True also
        polls, inside the same native call, the pinned loss slots of the launched steps and leaves the
        values on their Loss objects."""
        batch = self.queue
        if not batch:
            return
        self.queue = []
        n = len(batch)
        it = self.it
        s0 = batch[0].slot
        addr = self._jobs_addr + s0 * self._job_size if s0 >= 0 else 0
        for i in range(1, n):
            if batch[i].slot != s0 + i:
                addr = 0
                break
        if not addr:                                       # not a run of consecutive slots: pack them
            for i, model in enumerate(batch):
                self._scratch[i] = model.job
            addr = self._scratch_addr
        if addr != self._it_jobs or n != self._it_n:       # ctypes field stores cost ~0.1 us each: only what changed
            it.jobs, it.njobs = addr, n
            self._it_jobs, self._it_n = addr, n
        stream = self.last_stream = self._stream or self.stream()
        self.launch_no = launch_no = self.launch_no + 1
        stage = self.pending_stage
        if stage is not None:
            self.pending_stage = None
            stage.guard = (self, launch_no)
            p = stage.pending_idx
            if p is not None:
                stage.pending_idx = None
                ld = stage.loader
                if ld is not self._it_loader:
                    it.X, it.Y, it.dx, it.dy = ld.x_ptr, ld.y_ptr, stage.dx, stage.dy
                    self._it_loader = ld
                if stage.B != self._it_B:
                    it.B = self._it_B = stage.B
                it.idx, it.out_x, it.out_y = p[1], stage.host_x_ptr, stage.host_y_ptr
                self._it_gather = True
            elif self._it_gather:
                it.idx, self._it_gather = None, False
        elif self._it_gather:
            it.idx, self._it_gather = None, False
        nwait = 0
        for model in batch:
            loss = model.last_loss
            if loss is not None:
                loss.launch_no = launch_no
                if wait and loss.on_host and loss.value is None:
                    it.wait_slot[nwait] = model.loss_ptr + 4 * loss.slot
                    nwait += 1
                    if model.gloss_wait:                 # fused exchange: the all-rank mean lands right after it
                        it.wait_slot[nwait] = model.gloss_ptr + 4 * loss.slot
                        nwait += 1
        if nwait != self._it_nwait:
            it.nwait = self._it_nwait = nwait
        rc = self._iteration(self._it_addr)
        if rc:
            _lib.check(rc, "ddpb_toy_iteration")
        if nwait:
            k = 0
            for model in batch:
                loss = model.last_loss
                if loss is not None and loss.on_host and loss.value is None:
                    loss.value = it.wait_value[k]
                    k += 1
                    if model.gloss_wait:
                        loss.gvalue = it.wait_value[k]
                        k += 1
            self.completed = launch_no
        for model in batch:
            model.after_launch(stream)

    # -- communicator ----------------------------------------------------------
    def ensure_comm(self, process_group=None):
        """Create the library-owned NCCL communicator: rank 0 draws the id, the c10d
        process group (whatever its backend) carries it to the other ranks."""
        import torch.distributed as dist
        if self.comm is not None or not dist.is_initialized():
            return self.comm
        world, rank = dist.get_world_size(process_group), dist.get_rank(process_group)
        self.comm_rank, self.comm_world = rank, world
        if world == 1:
            return None
        import os
        nccl_path = os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "nccl", "lib", "libnccl.so.2")
        _lib.check(self.lib.ddpb_nccl_load(os.path.abspath(nccl_path).encode() if os.path.exists(nccl_path) else None),
                   "ddpb_nccl_load")
        uid = ctypes.create_string_buffer(128)
        if rank == 0:
            _lib.check(self.lib.ddpb_get_unique_id(uid), "ddpb_get_unique_id")
        box = [uid.raw]
        dist.broadcast_object_list(box, src=0, group=process_group)
        handle = ctypes.c_void_p()
        _lib.check(self.lib.ddpb_comm_create(box[0], rank, world, self.device, ctypes.byref(handle)), "ddpb_comm_create")
        self.comm = handle
        return self.comm

    def destroy_comm(self):
        if self.comm is not None:
            _lib.check(self.lib.ddpb_comm_destroy(self.comm), "ddpb_comm_destroy")
            self.comm = None
