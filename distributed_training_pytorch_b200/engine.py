"""Per-device launch engine: the lazy step queue, the NCCL communicator, launch counters.

The reference issues ~140 library kernels per iteration from Python (demo.py:104-111).
Here the front-end objects (wrapper, loss, optimizer) only *describe* work; the engine
launches it when a result is needed (loss.cpu(), a second step of the same model, an
explicit flush).  All models of the process that are pending at that moment go out in ONE
ddpb_toy_step launch (blockIdx.y = model), so demo.py's two models cost one launch per
iteration on a single rank.
"""
import ctypes

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
        self.queue = []                                    # FlatModel objects with a described, unlaunched job
        self.jobs = (_lib.ToyJob * _lib.MAX_JOBS)()
        self.launches = 0                                  # kernels this engine has launched (bench: gpu_launches)
        self.pending_copy = None                           # a data._Stage whose H2D copy rides with the next launch
        self.last_stream = None                            # stream of the most recent launch
        self._stream = None
        self.comm = None                                   # NCCL communicator handle (void*)
        self.comm_rank, self.comm_world = 0, 1

    # -- streams -------------------------------------------------------------
    def stream(self):
        """cudaStream_t the engine launches on: torch's current stream of the device when the
        engine was created (normally the default stream), cached because querying torch costs
        more than describing a whole step.  Call use_stream() to move the engine elsewhere."""
        s = self._stream
        if s is None:
            s = self._stream = torch.cuda.current_stream(self.device).cuda_stream
        return s

    def use_stream(self, stream=None):
        """Launch on `stream` (a torch.cuda.Stream) from now on; None = re-read torch's current stream."""
        self.flush()
        self._stream = None if stream is None else stream.cuda_stream

    def sync(self):
        _lib.check(self.lib.ddpb_stream_sync(self.stream()), "ddpb_stream_sync")

    # -- lazy queue ------------------------------------------------------------
    def enqueue(self, model):
        if model in self.queue:
            self.flush()
        self.queue.append(model)
        if len(self.queue) == _lib.MAX_JOBS:
            self.flush()

    def attach_copy(self, stage):
        """The next launch copies `stage` (pinned (x|y) block) to its device twin first."""
        if self.pending_copy is not None and self.pending_copy is not stage:
            self.pending_copy.copy_now()
        self.pending_copy = stage

    def drop_pending_copy(self, stage):
        if self.pending_copy is stage:
            self.pending_copy = None

    def flush(self):
        """Launch everything described so far (one kernel for all pending toy jobs)."""
        if not self.queue:
            return
        batch, self.queue = self.queue, []
        jobs = self.jobs
        for i, model in enumerate(batch):
            jobs[i] = model.job
        stream = self.last_stream = self.stream()
        stage = self.pending_copy
        if stage is not None and not stage.copied:
            rc = self.lib.ddpb_toy_step_h2d(jobs, len(batch), stage.dev_x_ptr, stage.host_ptr, stage.nbytes, stream)
            stage.copied = True
        else:
            rc = self.lib.ddpb_toy_step(jobs, len(batch), stream)
        self.pending_copy = None
        if rc:
            _lib.check(rc, "ddpb_toy_step")
        self.launches += 1
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
