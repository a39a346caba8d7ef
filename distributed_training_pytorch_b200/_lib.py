"""ctypes binding of csrc/libddp_b200.so (declarations in include/ddp_b200.h).

There is no fallback: if the library is missing or a call fails, the caller gets an
exception.  Nothing in this package computes the training step on the CPU.
"""
import ctypes
import os

from .build import LIB_PATH

c_void_p, c_int, c_int32, c_int64, c_uint64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint64
c_float, c_double, c_char_p = ctypes.c_float, ctypes.c_double, ctypes.c_char_p

MODE_FWD, MODE_GRAD, MODE_STEP = 0, 1, 2
OPT_ADAM, OPT_SGD = 0, 1
MAX_JOBS = 4
MAX_PEERS = 8
XCHG_ROW = 376
XCHG_WINDOW_BYTES = 2 * MAX_PEERS * XCHG_ROW * 8
BUCKET_MAX = 1024
BUCKET_FLAGS_BYTES = 2 * BUCKET_MAX * MAX_PEERS * 4 + BUCKET_MAX * 4
TOY_HIDDEN, TOY_PARAMS = 10, 371
LOSS_SENTINEL = 0x7FC0DEAD          # a NaN payload no kernel produces: "slot not written yet"
LOSS_TIMEOUT_BITS = 0x7FC0DEAE      # DDPB_LOSS_TIMEOUT_BITS: the step gave up waiting for a peer


class DdpB200Error(RuntimeError):
    pass


class OptState(ctypes.Structure):
    _fields_ = [("step", ctypes.c_longlong), ("seq", ctypes.c_longlong), ("reserved", ctypes.c_longlong * 2)]


class ToyJob(ctypes.Structure):
    _fields_ = [("params", c_void_p), ("grads", c_void_p), ("m", c_void_p), ("v", c_void_p),
                ("state", c_void_p), ("bias_table", c_void_p), ("loss_ring", c_void_p), ("x", c_void_p), ("y", c_void_p),
                ("idx", c_void_p), ("out", c_void_p), ("global_loss_ring", c_void_p),
                ("B", c_int32), ("mode", c_int32), ("opt", c_int32), ("loss_ring_n", c_int32),
                ("idx_rows", c_int32), ("idx_stride", c_int32), ("bias_table_n", c_int32), ("n_data", c_int32),
                ("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("momentum", c_double),
                ("peer_data", c_void_p * 8), ("peer_flags", c_void_p * 8),
                ("world", c_int32), ("rank", c_int32), ("inv_world", c_float), ("reserved2", c_int32)]


class ToyIter(ctypes.Structure):          # ddpb_toy_iter: one host call per iteration (gather, launch, loss wait)
    _fields_ = [("jobs", c_void_p), ("stream", c_void_p), ("X", c_void_p), ("Y", c_void_p), ("idx", c_void_p),
                ("out_x", c_void_p), ("out_y", c_void_p), ("wait_slot", c_void_p * (2 * MAX_JOBS)),
                ("B", c_int64), ("spin_ns", c_int64),
                ("njobs", c_int32), ("dx", c_int32), ("dy", c_int32), ("nwait", c_int32),
                ("sentinel_bits", ctypes.c_uint32), ("wait_value", c_float * (2 * MAX_JOBS)), ("reserved", c_int32)]


class SegJob(ctypes.Structure):
    _fields_ = [("params", c_void_p), ("grads", c_void_p), ("m", c_void_p), ("v", c_void_p), ("state", c_void_p),
                ("bias_table", c_void_p), ("loss_ring", c_void_p), ("x", c_void_p), ("y", c_void_p),
                ("d_out", c_void_p), ("out", c_void_p), ("d_in", c_void_p),
                ("B", c_int32), ("mode", c_int32), ("opt", c_int32), ("nl", c_int32), ("dims", c_int32 * 4),
                ("act_last", c_int32), ("has_loss", c_int32), ("loss_ring_n", c_int32), ("bias_table_n", c_int32),
                ("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double), ("momentum", c_double)]


class BucketJob(ctypes.Structure):          # ddpb_bucket_job
    _fields_ = [("window", c_void_p * MAX_PEERS), ("grads_off", c_int64), ("params_off", c_int64),
                ("m", c_void_p), ("v", c_void_p), ("bias_table", c_void_p), ("state", c_void_p),
                ("start", c_int64), ("n", c_int64), ("ticket", c_int64),
                ("bias_table_n", c_int32), ("world", c_int32), ("rank", c_int32), ("bucket_index", c_int32),
                ("mode", c_int32), ("advance", c_int32),
                ("beta1", c_double), ("beta2", c_double), ("eps", c_double), ("grad_scale", c_float), ("reserved", c_int32)]


_SIGNATURES = {
    "ddpb_version": (c_int, []),
    "ddpb_last_error": (c_char_p, []),
    "ddpb_device_info": (c_int, [c_int, c_void_p, c_void_p, c_void_p]),
    "ddpb_launch_count": (ctypes.c_longlong, []),
    "ddpb_randperm": (c_int, [c_uint64, c_int64, c_void_p]),
    "ddpb_sampler_indices": (c_int64, [c_int64, c_int, c_int, c_int, c_uint64, c_uint64, c_int, c_void_p, c_int64]),
    "ddpb_sampler_epochs_i32": (c_int, [c_int64, c_int, c_int, c_int, c_uint64, c_uint64, c_int, c_int, c_void_p, c_int64]),
    "ddpb_sampler_epochs_i64": (c_int, [c_int64, c_int, c_int, c_int, c_uint64, c_uint64, c_int, c_int, c_void_p, c_int64]),
    "ddpb_host_gather": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "ddpb_gather_batch": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ddpb_toy_step": (c_int, [c_void_p, c_int, c_void_p]),
    "ddpb_toy_step_h2d": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int64, c_void_p]),
    "ddpb_toy_step_virtual_ranks": (c_int, [c_void_p, c_int, c_void_p]),
    "ddpb_toy_iteration": (c_int, [c_void_p]),
    "ddpb_linear_fwd": (c_int, [c_void_p] * 5 + [c_int] * 5 + [c_void_p]),
    "ddpb_linear_dgrad": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "ddpb_linear_wgrad": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_void_p]),
    "ddpb_linear_wgrad_adam": (c_int, [c_void_p] * 10 + [c_int] * 4 + [c_double] * 3 + [c_void_p, c_int, c_void_p, c_void_p]),
    "ddpb_wide_first_fwd": (c_int, [c_void_p] * 6 + [c_int] * 3 + [c_void_p]),
    "ddpb_wide_last_fwd": (c_int, [c_void_p] * 8 + [c_int] * 2 + [c_void_p]),
    "ddpb_wide_last_bwd": (c_int, [c_void_p] * 10 + [c_int] * 3 + [c_void_p]),
    "ddpb_wide_first_bwd": (c_int, [c_void_p] * 5 + [c_int] * 2 + [c_void_p]),
    "ddpb_segment_step": (c_int, [c_void_p, c_void_p]),
    "ddpb_wait_f32": (c_int, [c_void_p, ctypes.c_uint32, c_int64, c_void_p, c_void_p]),
    "ddpb_fused_avg_adam": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float,
                                    c_double, c_double, c_double, c_double, c_int64, c_void_p]),
    "ddpb_fused_avg_adam_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float,
                                        c_double, c_double, c_double, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "ddpb_fused_avg_adam_dev_grid": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float,
                                             c_double, c_double, c_double, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p]),
    "ddpb_fused_avg_sgd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_float, c_double, c_double, c_int64, c_void_p]),
    "ddpb_fused_avg_sgd_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_float, c_double, c_double, c_void_p, c_int,
                                       c_void_p]),
    "ddpb_bucket_reduce_adam": (c_int, [c_void_p, c_void_p]),
    "ddpb_nccl_load": (c_int, [c_char_p]),
    "ddpb_get_unique_id": (c_int, [c_void_p]),
    "ddpb_comm_create": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p]),
    "ddpb_comm_destroy": (c_int, [c_void_p]),
    "ddpb_broadcast_f32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "ddpb_allreduce_sum_f32": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ddpb_graph_begin": (c_int, [c_void_p]),
    "ddpb_graph_end": (c_int, [c_void_p, c_void_p]),
    "ddpb_graph_launch": (c_int, [c_void_p, c_void_p]),
    "ddpb_graph_destroy": (c_int, [c_void_p]),
    "ddpb_peer_alloc": (c_int, [c_int64, c_void_p, c_void_p]),
    "ddpb_peer_open": (c_int, [c_void_p, c_void_p]),
    "ddpb_peer_close": (c_int, [c_void_p]),
    "ddpb_peer_free": (c_int, [c_void_p]),
    "ddpb_enable_peer": (c_int, [c_int, c_int]),
    "ddpb_stream_sync": (c_int, [c_void_p]),
    "ddpb_memcpy_h2d_async": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ddpb_memcpy_d2h_async": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
}

_lib = None


def exported_symbols():
    """Names every build of the library must export (checked by the CPU test-suite)."""
    return sorted(_SIGNATURES)


def lib():
    """The loaded library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DdpB200Error(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "This package has no CPU or PyTorch fallback for the training step.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = restype, argtypes
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().ddpb_last_error().decode(errors="replace")
        raise DdpB200Error(f"{what or 'libddp_b200'} failed (code {rc}): {msg}")


def adam_bias_table(lr, beta1, beta2, max_steps=1 << 22):
    """[n,2] float32 rows {-(lr/(1-beta1^t)), sqrt(1-beta2^t)} for t = 1..n, computed with
    Python doubles exactly like torch:optim/adam.py:531-536, n = first step at which both
    corrections have converged to 1.0 (later steps reuse the last row)."""
    import numpy as np
    rows = []
    t = 0
    while True:
        t += 1
        bc1 = 1 - beta1 ** t
        bc2 = 1 - beta2 ** t
        rows.append((-(lr / bc1), bc2 ** 0.5))
        if bc1 == 1.0 and bc2 == 1.0:
            break
        if t >= max_steps:
            raise DdpB200Error(f"Adam bias corrections for betas=({beta1},{beta2}) do not converge within {max_steps} steps")
    return np.asarray(rows, dtype=np.float64).astype(np.float32)
