"""ctypes binding of oracle/ddp_oracle.c (TEST INFRASTRUCTURE ONLY)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libddp_oracle.so")
TOY_DIMS = (2, 10, 10, 10, 10, 1)      # toy_model_and_data.py:13-21


def build_oracle(force=False):
    src = os.path.join(HERE, "ddp_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-s", "-C", HERE, "-B"], check=True)
    return LIB


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


class Oracle:
    def __init__(self):
        self.lib = ctypes.CDLL(build_oracle())
        L = self.lib
        L.oracle_randperm.argtypes = [ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
        L.oracle_sampler_indices.restype = ctypes.c_int64
        L.oracle_sampler_indices.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                                             ctypes.c_void_p, ctypes.c_int64]
        L.oracle_param_count.restype = ctypes.c_int64
        L.oracle_mlp_loss_grad.restype = ctypes.c_float
        L.oracle_mse.restype = ctypes.c_float
        L.oracle_mse.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        L.oracle_adam_step.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int64] + [ctypes.c_double] * 4 + [ctypes.c_int64]
        L.oracle_sgd_step.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int64, ctypes.c_double, ctypes.c_double, ctypes.c_int64]
        L.oracle_ddp_average.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p]

    # -- sampler ----------------------------------------------------------
    def randperm(self, seed, n):
        out = np.empty(n, np.int64)
        if self.lib.oracle_randperm(seed, n, out.ctypes.data):
            raise ValueError("oracle_randperm: n out of the restated range")
        return out

    def sampler_indices(self, n, world, rank, shuffle=True, seed=0, epoch=0, drop_last=False):
        cap = (n + world - 1) // world + 1
        out = np.empty(cap, np.int64)
        k = self.lib.oracle_sampler_indices(n, world, rank, int(shuffle), seed, epoch, int(drop_last),
                                            out.ctypes.data, cap)
        if k < 0:
            raise ValueError("oracle_sampler_indices failed")
        return out[:k].copy()

    # -- model ------------------------------------------------------------
    def param_count(self, dims):
        d = np.asarray(dims, np.int32)
        return int(self.lib.oracle_param_count(_ip(d), len(dims) - 1))

    def forward(self, params, dims, x, act_last=False):
        d = np.asarray(dims, np.int32); x = _f32(x); params = _f32(params)
        B = x.shape[0]
        out = np.empty((B, dims[-1]), np.float32)
        self.lib.oracle_mlp_forward(_fp(params), _ip(d), len(dims) - 1, int(act_last), _fp(x), B, _fp(out))
        return out

    def mse(self, o, t):
        o = _f32(o); t = _f32(t)
        dO = np.empty_like(o)
        loss = self.lib.oracle_mse(o.ctypes.data, t.ctypes.data, o.size, dO.ctypes.data)
        return float(loss), dO

    def loss_grad(self, params, dims, x, y):
        d = np.asarray(dims, np.int32); x = _f32(x); y = _f32(y); params = _f32(params)
        g = np.empty_like(params)
        loss = self.lib.oracle_mlp_loss_grad(_fp(params), _ip(d), len(dims) - 1, _fp(x), _fp(y),
                                             x.shape[0], _fp(g))
        return float(loss), g

    def segment_backward(self, params, dims, x, d_out, act_last, want_din=True):
        d = np.asarray(dims, np.int32); x = _f32(x); params = _f32(params); d_out = _f32(d_out)
        g = np.empty_like(params)
        din = np.empty((x.shape[0], dims[0]), np.float32) if want_din else None
        self.lib.oracle_segment_backward(_fp(params), _ip(d), len(dims) - 1, int(act_last), _fp(x),
                                         x.shape[0], _fp(d_out), _fp(g),
                                         _fp(din) if want_din else None)
        return g, din

    # -- reducer / optimizer ----------------------------------------------
    def ddp_average(self, rank_grads):
        rg = _f32(rank_grads)
        out = np.empty(rg.shape[1], np.float32)
        self.lib.oracle_ddp_average(rg.ctypes.data, rg.shape[0], rg.shape[1], out.ctypes.data)
        return out

    def adam_step(self, p, g, m, v, step, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8):
        """In place on p, m, v (float32 contiguous)."""
        g = _f32(g)
        self.lib.oracle_adam_step(p.ctypes.data, g.ctypes.data, m.ctypes.data, v.ctypes.data, p.size,
                                  lr, beta1, beta2, eps, step)

    def sgd_step(self, p, g, buf, step, lr, momentum=0.0):
        g = _f32(g)
        self.lib.oracle_sgd_step(p.ctypes.data, g.ctypes.data, buf.ctypes.data, p.size, lr, momentum, step)

    # -- whole loop ---------------------------------------------------------
    def train_loop(self, params, dims, data_x, data_y, world, iters, batch=256, mode="distributed", lr=1e-3):
        """params: [nmodels, P] (copied).  Returns (final_params, losses[iters, nmodels])."""
        d = np.asarray(dims, np.int32)
        p = _f32(params).copy()
        x = _f32(data_x); y = _f32(data_y)
        losses = np.empty((iters, p.shape[0]), np.float32)
        rc = self.lib.oracle_train_loop(_fp(p), p.shape[0], _ip(d), len(dims) - 1, _fp(x), _fp(y),
                                        ctypes.c_int64(x.shape[0]), world, batch,
                                        0 if mode == "distributed" else 1, iters,
                                        ctypes.c_double(lr), _fp(losses))
        if rc:
            raise RuntimeError("oracle_train_loop failed")
        return p, losses
