"""optim.Adam / optim.SGD with the call shapes of demo.py:80-81,100-101,110-111.

The optimizer owns no kernels of its own on the single-rank toy path: `step()` marks the
pending backward of its model as "update too", and the fused kernel applies torch's Adam
(op order of torch:optim/adam.py:457-547) in the same launch.  On the multi-rank path it
runs the fused average+Adam pass over the all-reduced bucket (csrc/optim.cu)."""
from . import _lib
from .ddp import owner_of


class _FlatOptimizer:
    kind = None

    def __init__(self, params, hyper):
        params = list(params)
        if not params:
            raise ValueError("optimizer got an empty parameter list")
        if isinstance(params[0], dict):
            raise NotImplementedError("parameter groups are not part of the reference path")
        backends = {id(owner_of(p)): owner_of(p) for p in params}
        if len(backends) != 1:
            raise NotImplementedError("one optimizer per wrapped model (demo.py:80-81); got parameters of "
                                      f"{len(backends)} models")
        self.backend = next(iter(backends.values()))
        if len(params) != self.backend.num_tensors:
            raise NotImplementedError("the optimizer must own all parameters of its model")
        self.defaults = dict(hyper)
        self.param_groups = [dict(params=params, **hyper)]
        self.backend.attach_optimizer(self.kind, hyper)

    def zero_grad(self, set_to_none=True):
        # The gradient bucket is persistent and fully overwritten by the next backward
        # (no accumulation across iterations in the reference loop), so this is bookkeeping only.
        self.backend.discard_pending_backward()

    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("closures are not part of the reference path")
        self.backend.request_step()

    def state_dict(self):
        return self.backend.optimizer_state_dict()


class Adam(_FlatOptimizer):
    kind = _lib.OPT_ADAM

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("weight_decay / amsgrad are not used by the reference (demo.py:80-81)")
        super().__init__(params, dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), eps=float(eps)))


class SGD(_FlatOptimizer):
    kind = _lib.OPT_SGD

    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0, weight_decay=0, nesterov=False):
        if dampening != 0 or weight_decay != 0 or nesterov:
            raise NotImplementedError("only plain / momentum SGD is implemented")
        super().__init__(params, dict(lr=float(lr), momentum=float(momentum)))
