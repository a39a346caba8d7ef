"""nn.MSELoss() with the call shape of demo.py:86,106-118: loss(out, target) returns an
object with .backward() and .detach().cpu().  Nothing is computed when it is built; the
mean-squared error comes out of the same fused kernel that produces the gradients."""
import torch

_Output = None          # ddp.Output, bound on first use (ddp imports this module's users)


class MSELoss:
    def __init__(self, reduction="mean"):
        if reduction != "mean":
            raise NotImplementedError("only reduction='mean' (the reference's nn.MSELoss()) is implemented")

    def __call__(self, output, target):
        global _Output
        if _Output is None:
            from .ddp import Output as _Output
        if type(output) is not _Output:
            raise TypeError("MSELoss expects the output of a distributed_training_pytorch_b200.DistributedDataParallel "
                            "forward; eager tensors are not supported (no PyTorch fallback)")
        return Loss(output, target)


class Loss:
    """A scalar that exists on the device once its step has been launched."""

    __slots__ = ("output", "target", "backend", "slot", "on_host", "value", "gvalue", "batch", "global_is_sum", "launch_no")

    def __init__(self, output, target):
        self.output, self.target = output, target
        self.backend = output.backend
        self.slot = None           # ring slot once described to the engine
        self.on_host = True        # whether that slot is in the pinned host ring or the HBM ring
        self.value = None          # the local mean loss (Python float) once it has been read
        self.gvalue = None         # the all-rank mean (fused exchange), when the launch call waited for it too
        self.batch = 0             # local batch size of the step
        self.global_is_sum = False # the reduced value that lands on the host is the squared-error SUM
        self.launch_no = 0         # the engine launch that carried the step (orders loader-block reuse)

    # -- autograd-like surface ----------------------------------------------
    def backward(self):
        self.backend.request_backward(self)

    def detach(self):
        return LossValue(self)

    def item(self):
        return float(self.detach().cpu())

    def __float__(self):
        return self.item()

    def cpu(self):
        return self.detach().cpu()

    def __add__(self, other):
        if isinstance(other, Loss):
            return LossSum([self, other])
        if isinstance(other, LossSum):
            return LossSum([self] + other.terms)
        return NotImplemented

    def __radd__(self, other):           # sum([l_X, l_Y]) starts from 0
        return self if (isinstance(other, (int, float)) and other == 0) else NotImplemented


class LossSum:
    """loss_X + loss_Y (demo_pytorch_lightning.py:27-33: one summed loss, one backward, then every
    optimizer steps).  d(sum)/d(loss_k) = 1, so backward() is each term's backward(); the value is the
    sum of the terms' values."""

    def __init__(self, terms):
        self.terms = list(terms)

    def __add__(self, other):
        if isinstance(other, Loss):
            return LossSum(self.terms + [other])
        if isinstance(other, LossSum):
            return LossSum(self.terms + other.terms)
        return NotImplemented

    __radd__ = __add__

    def backward(self):
        for t in self.terms:
            t.backward()

    def detach(self):
        return self

    def cpu(self):
        return torch.stack([t.detach().cpu() for t in self.terms]).sum()

    def item(self):
        return float(self.cpu())

    def __float__(self):
        return self.item()


class LossValue:
    """Result of loss.detach(): still lazy; .cpu() is the synchronisation point
    (demo.py:115,117), exactly where the reference blocks on its D2H copy."""

    __slots__ = ("_loss",)

    def __init__(self, loss):
        self._loss = loss

    def cpu(self):
        return self._loss.backend.read_loss(self._loss)

    def global_mean(self):
        """0-dim CPU tensor: the mean loss over ALL ranks, sum_r(loss_r * B) / (B * W) -- what
        demo.py:112-121 computes with two gloo all-reduces per step.  Here it is a by-product of the
        gradient exchange (the squared-error sum is element 371 of the bucket): no host collective.
        Like the reference's formula it assumes the same batch size on every rank (demo.py:113)."""
        backend = self._loss.backend
        if not hasattr(backend, "read_global_loss"):
            raise NotImplementedError(f"{type(backend).__name__} does not carry the loss in its gradient bucket")
        return backend.read_global_loss(self._loss)

    def item(self):
        return float(self.cpu())

    def __float__(self):
        return self.item()

    def cuda(self, *a, **k):
        return self.device_tensor()

    def device_tensor(self):
        """0-dim CUDA tensor holding the loss."""
        return self._loss.backend.loss_device_view(self._loss)

    # wandb.log({'loss/loss': loss.detach()}) (demo_one_model_multi_gpu.py:130) stringifies / converts
    def __repr__(self):
        return f"LossValue({self.item():.6g})"

    def __array__(self, dtype=None, copy=None):
        import numpy as np
        return np.asarray(self.item(), dtype=dtype or np.float32)
