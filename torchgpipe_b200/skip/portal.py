"""Portals carry a skip tensor *outside* the autograd graph.

If a skip tensor travelled through the graph as part of the micro-batch it
would be copied through every intermediate stage. A portal instead keeps the
tensor in a Python object; three tiny autograd nodes tie it to the main lane:

* ``blue``   (at the stash)  : hides the tensor, emits a phony to ``join``;
  in backward it injects the skip gradient back into the stash's lane.
* ``copy``   (at the fence)  : moves the hidden tensor source->destination
  device directly on the side streams; backward moves the gradient back.
* ``orange`` (at the pop)    : takes the tensor out; backward parks the
  gradient in the portal.

``tensor_life`` counts how many more times the tensor will be taken
(stash -> pop, plus once more for each checkpoint recomputation); when it
reaches zero the reference is dropped so memory is released as early as
possible. Contract: reference ``torchgpipe/skip/portal.py:22-227``.
"""
from typing import Optional

import torch
from torch import Tensor

from torchgpipe_b200.autograd.edges import get_phony
from torchgpipe_b200.autograd.transfer import move_tensors
from torchgpipe_b200.runtime.streams import AbstractStream, get_device

__all__ = ['Portal']


class Portal:
    """Holder of one skip tensor (and, transiently, of its gradient)."""

    def __init__(self, tensor: Optional[Tensor], tensor_life: int) -> None:
        self.put_tensor(tensor, tensor_life)
        self.grad: Optional[Tensor] = None

    # -- graph attachment points -----------------------------------------
    def blue(self) -> Tensor:
        """Hides the tensor; returns a phony to be joined into the main lane."""
        tensor = self.use_tensor()
        if tensor is None:
            return get_phony(torch.device('cpu'), requires_grad=False)
        return PortalBlue.apply(self, tensor)

    def orange(self, phony: Tensor) -> Optional[Tensor]:
        """Takes the tensor back out; ``phony`` must be forked from the main lane."""
        self.check_tensor_life()
        if self.tensor is None:
            return self.use_tensor()
        return PortalOrange.apply(self, phony)

    def copy(self, prev_stream: AbstractStream, next_stream: AbstractStream,
             phony: Tensor) -> Tensor:
        """Moves the hidden tensor to the device of ``next_stream``."""
        if self.tensor is None:
            return get_phony(torch.device('cpu'), requires_grad=False)
        return PortalCopy.apply(self, prev_stream, next_stream, phony)

    # -- tensor life-cycle -------------------------------------------------
    def check_tensor_life(self) -> None:
        if self.tensor_life <= 0:
            raise RuntimeError('tensor in portal has been removed')

    def put_tensor(self, tensor: Optional[Tensor], tensor_life: int) -> None:
        self.tensor_life = tensor_life
        self.tensor = tensor if tensor_life > 0 else None

    def use_tensor(self) -> Optional[Tensor]:
        """Returns the tensor and spends one life; drops it at zero."""
        self.check_tensor_life()
        tensor = self.tensor
        self.tensor_life -= 1
        if self.tensor_life <= 0:
            self.tensor = None
        return tensor

    # -- gradient hand-over ------------------------------------------------
    def put_grad(self, grad: Tensor) -> None:
        self.grad = grad

    def use_grad(self) -> Tensor:
        if self.grad is None:
            raise RuntimeError('grad in portal has been removed or never set')
        grad, self.grad = self.grad, None
        return grad


class PortalBlue(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, tensor: Tensor) -> Tensor:  # type: ignore[override]
        ctx.portal = portal
        return get_phony(tensor.device, requires_grad=False).detach()

    @staticmethod
    def backward(ctx, grad_phony: Tensor):  # type: ignore[override]
        return None, ctx.portal.use_grad()


class PortalOrange(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, phony: Tensor) -> Tensor:  # type: ignore[override]
        ctx.portal = portal
        tensor = portal.use_tensor()
        assert tensor is not None
        return tensor.detach()

    @staticmethod
    def backward(ctx, grad: Tensor):  # type: ignore[override]
        ctx.portal.put_grad(grad)
        return None, None


class PortalCopy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, prev_stream: AbstractStream,  # type: ignore[override]
                next_stream: AbstractStream, phony: Tensor) -> Tensor:
        ctx.portal = portal
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream
        assert portal.tensor is not None
        portal.tensor, = move_tensors((portal.tensor,), prev_stream, next_stream)
        return get_phony(get_device(next_stream), requires_grad=False).detach()

    @staticmethod
    def backward(ctx, grad_phony: Tensor):  # type: ignore[override]
        portal = ctx.portal
        assert portal.grad is not None
        portal.grad, = move_tensors((portal.grad,), ctx.next_stream, ctx.prev_stream)
        return None, None, None, None
