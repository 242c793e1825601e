"""Zero-byte ordering edges in the autograd graph.

A *phony* is an empty tensor. Routing a phony from one autograd lane into
another (``fork`` on the source lane, ``join`` on the destination lane) makes
the backward of the destination lane run before the backward of the source
lane without saving or moving any data. The pipeline uses these edges to make
micro-batch ``i`` finish its backward before ``i-1`` on every stage, to attach
recomputation, and to tie skip portals to the main lane.

Behavioural contract follows reference ``torchgpipe/dependency.py:12-48`` and
``torchgpipe/phony.py:15-43``.
"""
from typing import Dict, Tuple

import torch
from torch import Tensor

from torchgpipe_b200.runtime.streams import default_stream, use_stream

__all__ = ['get_phony', 'fork', 'join', 'Fork', 'Join']

_PHONY_CACHE: Dict[Tuple[torch.device, bool], Tensor] = {}


def get_phony(device: torch.device, *, requires_grad: bool) -> Tensor:
    """Returns the cached 0-element tensor for ``(device, requires_grad)``.

    The phony is allocated on the device's default stream so that no side
    stream ever owns it. An autograd Function that hands a cached phony back
    to autograd must ``detach()`` it first, otherwise the engine would attach
    a ``grad_fn`` to the shared object.
    """
    key = (device, requires_grad)
    phony = _PHONY_CACHE.get(key)
    if phony is None:
        with use_stream(default_stream(device)):
            phony = torch.empty(0, device=device, requires_grad=requires_grad)
        _PHONY_CACHE[key] = phony
    return phony


class Fork(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input: Tensor):  # type: ignore[override]
        phony = get_phony(input.device, requires_grad=False)
        return input.detach(), phony.detach()

    @staticmethod
    def backward(ctx, grad_input: Tensor, grad_phony: Tensor):  # type: ignore[override]
        # The phony branch carries no gradient; it only delays this node until
        # the joined lane has been back-propagated.
        return grad_input


class Join(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input: Tensor, phony: Tensor):  # type: ignore[override]
        return input.detach()

    @staticmethod
    def backward(ctx, grad_input: Tensor):  # type: ignore[override]
        return grad_input, None


def fork(input: Tensor) -> Tuple[Tensor, Tensor]:
    """Splits a phony branch off the lane of ``input``.

    No autograd node is created when gradients are off or ``input`` does not
    require grad; a shared non-differentiable phony is returned instead.
    """
    if torch.is_grad_enabled() and input.requires_grad:
        input, phony = Fork.apply(input)
        return input, phony
    return input, get_phony(input.device, requires_grad=False)


def join(input: Tensor, phony: Tensor) -> Tensor:
    """Merges a phony branch into the lane of ``input``."""
    if torch.is_grad_enabled() and (input.requires_grad or phony.requires_grad):
        return Join.apply(input, phony)
    return input
