"""Small helpers for the distributed mode."""
from typing import Optional, Tuple, Union

import torch
from torch import Tensor

__all__ = ['to']

Tensors = Tuple[Optional[Tensor], ...]
TensorOrTensors = Union[Tensor, Tensors, None]


def _move_one(device: Optional[torch.device], value: Optional[Tensor]) -> Optional[Tensor]:
    if value is None:
        return None
    moved = value.detach()
    if device is not None:
        moved = moved.to(device)
    if value.requires_grad:
        moved.requires_grad_()
    return moved


def to(device: Optional[torch.device], value: TensorOrTensors) -> TensorOrTensors:
    """Detaches ``value`` (tensor, tuple with optional ``None`` members, or ``None``),
    moves it to ``device`` and restores the ``requires_grad`` flag of every member.

    Contract: reference ``torchgpipe/distributed/utils.py:6-22``.
    """
    if value is None:
        return None
    if isinstance(value, tuple):
        return tuple(_move_one(device, v) for v in value)
    return _move_one(device, value)
