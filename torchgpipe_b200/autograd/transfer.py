"""Stream-aware device-to-device transfer and stream ordering as autograd nodes.

``Copy`` moves every tensor of a micro-batch from the device of one side
stream to the device of another while both side streams are current, so on
B200 the copy is an asynchronous peer DMA over NVLink 5 that occupies no SM and
overlaps with the neighbouring micro-batch's kernels. ``Wait`` orders two
streams (``cudaStreamWaitEvent``) without touching data. Both are mirrored in
backward so gradients take the reverse path on the same streams.

Contract follows reference ``torchgpipe/copy.py:27-107`` (all tensors of a
tuple are handled: regression ``tests/test_bugs.py:65-105``).
"""
from typing import Optional, Tuple

import torch
from torch import Tensor

from torchgpipe_b200.runtime.streams import (AbstractStream, current_stream, get_device,
                                             record_stream, use_stream, wait_stream)

__all__ = ['Copy', 'Wait', 'move_tensors']


def move_tensors(tensors: Tuple[Tensor, ...],
                 src_stream: AbstractStream,
                 dst_stream: AbstractStream,
                 ) -> Tuple[Tensor, ...]:
    """Copies ``tensors`` to the device of ``dst_stream`` on the two side streams.

    Source blocks are recorded on the source side stream and destination
    blocks on the destination device's *current* (compute) stream, so the
    caching allocator cannot hand either block out again while a stream that
    still uses it is behind.
    """
    dst_device = get_device(dst_stream)
    consumer_stream = current_stream(dst_device)
    moved = []
    with use_stream(src_stream), use_stream(dst_stream):
        for x in tensors:
            y = x.to(dst_device, non_blocking=True)
            moved.append(y)
            record_stream(x, src_stream)
            record_stream(y, consumer_stream)
    return tuple(moved)


class Copy(torch.autograd.Function):
    """``Copy.apply(prev_stream, next_stream, *tensors) -> tensors on next device``."""

    @staticmethod
    def forward(ctx, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                *input: Tensor) -> Tuple[Tensor, ...]:
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream
        return move_tensors(input, prev_stream, next_stream)

    @staticmethod
    def backward(ctx, *grad_output: Tensor) -> Tuple[Optional[Tensor], ...]:  # type: ignore[override]
        grads = move_tensors(grad_output, ctx.next_stream, ctx.prev_stream)
        return (None, None) + grads


class Wait(torch.autograd.Function):
    """``Wait.apply(prev_stream, next_stream, *tensors)``: next waits for prev.

    Backward makes ``prev_stream`` wait for ``next_stream``. One node serves
    the whole tuple so that *all* members are ordered, not just the first.
    """

    @staticmethod
    def forward(ctx, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                *input: Tensor) -> Tuple[Tensor, ...]:
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream
        wait_stream(next_stream, prev_stream)
        return tuple(x.detach() for x in input)

    @staticmethod
    def backward(ctx, *grad_input: Tensor) -> Tuple[Optional[Tensor], ...]:  # type: ignore[override]
        wait_stream(ctx.prev_stream, ctx.next_stream)
        return (None, None) + grad_input
