"""Flat gradient buckets in symmetric memory.

``GradientBucket(params, group)`` makes every ``param.grad`` a view into one flat fp32 buffer
that all data-parallel replicas of the stage can address, so the backward pass accumulates
straight into the memory the all-reduce kernel works on (no pack/unpack copies).
"""
from typing import Any, Iterable, List, Optional

import torch
import torch.distributed as dist
from torch import Tensor, nn

from torchgpipe_b200.parallel.allreduce import SymmetricBuffer, fused_allreduce_

__all__ = ['GradientBucket']


class GradientBucket:
    """All gradients of one pipeline stage as one symmetric buffer."""

    def __init__(self, params: Iterable[nn.Parameter], group: Any = None,
                 use_symmetric: bool = True) -> None:
        self.params: List[nn.Parameter] = [p for p in params if p.requires_grad]
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        total = sum(p.numel() for p in self.params)
        device = self.params[0].device
        self.symmetric: Optional[SymmetricBuffer] = None
        if use_symmetric and device.type == 'cuda' and self.world > 1:
            self.symmetric = SymmetricBuffer(total, self.group, device)
            self.flat = self.symmetric.tensor
        else:
            self.flat = torch.zeros(total, dtype=torch.float32, device=device)
        self.attach()

    def attach(self) -> None:
        """(Re)binds every ``param.grad`` to its slice of the flat buffer (call after
        ``zero_grad(set_to_none=True)``; use :meth:`zero` instead of ``zero_grad``)."""
        offset = 0
        for p in self.params:
            n = p.numel()
            view = self.flat[offset:offset + n].view(p.shape)
            p.grad = view
            offset += n

    def zero(self) -> None:
        self.flat.zero_()

    def allreduce(self, scale: Optional[float] = None, multicast: Optional[bool] = None,
                  ) -> Tensor:
        """``grad <- scale * sum over replicas`` (default scale 1/replicas), on the current stream."""
        if scale is None:
            scale = 1.0 / self.world
        if self.world == 1:
            if scale != 1.0:
                self.flat.mul_(scale)
            return self.flat
        if self.symmetric is not None:
            return fused_allreduce_(self.symmetric, scale, multicast=multicast)
        dist.all_reduce(self.flat, group=self.group)        # library baseline / fallback
        if scale != 1.0:
            self.flat.mul_(scale)
        return self.flat

    def check(self) -> None:
        """Raises if the fused all-reduce of this bucket timed out (see ``SymmetricBuffer.check``);
        a no-op on the NCCL fallback, whose own watchdog reports failures."""
        if self.symmetric is not None:
            self.symmetric.check()
