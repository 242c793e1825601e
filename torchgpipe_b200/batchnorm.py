"""BatchNorm whose running statistics follow the *mini*-batch.

Micro-batching changes what ``nn.BatchNorm`` sees: ``chunks`` small batches
instead of one large one, so running estimates drift from what un-pipelined
training would produce. ``DeferredBatchNorm`` normalises each micro-batch with
that micro-batch's own statistics (exactly like ``nn.BatchNorm`` in training
mode) but accumulates per-channel sum and sum of squares across the
micro-batches of a mini-batch and commits one running-statistics update when
all ``chunks`` micro-batches have been seen. Nothing is tracked while a
checkpoint is being recomputed.

Contract: reference ``torchgpipe/batchnorm.py:24-155`` - extra buffers ``sum``
and ``sum_squares`` are part of ``state_dict``; the committed variance is the
biased ``E[x^2] - E[x]^2`` (reference ``:70-71``).

On a B200 the per-channel sum / sum-of-squares come from one fused pass
(``ops.channel_moments``) instead of two full-tensor reductions.
"""
from typing import Optional, Tuple, TypeVar, cast

import torch
from torch import Tensor, nn
import torch.nn.functional as F
from torch.nn.modules.batchnorm import _BatchNorm

from torchgpipe_b200.checkpoint import is_recomputing

__all__ = ['DeferredBatchNorm', 'MiniBatchMoments']

TModule = TypeVar('TModule', bound=nn.Module)


def _channel_moments(input: Tensor) -> Tuple[Tensor, Tensor]:
    """Per-channel (sum, sum of squares) over every dim but 1."""
    if input.is_cuda and input.dim() == 4:
        try:
            from torchgpipe_b200.ops import fused
            if fused.available():
                return fused.channel_moments(input)
        except ImportError:
            pass
    dims = [0] + list(range(2, input.dim()))
    return input.sum(dims), (input * input).sum(dims)


class MiniBatchMoments:
    """The bookkeeping of one deferred BatchNorm: how many micro-batches and elements have been
    folded into the owner's ``sum`` / ``sum_squares`` buffers, and the commit into its running
    estimates once the mini-batch is complete.

    Shared by :class:`DeferredBatchNorm` (torch arithmetic) and the fused layers' BatchNorm
    (``ops.modules.AbsorbedBatchNorm2d``, whose moments come out of the convolution epilogue).
    The tensors stay buffers of the owner so that ``state_dict`` has the reference's keys; only
    the two host-side counts live here.
    """

    __slots__ = ('owner', 'chunks', 'elements', 'seen')

    def __init__(self, owner: _BatchNorm, chunks: int) -> None:
        self.owner = owner
        self.chunks = chunks
        self.elements = 0       # per channel, over the micro-batches seen so far
        self.seen = 0           # micro-batches of the current mini-batch

    def absorb(self, total: Tensor, total_of_squares: Tensor, elements: int) -> None:
        """One micro-batch's per-channel moments; commits when it was the mini-batch's last."""
        bn = self.owner
        with torch.no_grad():
            bn.sum += total
            bn.sum_squares += total_of_squares
            self.elements += elements
            self.seen += 1
            if self.seen == self.chunks:
                self._commit()

    def _commit(self) -> None:
        bn = self.owner
        bn.num_batches_tracked += 1
        # momentum None: cumulative average over the mini-batches seen since reset
        weight = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
        mean = bn.sum / self.elements
        biased_var = bn.sum_squares / self.elements - mean * mean   # as nn.BatchNorm normalises
        bn.running_mean.lerp_(mean, weight)
        bn.running_var.lerp_(biased_var, weight)
        bn.sum.zero_()
        bn.sum_squares.zero_()
        self.elements = self.seen = 0


class DeferredBatchNorm(_BatchNorm):
    """BatchNorm that commits running statistics once per ``chunks`` micro-batches."""

    sum: Tensor
    sum_squares: Tensor

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: Optional[float] = 0.1,
                 affine: bool = True, chunks: int = 1) -> None:
        super().__init__(num_features, eps, momentum, affine, track_running_stats=True)
        self.register_buffer('sum', torch.zeros_like(self.running_mean))
        self.register_buffer('sum_squares', torch.zeros_like(self.running_var))
        self.chunks = chunks
        self.moments = MiniBatchMoments(self, chunks)

    # the reference's attribute names for the two counts
    @property
    def counter(self) -> int:
        return self.moments.elements

    @property
    def tracked(self) -> int:
        return self.moments.seen

    def _check_input_dim(self, input: Tensor) -> None:
        # Any (N, C, *) input is fine: 1-d, 2-d and 3-d BatchNorm alike.
        if input.dim() <= 2:
            raise ValueError('expected at least 3D input (got %dD input)' % input.dim())

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        self._check_input_dim(input)
        running = not self.training
        if self.training and not is_recomputing():
            # a recomputation sees the same micro-batch a second time: it must not count twice
            self.moments.absorb(*_channel_moments(input), input.numel() // input.size(1))
        # training: this micro-batch's own statistics, exactly like nn.BatchNorm would
        return F.batch_norm(input, self.running_mean if running else None,
                            self.running_var if running else None, self.weight, self.bias,
                            training=not running, momentum=0.0, eps=self.eps)

    @classmethod
    def convert_deferred_batch_norm(cls, module: TModule, chunks: int = 1) -> TModule:
        """Replaces every tracking ``nn.BatchNorm*`` under ``module`` (recursively).

        Parameters and running buffers are shared with the original layers.
        BatchNorm layers with ``track_running_stats=False`` are left alone. A
        module that already is a ``DeferredBatchNorm`` with the same ``chunks``
        is returned unchanged::

            model = DeferredBatchNorm.convert_deferred_batch_norm(resnet101(), chunks=8)
        """
        if isinstance(module, DeferredBatchNorm) and module.chunks is chunks:
            return cast(TModule, module)

        # A BatchNorm whose arithmetic lives in a fused sm_100a layer keeps running there; it
        # only switches to mini-batch statistics tracking.
        if hasattr(module, 'make_deferred'):
            if getattr(module, 'track_running_stats', False):
                module.make_deferred(chunks)  # type: ignore[operator]
            return cast(TModule, module)

        converted: nn.Module = module

        if isinstance(module, _BatchNorm) and module.track_running_stats:
            converted = DeferredBatchNorm(module.num_features, module.eps, module.momentum,
                                          module.affine, chunks)
            if module.affine:
                converted.register_parameter('weight', module.weight)
                converted.register_parameter('bias', module.bias)
            converted.register_buffer('running_mean', module.running_mean)
            converted.register_buffer('running_var', module.running_var)
            converted.register_buffer('num_batches_tracked', module.num_batches_tracked)

        for name, child in module.named_children():
            converted.add_module(name, cls.convert_deferred_batch_norm(child, chunks))

        return cast(TModule, converted)
