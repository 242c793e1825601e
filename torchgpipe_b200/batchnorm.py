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
from typing import Optional, TypeVar, cast

import torch
from torch import Tensor, nn
import torch.nn.functional as F
from torch.nn.modules.batchnorm import _BatchNorm

from torchgpipe_b200.checkpoint import is_recomputing

__all__ = ['DeferredBatchNorm']

TModule = TypeVar('TModule', bound=nn.Module)


def _channel_moments(input: Tensor):
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


class DeferredBatchNorm(_BatchNorm):
    """BatchNorm that commits running statistics once per ``chunks`` micro-batches."""

    sum: Tensor
    sum_squares: Tensor

    def __init__(self,
                 num_features: int,
                 eps: float = 1e-5,
                 momentum: Optional[float] = 0.1,
                 affine: bool = True,
                 chunks: int = 1,
                 ) -> None:
        super().__init__(num_features, eps, momentum, affine, track_running_stats=True)

        self.register_buffer('sum', torch.zeros_like(self.running_mean))
        self.register_buffer('sum_squares', torch.zeros_like(self.running_var))

        self.counter = 0     # elements per channel accumulated so far
        self.tracked = 0     # micro-batches accumulated so far
        self.chunks = chunks

    def _check_input_dim(self, input: Tensor) -> None:
        # Any (N, C, *) input is fine: 1-d, 2-d and 3-d BatchNorm alike.
        if input.dim() <= 2:
            raise ValueError('expected at least 3D input (got %dD input)' % input.dim())

    def _track(self, input: Tensor) -> bool:
        """Adds one micro-batch to the accumulators; True when the mini-batch is complete."""
        with torch.no_grad():
            s, ss = _channel_moments(input)
            self.sum += s
            self.sum_squares += ss

        self.counter += input.numel() // input.size(1)
        self.tracked += 1
        return self.tracked == self.chunks

    def _commit(self) -> None:
        """Folds the accumulated mini-batch statistics into the running estimates."""
        self.num_batches_tracked += 1
        if self.momentum is None:
            factor = 1.0 / float(self.num_batches_tracked)   # cumulative average
        else:
            factor = self.momentum                           # exponential average

        mean = self.sum / self.counter
        var = self.sum_squares / self.counter - mean ** 2

        self.running_mean.mul_(1 - factor).add_(mean * factor)
        self.running_var.mul_(1 - factor).add_(var * factor)

        self.sum.zero_()
        self.sum_squares.zero_()
        self.counter = 0
        self.tracked = 0

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        self._check_input_dim(input)
        if not self.training:
            return F.batch_norm(input, self.running_mean, self.running_var,
                                self.weight, self.bias,
                                training=False, momentum=0.0, eps=self.eps)

        if not is_recomputing():
            if self._track(input):
                self._commit()

        # Normalise with this micro-batch's own statistics.
        return F.batch_norm(input, None, None, self.weight, self.bias,
                            training=True, momentum=0.0, eps=self.eps)

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
