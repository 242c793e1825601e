"""nn.Module front-ends of the fused sm_100a operators."""
from typing import Optional

import torch
from torch import Tensor, nn
import torch.nn.functional as F

from torchgpipe_b200.ops import functional as TF

__all__ = ['Absorbed', 'ConvInstNormLeaky', 'Conv2dTC']

_CL = torch.channels_last


class Absorbed(nn.Module):
    """Placeholder for a layer whose work now happens inside a preceding fused layer.

    Keeps the position and name of the original layer so ``balance`` lists and layer counts are
    unchanged by :func:`~torchgpipe_b200.ops.convert_to_sm100`.
    """

    def __init__(self, original: nn.Module, into: str) -> None:
        super().__init__()
        self.original_repr = repr(original)
        self.into = into

    def extra_repr(self) -> str:
        return f'{self.original_repr} -> fused into {self.into!r}'

    def forward(self, input):  # type: ignore[override]
        return input


class Conv2dTC(nn.Module):
    """Stride-1 ``nn.Conv2d`` (no bias) on the tcgen05 kernels; same ``weight`` parameter."""

    def __init__(self, conv: nn.Conv2d) -> None:
        super().__init__()
        self.in_channels, self.out_channels = conv.in_channels, conv.out_channels
        self.kernel_size, self.padding = conv.kernel_size, conv.padding
        self.weight = conv.weight
        self.weight.data = self.weight.data.contiguous(memory_format=_CL)
        self.bias = conv.bias
        self._cache: dict = {}

    def refresh_derived_weights(self) -> None:
        """Updates the cached dgrad weights in place after an optimiser step (CUDA-graph mode)."""
        TF.refresh_dgrad_weight(self.weight, self._cache)

    def extra_repr(self) -> str:
        return (f'{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, '
                f'padding={self.padding}, tcgen05/tf32')

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if input.is_cuda and input.dtype == torch.float32 and TF.conv_supported(
                self.in_channels, self.out_channels, 1, 1, 1):
            y, _ = TF.conv2d(input, self.weight, self.padding, cache=self._cache)
            if self.bias is not None:
                y = y + self.bias.view(1, -1, 1, 1)
            return y
        return F.conv2d(input, self.weight, self.bias, padding=self.padding)


class ConvInstNormLeaky(nn.Module):
    """``Conv2d(bias=False) -> Dropout2d(p) -> InstanceNorm2d -> LeakyReLU`` as two kernels.

    The convolution's epilogue accumulates the per-(image, channel) sum and sum of squares, so
    the normalisation needs one read and one write of the activation; the channel dropout is a
    per-(n,c) factor folded into the normalisation coefficient (its mask comes from PyTorch's
    CUDA generator, so checkpoint recomputation reproduces it). Reference: the unfused cell of
    ``benchmarks/models/unet/__init__.py:42-48``.
    """

    def __init__(self, conv: nn.Conv2d, p: float, eps: float, slope: float) -> None:
        super().__init__()
        self.in_channels, self.out_channels = conv.in_channels, conv.out_channels
        self.kernel_size, self.padding = conv.kernel_size, conv.padding
        self.weight = conv.weight
        self.weight.data = self.weight.data.contiguous(memory_format=_CL)
        self.p, self.eps, self.slope = float(p), float(eps), float(slope)
        self._cache: dict = {}

    def extra_repr(self) -> str:
        return (f'{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, '
                f'padding={self.padding}, dropout2d={self.p}, eps={self.eps}, '
                f'leaky={self.slope}, tcgen05/tf32')

    def refresh_derived_weights(self) -> None:
        """Updates the cached dgrad weights in place after an optimiser step (CUDA-graph mode)."""
        TF.refresh_dgrad_weight(self.weight, self._cache)

    def _drop_scale(self, n: int, device: torch.device) -> Optional[Tensor]:
        if not self.training or self.p == 0.0:
            return None
        keep = torch.rand(n, self.out_channels, device=device) >= self.p
        return keep.to(torch.float32).mul_(1.0 / (1.0 - self.p))

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if not (input.is_cuda and input.dtype == torch.float32):
            y = F.conv2d(input, self.weight, None, padding=self.padding)
            y = F.dropout2d(y, self.p, self.training)
            return F.leaky_relu(F.instance_norm(y, eps=self.eps), self.slope)

        scale = self._drop_scale(input.size(0), input.device)
        if TF.conv_supported(self.in_channels, self.out_channels, 1, 1, 1):
            y, stats = TF.conv2d(input, self.weight, self.padding, stats_mode=1,
                                 cache=self._cache)
        else:
            # e.g. the 3-channel stem: library convolution, statistics by the moments kernel
            x = input.contiguous(memory_format=_CL)
            y, stats = F.conv2d(x, self.weight, None, padding=self.padding), None
        return TF.instnorm_leaky(y, stats, scale, self.eps, self.slope)
