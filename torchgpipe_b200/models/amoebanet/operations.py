"""AmoebaNet-D primitive operations and genotype tables.

Follows reference ``benchmarks/models/amoebanet/operations.py:24-114`` and
``genotype.py:8-66``. ``NORMAL_CONCAT = [0, 3, 4, 6]`` is the TensorFlow
implementation's choice (the paper says [4, 5, 6]); it is what reproduces the
parameter counts of the GPipe paper.
"""
from typing import Any, Callable, List, Tuple

import torch
from torch import Tensor, nn

__all__ = ['Operation', 'FactorizedReduce', 'relu_conv_bn', 'NORMAL_OPERATIONS',
           'NORMAL_CONCAT', 'REDUCTION_OPERATIONS', 'REDUCTION_CONCAT']


def relu_conv_bn(in_channels: int, out_channels: int, kernel_size: Any = 1, stride: Any = 1,
                 padding: Any = 0) -> nn.Sequential:
    """Pre-activation unit used everywhere in AmoebaNet: ReLU -> conv (no bias) -> BN."""
    return nn.Sequential(
        nn.ReLU(inplace=False),
        nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding, bias=False),
        nn.BatchNorm2d(out_channels),
    )


class Operation(nn.Module):
    """Named wrapper so ``repr`` shows which genotype op a module implements."""

    def __init__(self, name: str, module: nn.Module) -> None:
        super().__init__()
        self.name = name
        self.module = module

    def __repr__(self) -> str:
        return f'{self.__class__.__name__}[{self.name}]'

    def forward(self, *args: Any) -> Any:  # type: ignore[override]
        return self.module(*args)


class FactorizedReduce(nn.Module):
    """Stride-2 reduction as two offset 1x1 stride-2 convs concatenated, then BN."""

    def __init__(self, in_channels: int, out_channels: int) -> None:
        super().__init__()
        self.relu = nn.ReLU(inplace=False)
        self.pad = nn.ZeroPad2d((0, 1, 0, 1))
        self.conv1 = nn.Conv2d(in_channels, out_channels // 2, 1, stride=2, bias=False)
        self.conv2 = nn.Conv2d(in_channels, out_channels // 2, 1, stride=2, bias=False)
        self.bn = nn.BatchNorm2d(out_channels)

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        x = self.relu(input)
        shifted = self.pad(x[:, :, 1:, 1:])
        return self.bn(torch.cat([self.conv1(x), self.conv2(shifted)], dim=1))


def none(channels: int, stride: int) -> Operation:
    body: nn.Module = nn.Identity() if stride == 1 else FactorizedReduce(channels, channels)
    return Operation('none', body)


def avg_pool_3x3(channels: int, stride: int) -> Operation:
    return Operation('avg_pool_3x3',
                     nn.AvgPool2d(3, stride=stride, padding=1, count_include_pad=False))


def max_pool_3x3(channels: int, stride: int) -> Operation:
    # Reference quirk (operations.py:57-59): this op is an *average* pool there.
    # Kept on purpose so the benchmarked network is the same network.
    return Operation('max_pool_3x3',
                     nn.AvgPool2d(3, stride=stride, padding=1, count_include_pad=False))


def max_pool_2x2(channels: int, stride: int) -> Operation:
    return Operation('max_pool_2x2', nn.MaxPool2d(2, stride=stride, padding=0))


def _chain(*units: nn.Sequential) -> nn.Sequential:
    """Concatenates ReLU-conv-BN units into one flat Sequential (indices 0..3k-1)."""
    return nn.Sequential(*[layer for unit in units for layer in unit])


def conv_1x7_7x1(channels: int, stride: int) -> Operation:
    c, q = channels, channels // 4
    return Operation('conv_1x7_7x1', _chain(
        relu_conv_bn(c, q, 1, 1, 0),
        relu_conv_bn(q, q, (1, 7), (1, stride), (0, 3)),
        relu_conv_bn(q, q, (7, 1), (stride, 1), (3, 0)),
        relu_conv_bn(q, c, 1, 1, 0),
    ))


def conv_1x1(channels: int, stride: int) -> Operation:
    return Operation('conv_1x1', relu_conv_bn(channels, channels, 1, stride, 0))


def conv_3x3(channels: int, stride: int) -> Operation:
    c, q = channels, channels // 4
    return Operation('conv_3x3', _chain(
        relu_conv_bn(c, q, 1, 1, 0),
        relu_conv_bn(q, q, 3, stride, 1),
        relu_conv_bn(q, c, 1, 1, 0),
    ))


OpFactory = Callable[[int, int], Operation]

# (index of the input state, op). Consecutive pairs are summed into a new state.
NORMAL_OPERATIONS: List[Tuple[int, OpFactory]] = [
    (1, conv_1x1), (1, max_pool_3x3),
    (1, none), (0, conv_1x7_7x1),
    (0, conv_1x1), (0, conv_1x7_7x1),
    (2, max_pool_3x3), (2, none),
    (1, avg_pool_3x3), (5, conv_1x1),
]
NORMAL_CONCAT = [0, 3, 4, 6]

REDUCTION_OPERATIONS: List[Tuple[int, OpFactory]] = [
    (0, max_pool_2x2), (0, max_pool_3x3),
    (2, none), (1, conv_3x3),
    (2, conv_1x7_7x1), (2, max_pool_3x3),
    (3, none), (1, max_pool_2x2),
    (2, avg_pool_3x3), (3, conv_1x1),
]
REDUCTION_CONCAT = [4, 5, 6]
