"""AmoebaNet-D ``(L, D)`` for ImageNet-sized inputs as an ``nn.Sequential`` of cells.

``amoebanetd(num_classes, num_layers=L, num_filters=D)``: stem (ReLU-conv3x3
s2-BN) + two reduction cells, then three groups of ``L/3`` normal cells
separated by reduction cells, then the classifier: ``L + 6`` top-level layers
((18, 256): 24 layers, 122.5 M parameters). Cells exchange a *tuple*
``(cell output, previous cell output)``, so stage boundaries carry two tensors.
Genotype, channel bookkeeping and layer names follow reference
``benchmarks/models/amoebanet/__init__.py:18-194``, ``genotype.py:8-66`` and
``operations.py:24-114`` - including the reference quirk that ``max_pool_3x3``
is an average pool (kept so FLOPs/bytes match the benchmark being compared).
"""
from collections import OrderedDict
from typing import Iterator, List, Tuple, Union

import torch
from torch import Tensor, nn

from torchgpipe_b200.models.amoebanet.operations import (NORMAL_CONCAT, NORMAL_OPERATIONS,
                                                         REDUCTION_CONCAT, REDUCTION_OPERATIONS,
                                                         FactorizedReduce, relu_conv_bn)

__all__ = ['amoebanetd', 'Cell', 'Stem', 'Classify']


class Stem(nn.Module):
    def __init__(self, channels: int) -> None:
        super().__init__()
        self.relu = nn.ReLU(inplace=False)
        self.conv = nn.Conv2d(3, channels, 3, stride=2, padding=1, bias=False)
        self.bn = nn.BatchNorm2d(channels)

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        return self.bn(self.conv(self.relu(input)))


class Classify(nn.Module):
    def __init__(self, channels_prev: int, num_classes: int) -> None:
        super().__init__()
        self.pool = nn.AvgPool2d(7)
        self.flat = nn.Flatten()
        self.fc = nn.Linear(channels_prev, num_classes)

    def forward(self, states: Tuple[Tensor, Tensor]) -> Tensor:  # type: ignore[override]
        x, _ = states
        return self.fc(self.flat(self.pool(x)))


class Cell(nn.Module):
    """One AmoebaNet-D cell: two input reductions, five summed op pairs, channel concat."""

    def __init__(self, channels_prev_prev: int, channels_prev: int, channels: int,
                 reduction: bool, reduction_prev: bool) -> None:
        super().__init__()

        self.reduce1 = relu_conv_bn(channels_prev, channels)

        self.reduce2: nn.Module = nn.Identity()
        if reduction_prev:
            self.reduce2 = FactorizedReduce(channels_prev_prev, channels)
        elif channels_prev_prev != channels:
            self.reduce2 = relu_conv_bn(channels_prev_prev, channels)

        genotype = REDUCTION_OPERATIONS if reduction else NORMAL_OPERATIONS
        self.concat = REDUCTION_CONCAT if reduction else NORMAL_CONCAT
        self.indices, factories = zip(*genotype)

        self.operations = nn.ModuleList(
            factory(channels, 2 if (reduction and source < 2) else 1)
            for source, factory in zip(self.indices, factories))

    def extra_repr(self) -> str:
        return f'indices: {self.indices}'

    def forward(self, input_or_states: Union[Tensor, Tuple[Tensor, Tensor]],  # type: ignore[override]
                ) -> Tuple[Tensor, Tensor]:
        if isinstance(input_or_states, tuple):
            s1, s2 = input_or_states
        else:
            s1 = s2 = input_or_states

        skip = s1
        states: List[Tensor] = [self.reduce1(s1), self.reduce2(s2)]

        ops, idx = self.operations, self.indices
        for k in range(0, len(ops), 2):
            states.append(ops[k](states[idx[k]]) + ops[k + 1](states[idx[k + 1]]))

        return torch.cat([states[i] for i in self.concat], dim=1), skip


def amoebanetd(num_classes: int = 10, num_layers: int = 4, num_filters: int = 512,
               ) -> nn.Sequential:
    """Builds AmoebaNet-D; ``num_layers`` must be a multiple of 3."""
    assert num_layers % 3 == 0
    repeat = num_layers // 3

    channels = num_filters // 4
    prev_prev = prev = channels
    was_reduction = False

    def make(reduction: bool, scale: int, count: int) -> Iterator[Cell]:
        nonlocal channels, prev_prev, prev, was_reduction
        channels *= scale
        for _ in range(count):
            cell = Cell(prev_prev, prev, channels, reduction, was_reduction)
            prev_prev, prev = prev, channels * len(cell.concat)
            was_reduction = reduction
            yield cell

    def reduction_cell() -> Cell:
        return next(make(True, 2, 1))

    layers: 'OrderedDict[str, nn.Module]' = OrderedDict()
    layers['stem1'] = Stem(channels)
    layers['stem2'] = reduction_cell()
    layers['stem3'] = reduction_cell()
    for group, tail in ((1, 'cell2_reduction'), (3, 'cell4_reduction'), (5, None)):
        for k, cell in enumerate(make(False, 1, repeat)):
            layers[f'cell{group}_normal{k + 1}'] = cell
        if tail is not None:
            layers[tail] = reduction_cell()
    layers['classify'] = Classify(prev, num_classes)
    return nn.Sequential(layers)
