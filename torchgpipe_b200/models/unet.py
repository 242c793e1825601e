"""Simplified U-Net ``(B, C)`` as a flat ``nn.Sequential``.

``unet(depth=5, num_convs=B, base_channels=C)``: each cell is ``B`` x
(conv3x3 no-bias -> Dropout2d(0.1) -> InstanceNorm2d -> LeakyReLU(0.01)); an
encoder level is cell -> stash -> MaxPool(2); a decoder level is nearest x2
upsample -> pop + pad-to-match + channel concat -> cell; one skip namespace
per level; a final 1x1 conv. (5, 64) has 241 layers / 232.7 M parameters.
Structure and names follow reference ``benchmarks/models/unet/__init__.py:20-148``.
"""
from collections import OrderedDict
from typing import Dict

import torch
from torch import Tensor, nn
import torch.nn.functional as F

from torchgpipe_b200.models.flatten import flatten_sequential
from torchgpipe_b200.skip import Namespace, pop, skippable, stash

__all__ = ['unet', 'Stash', 'PopCat']


@skippable(stash=['skip'])
class Stash(nn.Module):
    def forward(self, input: Tensor):  # type: ignore[override]
        yield stash('skip', input)
        return input


@skippable(pop=['skip'])
class PopCat(nn.Module):
    """Pops the encoder feature map and concatenates it behind the upsampled one."""

    def forward(self, input: Tensor):  # type: ignore[override]
        skipped = yield pop('skip')
        dh = skipped.size(2) - input.size(2)
        dw = skipped.size(3) - input.size(3)
        if dh or dw:
            input = F.pad(input, [0, dw, 0, dh])    # grow right / bottom to match
        return torch.cat((input, skipped), dim=1)


def conv_dropout_norm_relu(in_channels: int, out_channels: int) -> nn.Sequential:
    return nn.Sequential(OrderedDict([
        ('conv', nn.Conv2d(in_channels, out_channels, 3, padding=1, bias=False)),
        ('dropout', nn.Dropout2d(p=0.1)),
        ('norm', nn.InstanceNorm2d(out_channels)),
        ('relu', nn.LeakyReLU(negative_slope=1e-2)),
    ]))


def stacked_convs(in_channels: int, hidden_channels: int, out_channels: int, num_convs: int,
                  ) -> nn.Sequential:
    if num_convs < 1:
        return nn.Sequential()
    widths = [in_channels] + [hidden_channels] * (num_convs - 1) + [out_channels]
    return nn.Sequential(*[conv_dropout_norm_relu(a, b) for a, b in zip(widths, widths[1:])])


def unet(depth: int = 5, num_convs: int = 5, base_channels: int = 64,
         input_channels: int = 3, output_channels: int = 1) -> nn.Sequential:
    """Builds the simplified U-Net used by the speed and memory benchmarks."""
    C = base_channels

    def cell(ch: Dict[str, int]) -> nn.Sequential:
        return stacked_convs(ch['in'], ch['mid'], ch['out'], num_convs)

    enc = [{'in': input_channels if i == 0 else C * 2 ** (i - 1),
            'mid': C * 2 ** i, 'out': C * 2 ** i} for i in range(depth)]
    mid = {'in': C * 2 ** (depth - 1), 'mid': C * 2 ** depth, 'out': C * 2 ** (depth - 1)}
    dec = [{'in': C * 2 ** (i + 1), 'mid': int(C * 2 ** (i - 1)), 'out': int(C * 2 ** (i - 1))}
           for i in range(depth)]

    namespaces = [Namespace() for _ in range(depth)]

    encoder = nn.Sequential(*[
        nn.Sequential(OrderedDict([
            ('encode', cell(enc[i])),
            ('skip', Stash().isolate(namespaces[i])),
            ('down', nn.MaxPool2d(2, stride=2)),
        ])) for i in range(depth)])

    bottleneck = nn.Sequential(cell(mid))

    decoder = nn.Sequential(*[
        nn.Sequential(OrderedDict([
            ('up', nn.Upsample(scale_factor=2)),
            ('skip', PopCat().isolate(namespaces[i])),
            ('decode', cell(dec[i])),
        ])) for i in reversed(range(depth))])

    segment = nn.Conv2d(dec[0]['out'], output_channels, kernel_size=1, bias=False)

    return flatten_sequential(nn.Sequential(OrderedDict([
        ('encoder', encoder),
        ('bottleneck', bottleneck),
        ('decoder', decoder),
        ('segment', segment),
    ])))
