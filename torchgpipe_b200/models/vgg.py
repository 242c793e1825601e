"""VGG-16 and a coarse (11-child) ResNet-101 as ``nn.Sequential`` - the two models of the
reference's multi-process demo (``benchmarks/distributed/accuracy/{vgg,resnet}/__init__.py``).

``vgg16``: 13 conv3x3+ReLU in five max-pooled groups, adaptive 7x7 average pool + flatten (one
child), three linear layers with dropout: 39 top-level children. ``resnet101_coarse``: stem
(conv, bn, relu, maxpool), four residual stages as single children, avgpool, flatten, fc - stages
are whole modules here, so a balance over 11 children cuts only at stage boundaries (the
fine-grained 370-layer variant is :func:`torchgpipe_b200.models.resnet101`).
"""
from typing import List

from torch import Tensor, nn

__all__ = ['vgg16', 'resnet101_coarse', 'resnet50_coarse']

_VGG16 = [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 'M', 512, 512, 512, 'M', 512, 512, 512, 'M']


def vgg16(num_classes: int = 1000, inplace: bool = False) -> nn.Sequential:
    layers: List[nn.Module] = []
    channels = 3
    for v in _VGG16:
        if v == 'M':
            layers.append(nn.MaxPool2d(kernel_size=2, stride=2))
        else:
            layers += [nn.Conv2d(channels, v, kernel_size=3, padding=1), nn.ReLU(inplace=inplace)]
            channels = v
    layers.append(nn.Sequential(nn.AdaptiveAvgPool2d((7, 7)), nn.Flatten()))
    layers += [nn.Linear(512 * 7 * 7, 4096), nn.ReLU(inplace=inplace), nn.Dropout(0.5),
               nn.Linear(4096, 4096), nn.ReLU(inplace=inplace), nn.Dropout(0.5),
               nn.Linear(4096, num_classes)]
    model = nn.Sequential(*layers)
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, 0, 0.01)
            nn.init.zeros_(m.bias)
    return model


class _Bottleneck(nn.Module):
    """Standard (non-sequentialised) bottleneck block with an in-module shortcut."""

    def __init__(self, inplanes: int, planes: int, stride: int, project: bool, inplace: bool):
        super().__init__()
        out = planes * 4
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, out, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(out)
        self.relu = nn.ReLU(inplace=inplace)
        self.downsample = nn.Sequential(nn.Conv2d(inplanes, out, 1, stride=stride, bias=False),
                                        nn.BatchNorm2d(out)) if project else None

    def forward(self, x: Tensor) -> Tensor:  # type: ignore[override]
        identity = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return self.relu(y + identity)


def _coarse_resnet(blocks: List[int], num_classes: int, inplace: bool) -> nn.Sequential:
    inplanes = 64

    def stage(planes: int, count: int, stride: int) -> nn.Sequential:
        nonlocal inplanes
        members = [_Bottleneck(inplanes, planes, stride, True, inplace)]
        inplanes = planes * 4
        members += [_Bottleneck(inplanes, planes, 1, False, inplace) for _ in range(count - 1)]
        return nn.Sequential(*members)

    return nn.Sequential(
        nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False), nn.BatchNorm2d(64),
        nn.ReLU(inplace=inplace), nn.MaxPool2d(3, stride=2, padding=1),
        stage(64, blocks[0], 1), stage(128, blocks[1], 2), stage(256, blocks[2], 2),
        stage(512, blocks[3], 2),
        nn.AdaptiveAvgPool2d((1, 1)), nn.Flatten(), nn.Linear(2048, num_classes))


def resnet101_coarse(num_classes: int = 1000, inplace: bool = False) -> nn.Sequential:
    return _coarse_resnet([3, 4, 23, 3], num_classes, inplace)


def resnet50_coarse(num_classes: int = 1000, inplace: bool = False) -> nn.Sequential:
    """(The reference's ``resnet50`` is a stub returning ``None``; this one is real.)"""
    return _coarse_resnet([3, 4, 6, 3], num_classes, inplace)
