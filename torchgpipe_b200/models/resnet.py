"""ResNet-101 as a flat ``nn.Sequential`` of 370 layers (44.5 M parameters).

Every bottleneck is unrolled into ``identity`` (stash) - conv1x1/BN/ReLU -
conv3x3(stride)/BN/ReLU - conv1x1/BN - ``residual`` (pop, optional 1x1+BN
downsample, add) - ReLU, with one skip namespace per block, so GPipe may cut a
block anywhere and the shortcut travels through a portal. Layer names and
structure follow reference ``benchmarks/models/resnet/__init__.py:20-92`` and
``bottleneck.py:31-79`` (torchvision v1.5 layout: stride on the 3x3 conv).
"""
from collections import OrderedDict
from typing import Any, List, Optional

from torch import Tensor, nn

from torchgpipe_b200.models.flatten import flatten_sequential
from torchgpipe_b200.skip import Namespace, pop, skippable, stash

__all__ = ['resnet101', 'build_resnet', 'bottleneck']


@skippable(stash=['identity'])
class Identity(nn.Module):
    """Stashes the block input for the shortcut and passes it on unchanged."""

    def forward(self, tensor: Tensor):  # type: ignore[override]
        yield stash('identity', tensor)
        return tensor


@skippable(pop=['identity'])
class Residual(nn.Module):
    """Adds the (optionally projected) block input back."""

    def __init__(self, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        self.downsample = downsample

    def forward(self, input: Tensor):  # type: ignore[override]
        identity = yield pop('identity')
        if self.downsample is not None:
            identity = self.downsample(identity)
        return input + identity


def bottleneck(inplanes: int, planes: int, stride: int = 1,
               downsample: Optional[nn.Module] = None, inplace: bool = False) -> nn.Sequential:
    """One bottleneck block as 11 sequential layers."""
    ns = Namespace()
    out_planes = planes * 4
    return nn.Sequential(OrderedDict([
        ('identity', Identity().isolate(ns)),
        ('conv1', nn.Conv2d(inplanes, planes, 1, bias=False)),
        ('bn1', nn.BatchNorm2d(planes)),
        ('relu1', nn.ReLU(inplace=inplace)),
        ('conv2', nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)),
        ('bn2', nn.BatchNorm2d(planes)),
        ('relu2', nn.ReLU(inplace=inplace)),
        ('conv3', nn.Conv2d(planes, out_planes, 1, bias=False)),
        ('bn3', nn.BatchNorm2d(out_planes)),
        ('residual', Residual(downsample).isolate(ns)),
        ('relu3', nn.ReLU(inplace=inplace)),
    ]))


def build_resnet(layers: List[int], num_classes: int = 1000, inplace: bool = False,
                 ) -> nn.Sequential:
    """Builds a bottleneck ResNet with ``layers`` blocks per stage."""
    width_in = 64

    def stage(planes: int, blocks: int, stride: int) -> nn.Sequential:
        nonlocal width_in
        projection = None
        if stride != 1 or width_in != planes * 4:
            projection = nn.Sequential(
                nn.Conv2d(width_in, planes * 4, 1, stride=stride, bias=False),
                nn.BatchNorm2d(planes * 4))
        members = [bottleneck(width_in, planes, stride, projection, inplace)]
        width_in = planes * 4
        members += [bottleneck(width_in, planes, inplace=inplace) for _ in range(blocks - 1)]
        return nn.Sequential(*members)

    model = nn.Sequential(OrderedDict([
        ('conv1', nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)),
        ('bn1', nn.BatchNorm2d(64)),
        ('relu', nn.ReLU()),
        ('maxpool', nn.MaxPool2d(3, stride=2, padding=1)),
        ('layer1', stage(64, layers[0], 1)),
        ('layer2', stage(128, layers[1], 2)),
        ('layer3', stage(256, layers[2], 2)),
        ('layer4', stage(512, layers[3], 2)),
        ('avgpool', nn.AdaptiveAvgPool2d((1, 1))),
        ('flat', nn.Flatten()),
        ('fc', nn.Linear(512 * 4, num_classes)),
    ]))
    model = flatten_sequential(model)

    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.constant_(m.weight, 1)
            nn.init.constant_(m.bias, 0)
    return model


def resnet101(**kwargs: Any) -> nn.Sequential:
    """ResNet-101: stages of 3, 4, 23, 3 bottlenecks."""
    return build_resnet([3, 4, 23, 3], **kwargs)
