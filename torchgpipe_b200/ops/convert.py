"""Module rewriter: nn layers -> fused sm_100a operators.

``convert_to_sm100`` walks the top-level children of a flat ``nn.Sequential`` and replaces
recognised runs of layers by one fused layer followed by :class:`Absorbed` placeholders, so the
length of the sequential, the child names and the parameter names stay exactly as they were
(``balance`` lists of the reference benchmarks remain valid; ``state_dict`` keys unchanged).
It is the analogue of ``DeferredBatchNorm.convert_deferred_batch_norm`` for the compute path.
"""
from collections import OrderedDict
from typing import List, Tuple

from torch import nn

from torchgpipe_b200.ops.modules import (Absorbed, AbsorbedBatchNorm2d, AmoebaCell,
                                         AmoebaFactorizedReduce, AmoebaStem,
                                         BatchNormResidualAct,
                                         Conv2dTC, ConvBatchNormAct, ConvInstNormLeaky,
                                         GlobalAvgPoolNHWC, LinearTC, MaxPool2x2, Pool2dNHWC,
                                         UpsamplePopCat)

__all__ = ['convert_to_sm100']


def _plain_conv(m: nn.Module) -> bool:
    """A convolution some native path covers (implicit GEMM, packed stem or streaming head)."""
    from torchgpipe_b200.ops import functional as TF
    return (isinstance(m, nn.Conv2d) and m.padding_mode == 'zeros'
            and not isinstance(m.padding, str)
            and TF.conv_kind(m.in_channels, m.out_channels, m.kernel_size, m.stride,
                             m.dilation, m.groups) is not None)


def _match_unet_cell(layers: List[Tuple[str, nn.Module]], i: int) -> bool:
    if i + 3 >= len(layers):
        return False
    conv, drop, norm, act = (layers[i + k][1] for k in range(4))
    return (_plain_conv(conv) and conv.bias is None
            and isinstance(drop, nn.Dropout2d)
            and isinstance(norm, nn.InstanceNorm2d) and not norm.affine
            and not norm.track_running_stats
            and isinstance(act, nn.LeakyReLU) and act.negative_slope > 0)


def _match_conv_bn(layers: List[Tuple[str, nn.Module]], i: int) -> int:
    """0 = no match, 2 = conv+bn, 3 = conv+bn+relu starting at ``layers[i]``."""
    if i + 1 >= len(layers):
        return 0
    conv, bn = layers[i][1], layers[i + 1][1]
    if not (isinstance(conv, nn.Conv2d) and conv.bias is None and conv.groups == 1
            and conv.dilation == (1, 1) and conv.padding_mode == 'zeros'
            and not isinstance(conv.padding, str) and conv.out_channels % 4 == 0):
        return 0
    if type(bn) is not nn.BatchNorm2d or bn.num_features != conv.out_channels:
        return 0
    if i + 2 < len(layers) and type(layers[i + 2][1]) is nn.ReLU:
        return 3
    return 2


def _is_maxpool2x2(m: nn.Module) -> bool:
    def two(v):
        return v == 2 or v == (2, 2)
    return (type(m) is nn.MaxPool2d and two(m.kernel_size) and two(m.stride)
            and m.padding in (0, (0, 0)) and m.dilation in (1, (1, 1)) and not m.ceil_mode
            and not m.return_indices)


def _match_upsample_popcat(layers: List[Tuple[str, nn.Module]], i: int) -> bool:
    """``nn.Upsample(scale_factor=2, nearest)`` followed by the U-Net ``PopCat`` skippable."""
    if i + 1 >= len(layers):
        return False
    up, nxt = layers[i][1], layers[i + 1][1]
    if not (type(up) is nn.Upsample and up.mode == 'nearest' and up.size is None
            and up.scale_factor in (2, 2.0, (2, 2), (2.0, 2.0))):
        return False
    from torchgpipe_b200.skip.skippable import Skippable
    return (isinstance(nxt, Skippable) and type(nxt.module).__name__ == 'PopCat'
            and nxt.poppable_names == frozenset(['skip']) and not nxt.stashable_names)


def _is_residual(m: nn.Module) -> bool:
    """The sequential ResNet's ``Residual`` skippable: pops ``identity``, optionally projects it
    with ``downsample`` and adds it to the input."""
    from torchgpipe_b200.skip.skippable import Skippable
    return (isinstance(m, Skippable) and type(m.module).__name__ == 'Residual'
            and m.poppable_names == frozenset(['identity']) and not m.stashable_names
            and hasattr(m.module, 'downsample'))


def _square(v) -> int:
    """The common value of an int-or-pair argument, -1 if the pair is not square."""
    if isinstance(v, (tuple, list)):
        return int(v[0]) if len(v) == 2 and v[0] == v[1] else -1
    return int(v)


def _convert_pool(m: nn.Module):
    """NHWC kernel front-end for a pooling layer, or ``None`` if it is not covered."""
    if type(m) is nn.MaxPool2d:
        k, s, p = _square(m.kernel_size), _square(m.stride or m.kernel_size), _square(m.padding)
        if (min(k, s) >= 1 and 0 <= p and 2 * p <= k <= 7 and _square(m.dilation) == 1
                and not m.ceil_mode and not m.return_indices):
            return Pool2dNHWC(k, s, p, 'max')
    if type(m) is nn.AvgPool2d:
        k, s, p = _square(m.kernel_size), _square(m.stride or m.kernel_size), _square(m.padding)
        if (min(k, s) >= 1 and 0 <= p and 2 * p <= k <= 7 and not m.ceil_mode
                and m.divisor_override is None):
            return Pool2dNHWC(k, s, p, 'avg_include_pad' if m.count_include_pad else 'avg')
    if type(m) is nn.AdaptiveAvgPool2d and m.output_size in (1, (1, 1)):
        return GlobalAvgPoolNHWC()
    return None


def _children(m: nn.Module) -> List[str]:
    return [name for name, _ in m.named_children()]


def _is_amoeba_stem(m: nn.Module) -> bool:
    """``relu -> conv -> bn`` module of AmoebaNet (reference amoebanet/__init__.py:47-58)."""
    return (type(m).__name__ == 'Stem' and _children(m) == ['relu', 'conv', 'bn']
            and type(m.relu) is nn.ReLU and type(m.conv) is nn.Conv2d and m.conv.bias is None
            and type(m.bn) is nn.BatchNorm2d and _plain_conv(m.conv))


def _is_amoeba_cell(m: nn.Module) -> bool:
    """AmoebaNet-D cell (reference amoebanet/__init__.py:97-136): two input reductions, a
    ModuleList of genotype operations, pair indices and the list of states to concatenate."""
    return (type(m).__name__ == 'Cell' and _children(m) == ['reduce1', 'reduce2', 'operations']
            and isinstance(m.operations, nn.ModuleList) and len(m.operations) % 2 == 0
            and hasattr(m, 'indices') and len(m.indices) == len(m.operations)
            and hasattr(m, 'concat'))


def _is_factorized_reduce(m: nn.Module) -> bool:
    """AmoebaNet's stride-2 reduction (reference amoebanet/operations.py:24-40)."""
    if type(m).__name__ != 'FactorizedReduce' or \
            _children(m) != ['relu', 'pad', 'conv1', 'conv2', 'bn']:
        return False
    convs_ok = all(type(c) is nn.Conv2d and c.kernel_size == (1, 1) and c.stride == (2, 2)
                   and c.padding == (0, 0) and c.bias is None and c.groups == 1
                   for c in (m.conv1, m.conv2))
    return (convs_ok and type(m.bn) is nn.BatchNorm2d and type(m.pad) is nn.ZeroPad2d
            and tuple(m.pad.padding) == (0, 1, 0, 1))


def _fused_residual(residual: nn.Module, bn, relu: bool) -> nn.Module:
    downsample = residual.module.downsample
    if downsample is not None:
        downsample = _convert_any(downsample)
    fused = BatchNormResidualAct(bn, downsample, relu)
    fused.namespaces = dict(residual.namespaces)        # same skip route as the original
    return fused


def _convert_sequential(module: nn.Sequential) -> nn.Sequential:
    layers = list(module.named_children())
    out: 'OrderedDict[str, nn.Module]' = OrderedDict()
    i = 0
    while i < len(layers):
        name, layer = layers[i]
        if _match_unet_cell(layers, i):
            conv, drop, norm, act = (layers[i + k][1] for k in range(4))
            out[name] = ConvInstNormLeaky(conv, drop.p, norm.eps, act.negative_slope)
            for k in range(1, 4):
                out[layers[i + k][0]] = Absorbed(layers[i + k][1], into=name)
            i += 4
            continue
        if _match_upsample_popcat(layers, i):
            up, popcat = layers[i][1], layers[i + 1][1]
            fused = UpsamplePopCat()
            fused.namespaces = dict(popcat.namespaces)      # same skip route as the original
            out[name] = Absorbed(up, into=layers[i + 1][0])
            out[layers[i + 1][0]] = fused
            i += 2
            continue
        if _is_maxpool2x2(layer):
            out[name] = MaxPool2x2()
            i += 1
            continue
        n = _match_conv_bn(layers, i)
        if n == 2 and i + 2 < len(layers) and _is_residual(layers[i + 2][1]):
            # conv, bn, residual [, relu]: the normalisation moves into the residual join
            conv, bn, residual = (layers[i + k][1] for k in range(3))
            relu = i + 3 < len(layers) and type(layers[i + 3][1]) is nn.ReLU
            bn.__class__ = AbsorbedBatchNorm2d
            out[name] = ConvBatchNormAct(conv, bn, relu=False, defer=True)
            out[layers[i + 1][0]] = bn
            out[layers[i + 2][0]] = _fused_residual(residual, bn, relu)
            if relu:
                out[layers[i + 3][0]] = Absorbed(layers[i + 3][1], into=layers[i + 2][0])
            i += 4 if relu else 3
            continue
        if _is_residual(layer):
            relu = i + 1 < len(layers) and type(layers[i + 1][1]) is nn.ReLU
            out[name] = _fused_residual(layer, None, relu)
            if relu:
                out[layers[i + 1][0]] = Absorbed(layers[i + 1][1], into=name)
            i += 2 if relu else 1
            continue
        if n:
            conv, bn = layers[i][1], layers[i + 1][1]
            bn.__class__ = AbsorbedBatchNorm2d      # same object: parameters, buffers, keys stay
            out[name] = ConvBatchNormAct(conv, bn, relu=(n == 3))
            out[layers[i + 1][0]] = bn
            if n == 3:
                out[layers[i + 2][0]] = Absorbed(layers[i + 2][1], into=name)
            i += n
            continue
        out[name] = _convert_any(layer)
        i += 1
    converted = nn.Sequential(out)
    converted.train(module.training)
    return converted


def _convert_any(module: nn.Module) -> nn.Module:
    """Recursive rewrite of an arbitrary module tree (names of all children are preserved)."""
    if isinstance(module, (ConvInstNormLeaky, ConvBatchNormAct, Conv2dTC, Absorbed,
                           AbsorbedBatchNorm2d, MaxPool2x2, UpsamplePopCat, LinearTC, Pool2dNHWC,
                           GlobalAvgPoolNHWC, BatchNormResidualAct)):
        return module
    if type(module) is nn.Sequential:
        return _convert_sequential(module)
    if isinstance(module, (AmoebaStem, AmoebaFactorizedReduce, AmoebaCell)):
        return module
    if _is_amoeba_cell(module):
        for name, child in list(module.named_children()):       # operations, reductions first
            new = _convert_any(child)
            if new is not child:
                setattr(module, name, new)
        return AmoebaCell(module)
    if _is_amoeba_stem(module):
        return AmoebaStem(module)
    if _is_factorized_reduce(module):
        return AmoebaFactorizedReduce(module)
    pool = _convert_pool(module)
    if pool is not None:
        return pool
    if _plain_conv(module):
        return Conv2dTC(module)         # type: ignore[arg-type]
    if type(module) is nn.Linear and module.in_features % 4 == 0 \
            and module.out_features % 4 == 0:
        return LinearTC(module)
    for name, child in list(module.named_children()):
        new = _convert_any(child)
        if new is not child:
            setattr(module, name, new)
    return module


def convert_to_sm100(module: nn.Module) -> nn.Module:
    """Returns ``module`` with its hot layers running on the hand-written sm_100a kernels.

    Recognised runs inside every (nested) ``nn.Sequential``:
    ``Conv2d, Dropout2d, InstanceNorm2d, LeakyReLU`` (U-Net cell) and ``Conv2d, BatchNorm2d
    [, ReLU]`` (ResNet / AmoebaNet), ``Conv2d, BatchNorm2d, Residual [, ReLU]`` (the tail of a
    sequential bottleneck: normalisation, shortcut add and ReLU in one kernel), max / average /
    global-average pooling; any other stride-1 convolution with suitable channel counts
    becomes a :class:`Conv2dTC`. Number, order and names of layers and parameters never change.
    """
    return _convert_any(module)
