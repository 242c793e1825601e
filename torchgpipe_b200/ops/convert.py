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

from torchgpipe_b200.ops.modules import Absorbed, Conv2dTC, ConvInstNormLeaky

__all__ = ['convert_to_sm100']


def _plain_conv(m: nn.Module) -> bool:
    return (isinstance(m, nn.Conv2d) and m.stride == (1, 1) and m.dilation == (1, 1)
            and m.groups == 1 and m.padding_mode == 'zeros' and not isinstance(m.padding, str))


def _match_unet_cell(layers: List[Tuple[str, nn.Module]], i: int) -> bool:
    if i + 3 >= len(layers):
        return False
    conv, drop, norm, act = (layers[i + k][1] for k in range(4))
    return (_plain_conv(conv) and conv.bias is None
            and isinstance(drop, nn.Dropout2d)
            and isinstance(norm, nn.InstanceNorm2d) and not norm.affine
            and not norm.track_running_stats
            and isinstance(act, nn.LeakyReLU) and act.negative_slope > 0)


def convert_to_sm100(module: nn.Module) -> nn.Module:
    """Returns ``module`` with its hot layers running on the hand-written sm_100a kernels."""
    if not isinstance(module, nn.Sequential):
        return module

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
        if _plain_conv(layer) and layer.in_channels % 4 == 0 and layer.out_channels % 32 == 0:
            out[name] = Conv2dTC(layer)
        else:
            out[name] = layer
        i += 1
    return nn.Sequential(out)
