"""CPU checks of the module rewriter: structure, names, and exact (float64) parity of the ATen
fall-back path every fused layer has. The sm_100a kernels themselves are tested in
``test_ops_gpu.py``."""
import copy

import pytest
import torch
from torch import nn

from torchgpipe_b200 import GPipe, ops
from torchgpipe_b200.models import amoebanetd, resnet101, unet
from torchgpipe_b200.models.resnet import build_resnet


def _max_rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


def test_resnet_conversion_keeps_layers_names_and_keys():
    model = resnet101(num_classes=10)
    keys, names = list(model.state_dict().keys()), [n for n, _ in model.named_children()]
    fused = ops.convert_to_sm100(model)
    assert list(fused.state_dict().keys()) == keys
    assert [n for n, _ in fused.named_children()] == names
    kinds = [type(m).__name__ for m in fused]
    assert kinds.count('BatchNormResidualAct') == 33 and 'Residual' not in kinds
    assert kinds.count('ConvBatchNormAct') == 100 and kinds.count('AbsorbedBatchNorm2d') == 100
    assert 'Pool2dNHWC' in kinds and 'GlobalAvgPoolNHWC' in kinds
    deferring = [m for m in fused if type(m).__name__ == 'ConvBatchNormAct' and m.defer]
    assert len(deferring) == 33


def test_converted_resnet_is_exact_in_float64_through_gpipe_with_checkpointing():
    torch.manual_seed(0)
    plain = build_resnet([1, 1, 1, 1], num_classes=10).double()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    n = len(fused)
    # stem (4) + conv1..conv3 of the first block = 12: the cut falls between a deferring
    # convolution and its BatchNorm, so the (conv, statistics) tuple crosses a stage boundary
    pipe = GPipe(fused, [12, n - 12], devices=['cpu', 'cpu'], chunks=2, checkpoint='always')
    x = torch.rand(4, 3, 64, 64, dtype=torch.float64)
    want = torch.cat([plain(x[:2]), plain(x[2:])])
    got = pipe(x)
    assert _max_rel(got, want) < 1e-9
    got.square().mean().backward()
    want.square().mean().backward()
    mine = {k.split('.', 2)[2]: v for k, v in pipe.named_parameters()}
    for name, p in plain.named_parameters():
        assert _max_rel(mine[name].grad, p.grad) < 1e-8, name
    # running statistics were updated once per micro-batch although every stage recomputed
    state = {k.split('.', 2)[2]: v for k, v in pipe.state_dict().items()}
    for name, b in plain.state_dict().items():
        assert torch.allclose(state[name].double(), b.double(), atol=1e-9), name


def test_eval_mode_uses_running_statistics():
    torch.manual_seed(1)
    plain = build_resnet([1, 1, 1, 1], num_classes=10)
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    x = torch.rand(2, 3, 64, 64)
    plain.eval()
    fused.eval()
    with torch.no_grad():
        assert torch.allclose(fused(x), plain(x), atol=1e-5)


def test_balance_that_separates_a_fused_layer_from_its_partner_is_refused():
    fused = ops.convert_to_sm100(resnet101(num_classes=10))
    n = len(fused)
    with pytest.raises(ValueError, match='fused run'):
        GPipe(fused, [4 + 9, n - 13], devices=['cpu', 'cpu'])      # bn3 | residual
    with pytest.raises(ValueError, match='fused run'):
        GPipe(fused, [4 + 2, n - 6], devices=['cpu', 'cpu'])       # conv1 | bn1
    # the reference's published balances are all fine
    for balance in ([135, 235], [44, 92, 124, 110], [26, 22, 33, 44, 44, 66, 66, 69]):
        GPipe(ops.convert_to_sm100(resnet101(num_classes=10)), balance,
              devices=['cpu'] * len(balance))


def test_plain_residual_without_batchnorm_partner():
    from torchgpipe_b200.models.resnet import Identity, Residual
    from torchgpipe_b200.skip import Namespace
    ns = Namespace()
    plain = nn.Sequential(Identity().isolate(ns), nn.Conv2d(8, 8, 3, padding=1),
                          Residual().isolate(ns), nn.ReLU())
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    assert type(fused[2]).__name__ == 'BatchNormResidualAct' and fused[2].module.bn is None
    assert type(fused[3]).__name__ == 'Absorbed'
    x = torch.rand(2, 8, 5, 5)
    a = GPipe(plain, [4], devices=['cpu'], chunks=1)(x)
    b = GPipe(fused, [4], devices=['cpu'], chunks=1)(x)
    assert torch.allclose(a, b, atol=1e-6)


def test_pooling_layers_are_rewritten_and_fall_back_on_cpu():
    seq = nn.Sequential(nn.MaxPool2d(3, 2, 1), nn.AvgPool2d(3, 1, 1, count_include_pad=False),
                        nn.AvgPool2d(2), nn.MaxPool2d(3, 2, 1, ceil_mode=True),
                        nn.AdaptiveAvgPool2d(1), nn.AdaptiveAvgPool2d(2))
    fused = ops.convert_to_sm100(copy.deepcopy(seq))
    kinds = [type(m).__name__ for m in fused]
    assert kinds == ['Pool2dNHWC', 'Pool2dNHWC', 'Pool2dNHWC', 'MaxPool2d', 'GlobalAvgPoolNHWC',
                     'AdaptiveAvgPool2d']
    assert (fused[1].mode, fused[2].mode) == ('avg', 'avg_include_pad')
    x = torch.rand(2, 4, 17, 17)
    assert torch.allclose(fused(x), seq(x))


def test_unet_and_amoebanet_conversion_keep_keys():
    for model in (unet(depth=2, num_convs=2, base_channels=8),
                  amoebanetd(num_classes=10, num_layers=3, num_filters=16)):
        keys = list(model.state_dict().keys())
        n = len(model)
        fused = ops.convert_to_sm100(model)
        assert list(fused.state_dict().keys()) == keys and len(fused) == n


def test_forward_convolution_plans_are_pinned():
    """Host-side decisions of the tcgen05 forward kernel (no GPU needed: 148 SMs assumed)."""
    from torchgpipe_b200.ops import _lib
    if _lib.load() is None:
        pytest.skip('native library not built')
    from torchgpipe_b200.ops.functional import conv2d_plan
    # narrow, high-resolution layer: haloed 8x16 tiles, filter bank resident in shared memory
    p = conv2d_plan(40, 64, 192, 192, 64)
    assert (p['BN'], p['BW'], p['BH'], p['BNB']) == (64, 8, 16, 1)
    assert p['halo'] and p['resident_weights'] and p['splits'] == 1 and p['stages'] == 2
    # 256 -> 64: the filter bank (576 KB) cannot stay resident, weights stream with the stages
    p = conv2d_plan(40, 256, 96, 96, 64)
    assert p['halo'] and not p['resident_weights']
    # too few tiles for the haloed variant: the per-tap kernel
    assert not conv2d_plan(2, 64, 16, 16, 64)['halo']
    # 12x12 images: row bands of several images fill 120 of 128 MMA rows; reduction split in 3
    p = conv2d_plan(40, 1024, 12, 12, 1024)
    assert (p['BN'], p['BW'], p['BH'], p['BNB']) == (256, 12, 2, 5)
    assert not p['halo'] and p['splits'] == 3 and p['splits'] * p['iters_per_split'] >= 288
    # per-image statistics make extra images per tile expensive: 24x24 keeps one image per tile
    assert conv2d_plan(40, 256, 24, 24, 512, stats_mode=1)['BNB'] == 1
    assert conv2d_plan(40, 256, 24, 24, 512, stats_mode=0)['BNB'] == 5
    # wide layers with many tiles: no split
    assert conv2d_plan(40, 256, 48, 48, 256)['splits'] == 1
    # 1x1 convolution / linear layer: never haloed
    p = conv2d_plan(40, 2048, 1, 1, 1000, 1, 1, (0, 0))
    assert not p['halo'] and p['Ho'] == 1 and p['tiles_k'] * p['BN'] >= 1000
    with pytest.raises(RuntimeError, match='C % 4'):
        conv2d_plan(1, 3, 8, 8, 64)
