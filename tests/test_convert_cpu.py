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


def test_conv_kind_routes_every_benchmark_convolution_to_a_native_path():
    from torchgpipe_b200.ops import functional as TF
    # (c_in, c_out, kernel, stride) -> path; reference call sites in the comments
    cases = [
        ((3, 64, 7, 2), 'stem'),            # resnet/__init__.py:55
        ((3, 64, 3, 2), 'stem'),            # amoebanet/__init__.py:54
        ((3, 64, 3, 1), 'stem'),            # unet/__init__.py:44, first cell
        ((64, 1, 1, 1), 'head'),            # unet/__init__.py:138
        ((256, 512, 1, 2), 'tc'),           # resnet/__init__.py:41
        ((128, 128, 3, 2), 'tc'),           # resnet/bottleneck.py:22,28
        ((64, 64, (1, 7), (1, 2)), 'tc'),   # amoebanet/operations.py:71-84
        ((64, 64, (7, 1), (2, 1)), 'tc'),
        ((64, 64, 3, 1), 'tc'),
    ]
    for (c_in, c_out, k, s), want in cases:
        assert TF.conv_kind(c_in, c_out, k, s, 1, 1) == want, (c_in, c_out, k, s)
    assert TF.conv_kind(64, 64, 3, 1, 2, 1) is None          # dilation
    assert TF.conv_kind(64, 64, 3, 1, 1, 2) is None          # groups
    assert TF.conv_kind(64, 64, 3, 4, 1, 1) is None          # stride 4x4
    assert TF.conv_kind(6, 64, 3, 1, 1, 1) is None           # 6 input channels
    assert TF.conv_kind(3, 64, (3, 9), 1, 1, 1) is None      # filter row longer than a K block


def test_amoebanet_conversion_fuses_cells_and_keeps_keys_and_numbers_on_cpu():
    import copy
    import torch
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import amoebanetd
    from torchgpipe_b200.ops.modules import AmoebaCell, AmoebaFactorizedReduce, AmoebaStem
    torch.manual_seed(0)
    # float64: with batch 2 and BatchNorm on 7x7 maps, fp32 summation order alone moves gradients
    # by 1 % - in double the fused and the plain model agree to 1e-13
    plain = amoebanetd(num_classes=10, num_layers=3, num_filters=64).double()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    assert len(fused) == len(plain)
    assert list(fused.state_dict().keys()) == list(plain.state_dict().keys())
    assert isinstance(fused[0], AmoebaStem)
    assert all(isinstance(layer, AmoebaCell) for layer in list(fused)[1:-1])
    assert any(isinstance(m, AmoebaFactorizedReduce) for m in fused.modules())
    # on CPU the fused modules take their ATen paths: shared ReLU and folded sums must not change
    # the numbers (outputs, gradients, running statistics)
    x = torch.randn(2, 3, 224, 224).double()
    a, b = plain(x), fused(x)
    assert torch.allclose(a, b, atol=1e-10)
    a.square().mean().backward()
    b.square().mean().backward()
    for (n, p), (_, q) in zip(plain.named_parameters(), fused.named_parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=1e-8, atol=1e-12), n
    for (n, p), (_, q) in zip(plain.named_buffers(), fused.named_buffers()):
        assert torch.allclose(p.double(), q.double(), atol=1e-10), n
    # converting twice is the identity
    assert ops.convert_to_sm100(fused)[1] is fused[1]


def test_graph_capture_key_separates_training_from_evaluation():
    import torch
    from torch import nn
    from torchgpipe_b200.runtime.graphs import _signature
    m = nn.Sequential(nn.Dropout(0.5))
    x = (torch.zeros(2, 3),)
    m.train()
    with torch.enable_grad():
        train_key = _signature(x, m)
    m.eval()
    with torch.no_grad():
        eval_key = _signature(x, m)
    with torch.enable_grad():
        eval_grad_key = _signature(x, m)
    assert len({train_key, eval_key, eval_grad_key}) == 3
    assert _signature(x, m, grad=True) == eval_grad_key


def test_link_bandwidth_comes_from_the_committed_measurement(monkeypatch):
    from torchgpipe_b200.utils import link_bandwidth_gbs
    monkeypatch.delenv('TGB_LINK_GBS', raising=False)
    assert 600.0 < link_bandwidth_gbs() < 900.0          # profiles/p2p_bandwidth.json (775.2)
    monkeypatch.setenv('TGB_LINK_GBS', '123.5')
    assert link_bandwidth_gbs() == 123.5


def test_absorbed_batchnorm_deferred_statistics_on_cpu():
    """The fused layer's BatchNorm under DeferredBatchNorm semantics (ATen path of the same
    bookkeeping the sm_100a path uses): running estimates move once per mini-batch and equal the
    full-batch ones."""
    import copy
    from torch import nn
    from torchgpipe_b200 import ops
    from torchgpipe_b200.batchnorm import DeferredBatchNorm, MiniBatchMoments
    torch.manual_seed(23)
    plain = nn.Sequential(nn.Conv2d(32, 64, 3, padding=1, bias=False), nn.BatchNorm2d(64))
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    fused = DeferredBatchNorm.convert_deferred_batch_norm(fused, chunks=2)
    bn = [m for m in fused.modules() if isinstance(m, ops.modules.AbsorbedBatchNorm2d)][0]
    assert bn.deferred_chunks == 2 and isinstance(bn.moments, MiniBatchMoments)
    assert {'sum', 'sum_squares'} <= set(dict(bn.named_buffers()))
    x = torch.randn(8, 32, 12, 12) * 3
    plain(x)
    first, second = x.chunk(2)
    before = bn.running_mean.clone()
    fused(first)
    assert bn.tracked == 1 and bn.counter == 4 * 12 * 12 and torch.equal(bn.running_mean, before)
    fused(second)
    assert int(bn.num_batches_tracked) == 1 and bn.tracked == 0 and bn.counter == 0
    assert torch.allclose(bn.running_mean, plain[1].running_mean, atol=1e-4)
    assert torch.allclose(bn.running_var, plain[1].running_var, rtol=2e-2, atol=1e-3)
    clone = copy.deepcopy(fused)
    cbn = [m for m in clone.modules() if isinstance(m, ops.modules.AbsorbedBatchNorm2d)][0]
    assert cbn.moments.owner is cbn
