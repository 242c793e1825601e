"""Numerics of the hand-written sm_100a kernels against plain PyTorch fp32 references."""
import copy

import pytest
import torch
from torch import nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CL = torch.channels_last


@pytest.fixture(autouse=True)
def full_fp32_reference():
    """Reference ops in true fp32 (no TF32), so the only TF32 rounding is in the kernel under test."""
    old_c, old_m = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old_c, old_m


def rel_err(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def test_native_library_is_loaded():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.ops import _lib
    assert ops.available(), _lib.load_error()


CONV_CASES = [
    # N, C, H, W, K, R, S, ph, pw
    (2, 32, 8, 16, 32, 1, 1, 0, 0),
    (2, 64, 16, 16, 64, 3, 3, 1, 1),
    (3, 96, 12, 12, 128, 3, 3, 1, 1),
    (5, 32, 6, 6, 256, 3, 3, 1, 1),
    (2, 32, 10, 20, 32, 3, 3, 1, 1),
    (2, 32, 8, 24, 64, 1, 7, 0, 3),
    (2, 32, 24, 8, 64, 7, 1, 3, 0),
    (1, 512, 8, 8, 256, 3, 3, 1, 1),
    (4, 64, 48, 48, 64, 3, 3, 1, 1),
    (2, 128, 24, 24, 32, 3, 3, 1, 1),
    (2, 32, 8, 8, 40, 1, 1, 0, 0),          # output channels not a multiple of the tile
    (1, 64, 8, 8, 72, 3, 3, 1, 1),
]


@pytest.mark.parametrize('case', CONV_CASES)
def test_conv2d_forward_backward_vs_fp32(case):
    from torchgpipe_b200.ops import functional as TF
    N, C, H, W, K, R, S, ph, pw = case
    torch.manual_seed(1)
    x = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL).requires_grad_()
    w = (torch.randn(K, C, R, S, device='cuda') / (C * R * S) ** 0.5).contiguous(
        memory_format=CL).requires_grad_()
    y, _ = TF.conv2d(x, w, (ph, pw))
    ref = F.conv2d(x, w, padding=(ph, pw))
    assert y.shape == ref.shape and y.is_contiguous(memory_format=CL)
    assert rel_err(y, ref) < 2e-3
    g = torch.randn_like(ref)
    dx, dw = torch.autograd.grad(y, (x, w), g)
    rdx, rdw = torch.autograd.grad(ref, (x, w), g)
    assert rel_err(dx, rdx) < 2e-3
    assert rel_err(dw, rdw) < 2e-3


# Shapes that reach the kernel variants the benchmark actually runs (the small cases above all
# land in the plain <32/64/128, 0, 0> kernels): haloed tiles (needs >= 4 x 148 tiles of 8 x 16
# pixels), resident filter bank, BN=256 with band tiles, forced split-K with the TMA reduce-add
# epilogue, and the tap-fused weight gradient with 128-wide tiles.
BIG_CONV_CASES = [
    # N, C, H, W, K, env
    (8, 64, 96, 96, 64, {}),                              # halo + resident weights, fused wgrad<64>
    (8, 64, 96, 96, 64, {'TGB_FPROP_RESW': '0'}),          # halo, streamed weights
    (8, 64, 96, 96, 64, {'TGB_FPROP_HALO': '0'}),          # same layer on the per-tap kernel
    (4, 128, 96, 96, 32, {}),                             # halo<32>, fused wgrad<128>
    (4, 32, 64, 64, 32, {'TGB_FPROP_HALO': '2'}),          # halo forced on a small grid
    (5, 1024, 12, 12, 1024, {}),                          # BN=256, band tiles (12x2 x 5 images)
    (5, 1024, 12, 12, 1024, {'TGB_FPROP_SPLITS': '3'}),    # split-K + TMA reduce-add
    (16, 256, 24, 24, 512, {}),                           # micro-batch 16 shape of the 8-GPU run
    (2, 2048, 6, 6, 2048, {}),                            # whole small images per tile
    # CTA pairs (tcgen05.mma.cta_group::2, M = 256): even / odd tile counts, several channel
    # tiles, split-K, statistics with partial tiles
    (5, 1024, 12, 12, 1024, {'TGB_FPROP_2CTA': '1'}),
    (16, 256, 24, 24, 512, {'TGB_FPROP_2CTA': '1'}),
    (3, 256, 20, 20, 256, {'TGB_FPROP_2CTA': '1'}),        # 3 x 4 = 12 tiles... odd rows
    (7, 512, 12, 12, 256, {'TGB_FPROP_2CTA': '1'}),        # odd number of pixel tiles
    (2, 2048, 6, 6, 2048, {'TGB_FPROP_2CTA': '1'}),
    (5, 1024, 12, 12, 1024, {'TGB_FPROP_2CTA': '1', 'TGB_FPROP_SPLITS': '3'}),
    (40, 256, 48, 48, 256, {'TGB_FPROP_2CTA': '1'}),       # many tiles per CTA pair
]


@pytest.mark.parametrize('case', BIG_CONV_CASES)
def test_conv2d_benchmark_variants_vs_fp32(case, monkeypatch):
    from torchgpipe_b200.ops import functional as TF
    N, C, H, W, K, env = case
    for key, value in env.items():
        monkeypatch.setenv(key, value)
    torch.manual_seed(11)
    x = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL).requires_grad_()
    w = (torch.randn(K, C, 3, 3, device='cuda') / (C * 9) ** 0.5).contiguous(
        memory_format=CL).requires_grad_()
    for mode in (0, 1, 2):
        y, stats = TF.conv2d(x, w, (1, 1), stats_mode=mode)
        ref = F.conv2d(x, w, padding=1)
        assert rel_err(y, ref) < 2e-3
        if mode:
            dims = (2, 3) if mode == 1 else (0, 2, 3)
            yd = y.detach().double()
            assert rel_err(stats[..., 0].double(), yd.sum(dims)) < 1e-3 or \
                torch.allclose(stats[..., 0].double(), yd.sum(dims), atol=0.5)
            assert rel_err(stats[..., 1].double(), (yd * yd).sum(dims)) < 1e-4
    g = torch.randn_like(ref)
    dx, dw = torch.autograd.grad(y, (x, w), g)
    rdx, rdw = torch.autograd.grad(ref, (x, w), g)
    assert rel_err(dx, rdx) < 2e-3
    assert rel_err(dw, rdw) < 2e-3


STRIDED_CASES = [
    # N, C, H, W, K, (R, S), (sh, sw), (ph, pw) - every strided shape of the benchmark models:
    # resnet/__init__.py:41 (1x1/2 downsample), bottleneck.py:22,28 (3x3/2), amoebanet
    # __init__.py:28 / operations.py:71-84 (1x7, 7x1, 3x3, 1x1 with stride), operations.py:31-32
    (2, 64, 16, 16, 64, (3, 3), (2, 2), (1, 1)),
    (3, 32, 15, 17, 64, (3, 3), (2, 2), (1, 1)),          # odd extents: unequal parity classes
    (2, 256, 14, 14, 512, (1, 1), (2, 2), (0, 0)),
    (2, 64, 15, 15, 128, (1, 1), (2, 2), (0, 0)),
    (2, 32, 12, 24, 32, (1, 7), (1, 2), (0, 3)),
    (2, 32, 24, 12, 32, (7, 1), (2, 1), (3, 0)),
    (2, 16, 20, 20, 32, (7, 7), (2, 2), (3, 3)),
    (4, 128, 28, 28, 128, (3, 3), (2, 2), (1, 1)),
    (2, 64, 9, 9, 64, (3, 3), (2, 2), (0, 0)),            # no padding
    (2, 36, 16, 16, 40, (3, 3), (2, 2), (1, 1)),          # partial channel tiles
]


@pytest.mark.parametrize('case', STRIDED_CASES)
def test_strided_conv2d_forward_backward_vs_fp32(case):
    from torchgpipe_b200.ops import functional as TF
    N, C, H, W, K, (R, S), stride, pad = case
    torch.manual_seed(5)
    x = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL).requires_grad_()
    w = (torch.randn(K, C, R, S, device='cuda') / (C * R * S) ** 0.5).contiguous(
        memory_format=CL).requires_grad_()
    y, stats = TF.conv2d(x, w, pad, stats_mode=2, stride=stride)
    ref = F.conv2d(x, w, stride=stride, padding=pad)
    assert y.shape == ref.shape and y.is_contiguous(memory_format=CL)
    assert rel_err(y, ref) < 2e-3
    assert rel_err(stats[:, 1], (ref * ref).sum((0, 2, 3))) < 2e-3
    g = torch.randn_like(ref)
    dx, dw = torch.autograd.grad(y, (x, w), g)
    rdx, rdw = torch.autograd.grad(ref, (x, w), g)
    assert rel_err(dx, rdx) < 2e-3
    assert rel_err(dw, rdw) < 2e-3


@pytest.mark.parametrize('hw', [14, 15])
def test_factorized_reduce_convs_write_channel_halves_in_place(hw):
    """The two halves of AmoebaNet's FactorizedReduce (reference operations.py:24-40): conv1 on
    the even pixels, conv2 on the odd ones (shift by one, zero-pad the far edge = a 1x1 stride-2
    convolution with padding -1), both written straight into their half of one output."""
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(6)
    N, C, H, W, K = 3, 64, hw, hw, 96
    x = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL).requires_grad_()
    w1 = (torch.randn(K // 2, C, 1, 1, device='cuda') / 8).contiguous(memory_format=CL)
    w2 = (torch.randn(K // 2, C, 1, 1, device='cuda') / 8).contiguous(memory_format=CL)
    w1.requires_grad_(), w2.requires_grad_()
    out, stats = TF.factorized_reduce_conv(x, w1, w2, {}, {})
    shifted = F.pad(x[:, :, 1:, 1:], (0, 1, 0, 1))
    ref = torch.cat([F.conv2d(x, w1, stride=2), F.conv2d(shifted, w2, stride=2)], dim=1)
    assert out.shape == ref.shape and rel_err(out, ref) < 2e-3
    assert rel_err(stats[:, 1], (ref * ref).sum((0, 2, 3))) < 2e-3
    g = torch.randn_like(ref)
    got = torch.autograd.grad(out, (x, w1, w2), g)
    want = torch.autograd.grad(ref, (x, w1, w2), g)
    for a, b in zip(got, want):
        assert rel_err(a, b) < 2e-3


STEM_CASES = [
    # N, C, H, W, K, (R, S), stride, pad, channels_last input
    (2, 3, 64, 64, 64, (7, 7), (2, 2), (3, 3), False),     # resnet/__init__.py:55
    (2, 3, 56, 40, 64, (3, 3), (2, 2), (1, 1), False),     # amoebanet/__init__.py:54
    (2, 3, 48, 48, 64, (3, 3), (1, 1), (1, 1), True),      # unet/__init__.py:44, first cell
    (3, 1, 33, 31, 32, (5, 5), (1, 1), (2, 2), False),
    (2, 4, 30, 30, 32, (3, 8), (2, 1), (1, 0), True),
]


@pytest.mark.parametrize('case', STEM_CASES)
def test_small_channel_stem_vs_fp32(case):
    from torchgpipe_b200.ops import functional as TF
    N, C, H, W, K, (R, S), stride, pad, cl = case
    torch.manual_seed(8)
    x = torch.randn(N, C, H, W, device='cuda')
    if cl:
        x = x.contiguous(memory_format=CL)
    w = (torch.randn(K, C, R, S, device='cuda') / (C * R * S) ** 0.5).requires_grad_()
    cache = {}
    y, stats = TF.stem_conv2d(x, w, pad, stride, stats_mode=1, cache=cache)
    ref = F.conv2d(x, w, stride=stride, padding=pad)
    assert y.shape == ref.shape
    assert rel_err(y, ref) < 2e-3
    assert rel_err(stats[..., 0], ref.sum((2, 3))) < 5e-3 or \
        torch.allclose(stats[..., 0], ref.sum((2, 3)), atol=0.05)
    g = torch.randn_like(ref)
    dw, = torch.autograd.grad(y, w, g, retain_graph=True)
    rdw, = torch.autograd.grad(ref, w, g, retain_graph=True)
    assert rel_err(dw, rdw) < 2e-3
    # gradient with respect to the image (the eager pipeline asks for it at the first stage)
    xg = x.detach().clone().requires_grad_()
    yg, _ = TF.stem_conv2d(xg, w, pad, stride)
    dx, = torch.autograd.grad(yg, xg, g)
    xr = x.detach().clone().requires_grad_()
    rdx, = torch.autograd.grad(F.conv2d(xr, w, stride=stride, padding=pad), xr, g)
    assert rel_err(dx, rdx) < 2e-3
    # cached packed weights follow an in-place parameter update
    with torch.no_grad():
        w.mul_(0.5)
    TF.refresh_stem_weight(w, cache)
    y2, _ = TF.stem_conv2d(x, w, pad, stride, cache=cache)
    assert rel_err(y2, 0.5 * ref) < 2e-3


@pytest.mark.parametrize('C,K,bias', [(64, 1, True), (32, 2, False), (256, 4, True), (8, 3, True)])
def test_narrow_head_conv_vs_fp32(C, K, bias):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(9)
    x = torch.randn(3, C, 19, 23, device='cuda').contiguous(memory_format=CL).requires_grad_()
    w = (torch.randn(K, C, 1, 1, device='cuda') / C ** 0.5).requires_grad_()
    b = torch.randn(K, device='cuda').requires_grad_() if bias else None
    y = TF.head_conv2d(x, w, b)
    ref = F.conv2d(x, w, b)
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-5)      # plain fp32 arithmetic
    g = torch.randn_like(ref)
    ins = (x, w) + ((b,) if bias else ())
    got = torch.autograd.grad(y, ins, g)
    want = torch.autograd.grad(ref, ins, g)
    for a, r in zip(got, want):
        assert rel_err(a, r) < 1e-4


@pytest.mark.parametrize('mode', [1, 2])
def test_conv_epilogue_statistics(mode):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(2)
    x = torch.randn(6, 64, 24, 24, device='cuda').contiguous(memory_format=CL)
    w = (torch.randn(128, 64, 3, 3, device='cuda') / 24).contiguous(memory_format=CL)
    y, stats = TF.conv2d_forward(x, w, (1, 1), stats_mode=mode)
    dims = (2, 3) if mode == 1 else (0, 2, 3)
    s1, s2 = y.sum(dims), (y * y).sum(dims)
    assert torch.allclose(stats[..., 0], s1, rtol=1e-4, atol=1e-2)
    assert torch.allclose(stats[..., 1], s2, rtol=1e-4, atol=1e-2)


def test_channel_moments():
    from torchgpipe_b200.ops import functional as TF
    x = torch.randn(5, 36, 17, 13, device='cuda').contiguous(memory_format=CL)
    m = TF.channel_moments(x)
    assert torch.allclose(m[:, 0], x.sum((0, 2, 3)), rtol=1e-4, atol=1e-3)
    assert torch.allclose(m[:, 1], (x * x).sum((0, 2, 3)), rtol=1e-4, atol=1e-3)
    mi = TF.channel_moments(x, per_image=True)
    assert torch.allclose(mi[..., 0], x.sum((2, 3)), rtol=1e-4, atol=1e-3)
    big = torch.randn(2, 2048, 6, 6, device='cuda').contiguous(memory_format=CL)  # C/4 > block
    mb = TF.channel_moments(big, per_image=True)
    assert torch.allclose(mb[..., 1], (big * big).sum((2, 3)), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize('C,H,W', [(64, 24, 24), (32, 9, 7), (2048, 6, 6)])
@pytest.mark.parametrize('dropout', [False, True])
def test_instnorm_dropout_leaky_fwd_bwd(C, H, W, dropout):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(3)
    N, p, eps, slope = 4, 0.25, 1e-5, 0.01
    y = (torch.randn(N, C, H, W, device='cuda') * 3 + 1).contiguous(memory_format=CL)
    y.requires_grad_()
    scale = None
    if dropout:
        scale = (torch.rand(N, C, device='cuda') >= p).float() / (1 - p)
    z = TF.instnorm_leaky(y, None, scale, eps, slope)

    yr = y.detach().clone().requires_grad_()
    d = yr * scale.view(N, C, 1, 1) if dropout else yr
    zr = F.leaky_relu(F.instance_norm(d, eps=eps), slope)
    assert torch.allclose(z, zr, rtol=1e-4, atol=1e-4)

    g = torch.randn_like(zr)
    dy, = torch.autograd.grad(z, y, g)
    dyr, = torch.autograd.grad(zr, yr, g)
    assert rel_err(dy, dyr) < 1e-3


def test_fused_unet_cell_matches_unfused_in_eval_and_train_without_dropout():
    from torchgpipe_b200.ops.modules import ConvInstNormLeaky
    torch.manual_seed(4)
    conv = nn.Conv2d(64, 128, 3, padding=1, bias=False).cuda()
    ref = nn.Sequential(copy.deepcopy(conv), nn.Dropout2d(0.0), nn.InstanceNorm2d(128),
                        nn.LeakyReLU(0.01)).cuda()
    fused = ConvInstNormLeaky(conv, 0.0, 1e-5, 0.01).cuda()
    x = torch.randn(3, 64, 20, 20, device='cuda', requires_grad=True)
    xr = x.detach().clone().requires_grad_()
    z, zr = fused(x), ref(xr)
    assert rel_err(z, zr) < 3e-3
    g = torch.randn_like(zr)
    z.backward(g)
    zr.backward(g)
    assert rel_err(x.grad, xr.grad) < 2e-2
    assert rel_err(fused.weight.grad, ref[0].weight.grad) < 2e-2


def test_stem_with_three_input_channels_runs_on_the_packed_window_path():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.ops.modules import ConvInstNormLeaky
    torch.manual_seed(5)
    conv = nn.Conv2d(3, 64, 3, padding=1, bias=False).cuda()
    ref = nn.Sequential(copy.deepcopy(conv), nn.InstanceNorm2d(64), nn.LeakyReLU(0.01)).cuda()
    fused = ConvInstNormLeaky(conv, 0.0, 1e-5, 0.01).cuda()
    x = torch.rand(2, 3, 32, 32, device='cuda')
    before = ops.launch_count()
    assert rel_err(fused(x), ref(x)) < 3e-3
    assert ops.launch_count() - before >= 4      # pack, conv, prepare, normalise: no library conv


def _tf32_bound_check(fused, plain, x, slack=3.0, floor=2e-3):
    """Compares ``fused`` with a float64 oracle of ``plain``. Stated bound: the fused model may be
    at most ``slack`` times as far from the oracle as the same plain model run by PyTorch with
    TF32 convolutions (the reference's arithmetic), plus ``floor``. That measures the kernels'
    rounding against the rounding the reference itself has, however ill-conditioned the net."""
    oracle = copy.deepcopy(plain).double()
    xd = x.double()
    ref = oracle(xd)
    ref.mean().backward()
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    base = plain(x)
    base.mean().backward()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    out = fused(x)
    out.mean().backward()
    assert rel_err(out.double(), ref) <= slack * rel_err(base.double(), ref) + floor
    worst = 0.0
    for (n, p), (_, q), (_, o) in zip(fused.named_parameters(), plain.named_parameters(),
                                      oracle.named_parameters()):
        if o.grad is None:
            continue
        e_fused, e_base = rel_err(p.grad.double(), o.grad), rel_err(q.grad.double(), o.grad)
        assert e_fused <= slack * e_base + floor, (n, e_fused, e_base)
        worst = max(worst, e_fused)
    return worst


def test_converted_unet_matches_plain_unet():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import unet
    torch.manual_seed(6)
    # eval mode: no dropout; InstanceNorm uses batch statistics in both modes
    plain = unet(depth=2, num_convs=2, base_channels=32).cuda().eval()
    fused = ops.convert_to_sm100(copy.deepcopy(plain)).cuda().eval()
    assert len(fused) == len(plain)
    assert list(fused.state_dict().keys()) == list(plain.state_dict().keys())
    x = torch.rand(2, 3, 32, 32, device='cuda')
    before = ops.launch_count()
    _tf32_bound_check(fused, plain, x)
    assert ops.launch_count() > before


def test_dropout_mask_is_reproduced_under_checkpointing():
    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models import unet

    def grads(checkpoint):
        torch.manual_seed(7)
        model = ops.convert_to_sm100(unet(depth=2, num_convs=2, base_channels=32))
        pipe = GPipe(model, [len(model)], devices=['cuda:0'], chunks=2, checkpoint=checkpoint)
        torch.manual_seed(8)
        x = torch.rand(4, 3, 32, 32, device='cuda')
        torch.cuda.manual_seed(9)
        out = pipe(x)
        out.mean().backward()
        return out.detach(), [p.grad.clone() for p in pipe.parameters()]

    out_a, g_a = grads('always')
    out_n, g_n = grads('never')
    assert rel_err(out_a, out_n) < 1e-2   # TF32 input rounding amplifies atomics-order noise
    for a, b in zip(g_a, g_n):
        assert F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item() > 0.99


def test_deferred_batch_norm_uses_fused_moments_on_channels_last():
    from torchgpipe_b200.batchnorm import DeferredBatchNorm
    bn = nn.BatchNorm2d(16).cuda()
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(copy.deepcopy(bn), chunks=2).cuda()
    x = (torch.randn(8, 16, 9, 9, device='cuda') * 2 + 3).contiguous(memory_format=CL)
    bn(x)
    for c in x.chunk(2):
        dbn(c)
    assert torch.allclose(bn.running_mean, dbn.running_mean, atol=1e-4)
    assert torch.allclose(bn.running_var, dbn.running_var, atol=1e-3)


def test_graft_smoke():
    import __graft_entry__ as entry
    entry.smoke()


def _single_rank_grads(cuda_graphs, seed=11):
    from torchgpipe_b200 import ops
    from torchgpipe_b200.distributed import DistributedGPipe, context
    from torchgpipe_b200.models import unet
    from torchgpipe_b200.ops.modules import ConvInstNormLeaky
    torch.manual_seed(seed)
    model = ops.convert_to_sm100(unet(depth=2, num_convs=2, base_channels=32))
    for m in model.modules():
        if isinstance(m, ConvInstNormLeaky):
            m.p = 0.0           # eager and graph-captured RNG streams differ: compare without dropout
    name = f'solo-{cuda_graphs}-{seed}'
    with context.worker(name, 3):
        pipe = DistributedGPipe(model, 0, {0: name}, [len(model)], 3, device=torch.device('cuda'),
                                checkpoint='always', cuda_graphs=cuda_graphs)
        torch.manual_seed(seed + 1)
        x = torch.rand(6, 3, 32, 32, device='cuda')
        results = []
        opt = torch.optim.SGD(pipe.model().parameters(), lr=0.05)
        for step in range(3):               # later steps replay the captured graphs
            outs = pipe.forward(x)
            losses = [o.square().mean() for o in outs]
            pipe.backward(losses)
            results.append((torch.stack([l.detach() for l in losses]),
                            [p.grad.clone() for p in pipe.model().parameters()]))
            # derived (dgrad) weights must have followed the previous optimiser step
            for m in pipe.model().modules():
                if isinstance(m, ConvInstNormLeaky) and 'wt' in m._cache:
                    want = m.weight.detach().flip(2, 3).permute(1, 0, 2, 3)
                    assert torch.equal(m._cache['wt'], want), 'stale dgrad weights'
            opt.step()
            opt.zero_grad()
    return results, pipe


def test_cuda_graph_stage_matches_eager_stage():
    eager, _ = _single_rank_grads(False)
    graphed, pipe = _single_rank_grads(True)
    assert pipe.native_launches_replayed() > 0
    # 3 steps x 3 micro-batches: the last micro-batch of every step skips its recomputation
    assert sum(st.recomputes_skipped for st in pipe._graphed.values()) == 3
    for step, ((le, ge), (lg, gg)) in enumerate(zip(eager, graphed)):
        assert rel_err(lg, le) < 1e-2           # same loss trajectory over optimiser steps
        if step == 0:                           # later steps: TF32 noise makes weights diverge
            for a, b in zip(gg, ge):
                assert F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item() > 0.995


def test_cuda_graph_recompute_reproduces_dropout_mask():
    from torchgpipe_b200.ops.modules import ConvInstNormLeaky
    from torchgpipe_b200.runtime.graphs import GraphedStage
    torch.manual_seed(12)
    cell = ConvInstNormLeaky(nn.Conv2d(32, 64, 3, padding=1, bias=False), 0.5, 1e-5, 0.01).cuda()
    staged = GraphedStage(lambda t: (cell(t[0]),), cell, torch.device('cuda'))
    x = torch.randn(4, 32, 16, 16, device='cuda').contiguous(memory_format=CL).requires_grad_()
    x2 = torch.randn_like(x).requires_grad_()
    masks = []
    for step in range(2):
        (out,), rng = staged.forward((x,))
        (out2,), rng2 = staged.forward((x2,))
        # the latest forward is still resident: its backward needs no recomputation ...
        gin2, = staged.backward((x2,), rng2, (torch.ones_like(out2),))
        assert staged.recomputes_skipped == step + 1
        kept = next(iter(staged._captured.values())).static_outputs[0]
        assert torch.equal(kept, out2)
        # ... the earlier one is recomputed and must draw the same dropout mask again
        gin, = staged.backward((x,), rng, (torch.ones_like(out),))
        assert staged.recomputes_skipped == step + 1
        recomputed = next(iter(staged._captured.values())).static_outputs[0]
        # dropped channels are exactly zero after the norm: same channels in both passes
        assert torch.equal(out == 0, recomputed == 0)
        assert rel_err(recomputed, out) < 1e-3
        assert torch.isfinite(gin).all() and torch.isfinite(gin2).all()
        masks.append((out == 0).all(dim=3).all(dim=2))
    assert masks[0].any() and not torch.equal(masks[0], masks[1])   # a fresh mask every step


@pytest.mark.parametrize('relu', [False, True])
@pytest.mark.parametrize('C,H,W', [(64, 14, 14), (36, 7, 5), (2048, 4, 4)])
def test_batchnorm_act_fwd_bwd(C, H, W, relu):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(21)
    N = 6
    y = (torch.randn(N, C, H, W, device='cuda') * 2 + 0.5).contiguous(memory_format=CL)
    y.requires_grad_()
    bn = nn.BatchNorm2d(C).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    w = bn.weight.detach().clone().requires_grad_()
    b = bn.bias.detach().clone().requires_grad_()
    z = TF.batchnorm_act(y, None, w, b, rm, rv, bn.eps, bn.momentum, relu)

    yr = y.detach().clone().requires_grad_()
    zr = bn(yr)
    zr = F.relu(zr) if relu else zr
    assert torch.allclose(z, zr, rtol=1e-4, atol=1e-4)
    assert torch.allclose(rm, bn.running_mean, atol=1e-5)
    assert torch.allclose(rv, bn.running_var, rtol=1e-4, atol=1e-5)

    g = torch.randn_like(zr)
    dy, dw, db = torch.autograd.grad(z, (y, w, b), g)
    dyr, dwr, dbr = torch.autograd.grad(zr, (yr, bn.weight, bn.bias), g)
    assert rel_err(dy, dyr) < 1e-3
    assert rel_err(dw, dwr) < 1e-3 and rel_err(db, dbr) < 1e-3


@pytest.mark.parametrize('stride', [1, 2])
def test_conv_bn_relu_module_matches_unfused(stride):
    from torchgpipe_b200 import ops
    torch.manual_seed(22)
    plain = nn.Sequential(nn.Conv2d(64, 128, 3, stride=stride, padding=1, bias=False),
                          nn.BatchNorm2d(128), nn.ReLU()).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    assert type(fused[0]).__name__ == 'ConvBatchNormAct'
    assert list(fused.state_dict().keys()) == list(plain.state_dict().keys())
    x = torch.randn(4, 64, 16, 16, device='cuda', requires_grad=True)
    xr = x.detach().clone().requires_grad_()
    z, zr = fused(x), plain(xr)
    assert rel_err(z, zr) < 3e-3
    assert torch.allclose(fused[1].running_mean, plain[1].running_mean, atol=1e-3)
    assert torch.allclose(fused[1].running_var, plain[1].running_var, rtol=1e-2, atol=1e-3)
    assert int(fused[1].num_batches_tracked) == 1
    g = torch.randn_like(zr)
    z.backward(g)
    zr.backward(g)
    assert rel_err(x.grad, xr.grad) < 2e-2
    assert rel_err(fused[0].weight.grad, plain[0].weight.grad) < 2e-2
    assert rel_err(fused[1].weight.grad, plain[1].weight.grad) < 2e-2
    fused.eval()
    plain.eval()
    assert rel_err(fused(x), plain(xr)) < 3e-3


def test_absorbed_batchnorm_deferred_statistics():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.batchnorm import DeferredBatchNorm
    torch.manual_seed(23)
    plain = nn.Sequential(nn.Conv2d(32, 64, 3, padding=1, bias=False), nn.BatchNorm2d(64)).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    fused = DeferredBatchNorm.convert_deferred_batch_norm(fused, chunks=2)
    assert fused[1].deferred_chunks == 2 and 'sum' in dict(fused[1].named_buffers())
    x = torch.randn(8, 32, 12, 12, device='cuda') * 3
    plain(x)                                    # full mini-batch statistics
    for c in x.chunk(2):
        fused(c)
    assert int(fused[1].num_batches_tracked) == 1 and fused[1].tracked == 0
    assert torch.allclose(fused[1].running_mean, plain[1].running_mean, atol=1e-3)
    assert torch.allclose(fused[1].running_var, plain[1].running_var, rtol=2e-2, atol=1e-3)


def test_converted_small_resnet_matches_plain():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models.resnet import build_resnet
    torch.manual_seed(24)
    plain = build_resnet([1, 1, 1, 1], num_classes=10).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    assert len(fused) == len(plain)
    x = torch.randn(4, 3, 64, 64, device='cuda')
    _tf32_bound_check(fused, plain, x, slack=4.0, floor=5e-3)


def test_converted_small_amoebanet_matches_plain():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import amoebanetd
    torch.manual_seed(25)
    plain = amoebanetd(num_classes=10, num_layers=3, num_filters=64).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    assert list(fused.state_dict().keys()) == list(plain.state_dict().keys())
    x = torch.randn(4, 3, 224, 224, device='cuda')
    _tf32_bound_check(fused, plain, x, slack=4.0, floor=5e-3)


def _kernel_names(fn):
    from torch.profiler import ProfilerActivity, profile
    fn()                                    # warm-up: lazy initialisation launches
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    return [e.key for e in prof.key_averages()]


@pytest.mark.parametrize('which', ['unet', 'resnet', 'amoebanet'])
def test_converted_models_launch_no_library_convolution(which):
    """Stems, strided layers and the 1-channel head included: a training step of a converted
    benchmark model contains no cuDNN / cuBLAS convolution or GEMM kernel."""
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import amoebanetd, unet
    from torchgpipe_b200.models.resnet import build_resnet
    torch.manual_seed(26)
    if which == 'unet':
        model, hw = unet(depth=2, num_convs=2, base_channels=32), 64
    elif which == 'resnet':
        model, hw = build_resnet([1, 1, 1, 1], num_classes=12), 64
    else:
        model, hw = amoebanetd(num_classes=12, num_layers=3, num_filters=64), 224
    model = ops.convert_to_sm100(model).cuda()
    x = torch.randn(4, 3, hw, hw, device='cuda')

    def step():
        model.zero_grad()
        model(x).square().mean().backward()

    names = _kernel_names(step)
    library = [n for n in names if any(tag in n.lower() for tag in (
        'cudnn', 'cutlass', 'xmma', 'implicit_convolve', 'wgrad', 'dgrad', 'gemm', 'cublas',
        'winograd', 'conv2d_grouped', 'nchwtonhwc', 'nhwctonchw')) and 'tgb::' not in n
        and not n.startswith('tgb') and 'dgrad_weight' not in n]
    assert not library, library
    assert any('conv_fprop_tf32_kernel' in n for n in names)


@pytest.mark.multigpu
def test_native_kernels_on_second_device_in_one_process():
    """Single-process GPipe drives several GPUs: per-device kernel attributes must be set on each."""
    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models import unet
    torch.manual_seed(30)
    model = ops.convert_to_sm100(unet(depth=2, num_convs=2, base_channels=64))
    n = len(model)
    pipe = GPipe(model, [n // 2, n - n // 2], devices=['cuda:0', 'cuda:1'], chunks=2)
    out = pipe(torch.rand(4, 3, 32, 32, device='cuda:0'))
    assert out.device == torch.device('cuda', 1)
    out.mean().backward()
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    assert all(torch.isfinite(p.grad).all() for p in pipe.parameters())


@pytest.mark.parametrize('shape', [(3, 64, 16, 16), (2, 36, 9, 7), (1, 8, 2, 2)])
def test_maxpool2x2_fwd_bwd(shape):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(31)
    x = torch.randn(*shape, device='cuda').contiguous(memory_format=CL).requires_grad_()
    xr = x.detach().clone().requires_grad_()
    y, yr = TF.maxpool2x2(x), F.max_pool2d(xr, 2, 2)
    assert torch.equal(y, yr)
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    assert torch.equal(x.grad, xr.grad)


@pytest.mark.parametrize('hx,wx,hs,ws', [(8, 8, 16, 16), (5, 6, 11, 13)])
def test_upsample2x_cat_fwd_bwd(hx, wx, hs, ws):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(32)
    x = torch.randn(3, 32, hx, wx, device='cuda').contiguous(memory_format=CL).requires_grad_()
    s = torch.randn(3, 16, hs, ws, device='cuda').contiguous(memory_format=CL).requires_grad_()
    xr, sr = x.detach().clone().requires_grad_(), s.detach().clone().requires_grad_()
    out = TF.upsample2x_cat(x, s)
    up = F.interpolate(xr, scale_factor=2.0, mode='nearest')
    up = F.pad(up, [0, ws - up.size(3), 0, hs - up.size(2)])
    ref = torch.cat((up, sr), dim=1)
    assert torch.equal(out, ref)
    g = torch.randn_like(ref)
    out.backward(g)
    ref.backward(g)
    assert torch.allclose(x.grad, xr.grad, atol=1e-6) and torch.equal(s.grad, sr.grad)


def test_converted_unet_uses_fused_pool_and_upcat_across_stages():
    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models import unet
    torch.manual_seed(33)
    plain = unet(depth=2, num_convs=1, base_channels=32).cuda().eval()
    fused = ops.convert_to_sm100(copy.deepcopy(plain)).eval()
    names = [type(m).__name__ for m in fused]
    assert 'MaxPool2x2' in names and 'UpsamplePopCat' in names
    x = torch.rand(2, 3, 32, 32, device='cuda')
    ref = plain(x)
    n = len(fused)
    pipe = GPipe(fused, [n // 2, n - n // 2], devices=['cuda:0', 'cuda:0'], chunks=2).eval()
    assert rel_err(pipe(x), ref) < 1e-2        # skip portals still route to the fused PopCat


def test_linear_on_conv_kernels():
    from torchgpipe_b200 import ops
    torch.manual_seed(40)
    plain = nn.Linear(2048, 1000).cuda()
    fused = ops.convert_to_sm100(nn.Sequential(copy.deepcopy(plain)))[0]
    assert type(fused).__name__ == 'LinearTC'
    x = torch.randn(40, 2048, device='cuda', requires_grad=True)
    xr = x.detach().clone().requires_grad_()
    y, yr = fused(x), plain(xr)
    assert rel_err(y, yr) < 2e-3
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    assert rel_err(x.grad, xr.grad) < 2e-3
    assert rel_err(fused.weight.grad, plain.weight.grad) < 2e-3
    assert rel_err(fused.bias.grad, plain.bias.grad) < 1e-5


POOL_CASES = [(3, 2, 1, 'max'), (3, 1, 1, 'max'), (2, 2, 0, 'max'), (2, 1, 0, 'max'),
              (3, 1, 1, 'avg'), (3, 2, 1, 'avg'), (3, 1, 1, 'avg_include_pad'), (2, 2, 0, 'avg')]


@pytest.mark.parametrize('k,s,p,mode', POOL_CASES)
@pytest.mark.parametrize('shape', [(3, 64, 16, 16), (2, 36, 9, 7)])
def test_pool2d_family_fwd_bwd(k, s, p, mode, shape):
    # Oracle: PyTorch fp32 on the CPU. (ATen's CUDA channels-last avg_pool2d backward with
    # count_include_pad=False disagrees with its own CPU kernel in this build; the kernel under
    # test matches the CPU result and the closed form conv_transpose(g / count).)
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(50)
    xr = torch.randn(*shape, requires_grad=True)
    x = xr.detach().cuda().contiguous(memory_format=CL).requires_grad_()
    y = TF.pool2d(x, k, s, p, mode)
    if mode == 'max':
        yr = F.max_pool2d(xr, k, s, p)
    else:
        yr = F.avg_pool2d(xr, k, s, p, count_include_pad=mode == 'avg_include_pad')
    assert y.shape == yr.shape
    assert torch.allclose(y.cpu(), yr, rtol=1e-6, atol=1e-6)
    g = torch.randn_like(yr)
    y.backward(g.cuda())
    yr.backward(g)
    assert torch.allclose(x.grad.cpu(), xr.grad, rtol=1e-5, atol=1e-5)


def test_max_pool_backward_routes_ties_like_aten():
    from torchgpipe_b200.ops import functional as TF
    x = torch.zeros(1, 4, 6, 6, device='cuda').contiguous(memory_format=CL).requires_grad_()
    xr = torch.zeros(1, 4, 6, 6, requires_grad=True)
    g = torch.arange(1., 37.).view(1, 1, 6, 6).expand(1, 4, 6, 6).contiguous()
    TF.pool2d(x, 3, 1, 1, 'max').backward(g.cuda())
    F.max_pool2d(xr, 3, 1, 1).backward(g)
    assert torch.equal(x.grad.cpu(), xr.grad)


@pytest.mark.parametrize('shape', [(5, 2048, 7, 7), (3, 36, 5, 3), (2, 100, 1, 1)])
def test_global_avgpool_fwd_bwd(shape):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(51)
    x = torch.randn(*shape, device='cuda').contiguous(memory_format=CL).requires_grad_()
    xr = x.detach().clone().requires_grad_()
    y, yr = TF.global_avgpool(x), F.adaptive_avg_pool2d(xr, 1)
    assert y.shape == yr.shape and torch.allclose(y, yr, rtol=1e-5, atol=1e-6)
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    assert torch.allclose(x.grad, xr.grad, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize('relu', [False, True])
def test_add_act_fwd_bwd(relu):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(52)
    a = torch.randn(3, 36, 9, 7, device='cuda').contiguous(memory_format=CL).requires_grad_()
    b = torch.randn(3, 36, 9, 7, device='cuda').contiguous(memory_format=CL).requires_grad_()
    ar, br = a.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
    z = TF.add_act(a, b, relu)
    zr = F.relu(ar + br) if relu else ar + br
    assert torch.equal(z, zr)
    g = torch.randn_like(zr)
    z.backward(g)
    zr.backward(g)
    assert torch.equal(a.grad, ar.grad) and torch.equal(b.grad, br.grad)


@pytest.mark.parametrize('relu', [False, True])
@pytest.mark.parametrize('C,H,W', [(256, 14, 14), (36, 7, 5), (2048, 4, 4)])
def test_batchnorm_add_act_fwd_bwd(C, H, W, relu):
    from torchgpipe_b200.ops import functional as TF
    torch.manual_seed(53)
    N = 6
    y = (torch.randn(N, C, H, W, device='cuda') * 2 + 0.5).contiguous(memory_format=CL)
    r = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL)
    y.requires_grad_()
    r.requires_grad_()
    bn = nn.BatchNorm2d(C).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    w = bn.weight.detach().clone().requires_grad_()
    b = bn.bias.detach().clone().requires_grad_()
    z = TF.batchnorm_add_act(y, None, w, b, r, rm, rv, bn.eps, bn.momentum, relu)

    yr, rr = y.detach().clone().requires_grad_(), r.detach().clone().requires_grad_()
    zr = bn(yr) + rr
    zr = F.relu(zr) if relu else zr
    assert torch.allclose(z, zr, rtol=1e-4, atol=1e-4)
    assert torch.allclose(rm, bn.running_mean, atol=1e-5)
    assert torch.allclose(rv, bn.running_var, rtol=1e-4, atol=1e-5)

    g = torch.randn_like(zr)
    dy, dr, dw, db = torch.autograd.grad(z, (y, r, w, b), g)
    dyr, drr, dwr, dbr = torch.autograd.grad(zr, (yr, rr, bn.weight, bn.bias), g)
    assert rel_err(dy, dyr) < 1e-3 and rel_err(dr, drr) < 1e-5
    assert rel_err(dw, dwr) < 1e-3 and rel_err(db, dbr) < 1e-3


def test_converted_bottleneck_tail_is_one_fused_join():
    """conv3, bn3, residual, relu3 -> deferring conv + BatchNormResidualAct; numerics vs plain."""
    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models.resnet import bottleneck
    torch.manual_seed(54)
    proj = nn.Sequential(nn.Conv2d(64, 256, 1, bias=False), nn.BatchNorm2d(256))
    plain = nn.Sequential(*bottleneck(64, 64, 1, proj).children(),
                          *bottleneck(256, 64).children()).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    names = [type(m).__name__ for m in fused]
    assert names.count('BatchNormResidualAct') == 2 and 'Residual' not in names
    assert list(fused.state_dict().keys()) == list(plain.state_dict().keys())
    x = torch.randn(8, 64, 14, 14, device='cuda')
    # through GPipe so that the skip route and checkpoint recomputation are exercised as well
    once = GPipe(copy.deepcopy(plain), [22], devices=['cuda:0'], chunks=2, checkpoint='never')
    once(x)
    gp = GPipe(plain, [22], devices=['cuda:0'], chunks=2, checkpoint='always')
    gf = GPipe(fused, [22], devices=['cuda:0'], chunks=2, checkpoint='always')
    out, ref = gf(x), gp(x)
    assert rel_err(out, ref) < 5e-3
    out.square().mean().backward()
    ref.square().mean().backward()
    for (n, p), (_, q) in zip(gf.named_parameters(), gp.named_parameters()):
        assert rel_err(p.grad, q.grad) < 8e-2, n       # TF32 in both, different rounding
    # Running statistics: the fused layers skip the update during recomputation, so they match
    # a plain run WITHOUT checkpointing (plain BatchNorm under checkpointing updates twice).
    for (n, p), (_, q) in zip(gf.named_buffers(), once.named_buffers()):
        assert torch.allclose(p.float(), q.float(), rtol=1e-3, atol=1e-3), n


@pytest.mark.parametrize('momentum', [0.1, None])
def test_fused_batchnorm_keeps_counter_and_cumulative_average_on_the_device(momentum):
    """num_batches_tracked is incremented inside the prepare kernel, and momentum=None (the
    cumulative moving average) is resolved from that device counter - no host read."""
    from torchgpipe_b200 import ops
    torch.manual_seed(31)
    plain = nn.Sequential(nn.Conv2d(8, 16, 3, padding=1, bias=False),
                          nn.BatchNorm2d(16, momentum=momentum), nn.ReLU()).cuda()
    fused = ops.convert_to_sm100(copy.deepcopy(plain))
    for step in range(3):
        x = torch.randn(4, 8, 10, 10, device='cuda') * (step + 1)
        a, b = fused(x), plain(x)
        assert rel_err(a, b) < 5e-3
    assert int(fused[1].num_batches_tracked) == 3 == int(plain[1].num_batches_tracked)
    assert torch.allclose(fused[1].running_mean, plain[1].running_mean, atol=2e-3)
    assert torch.allclose(fused[1].running_var, plain[1].running_var, rtol=1e-2, atol=2e-3)
