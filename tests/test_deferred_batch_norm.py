from copy import deepcopy
from itertools import chain

import pytest
import torch
from torch import nn, optim

from torchgpipe_b200.batchnorm import DeferredBatchNorm
from torchgpipe_b200.checkpoint import checkpoint

CHUNKS = 4


def tilt_dist(input):
    # make per-chunk statistics differ from the full batch's
    rgb = input.transpose(0, 1)
    rgb[0] *= 1
    rgb[1] *= 10
    rgb[2] *= 100
    means = torch.tensor([[[[2.]]], [[[5.]]], [[[7.]]]])
    input = input + means.transpose(0, 1) * torch.arange(input.size(0)).view(-1, 1, 1, 1)
    return input


def chunked_forward(model, input, chunks=CHUNKS):
    return torch.cat([model(c) for c in input.chunk(chunks)])


@pytest.mark.parametrize('chunks', [1, 4])
@pytest.mark.parametrize('input_requires_grad', [True, False])
def test_output_and_grad_equal_batchnorm_per_chunk(chunks, input_requires_grad):
    bn = nn.BatchNorm2d(3)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=chunks)

    input = torch.rand(16, 3, 224, 224)
    input = tilt_dist(input)
    a = input.clone().requires_grad_(input_requires_grad)
    b = input.clone().requires_grad_(input_requires_grad)

    out1 = chunked_forward(bn, a, chunks)
    out2 = chunked_forward(dbn, b, chunks)
    assert torch.allclose(out1, out2, atol=1e-4)

    out1.mean().backward()
    out2.mean().backward()
    assert torch.allclose(bn.weight.grad, dbn.weight.grad, atol=1e-4)
    if input_requires_grad:
        assert torch.allclose(a.grad, b.grad, atol=1e-4)


@pytest.mark.parametrize('momentum', [0.1, None])
def test_running_stats_equal_full_batch(momentum):
    bn = nn.BatchNorm2d(3, momentum=momentum)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)

    input = tilt_dist(torch.rand(16, 3, 224, 224))
    bn(input)
    chunked_forward(dbn, input)
    assert torch.allclose(bn.running_mean, dbn.running_mean, atol=1e-4)
    assert torch.allclose(bn.running_var, dbn.running_var, atol=1e-4)
    assert int(dbn.num_batches_tracked) == 1
    assert dbn.sum.abs().sum() == 0 and dbn.counter == 0 and dbn.tracked == 0


def test_convert_rules():
    bn = nn.BatchNorm2d(3, track_running_stats=True)
    converted = DeferredBatchNorm.convert_deferred_batch_norm(bn, chunks=2)
    assert isinstance(converted, DeferredBatchNorm)
    assert converted.weight is bn.weight and converted.running_mean is bn.running_mean

    free = nn.BatchNorm2d(3, track_running_stats=False)
    assert DeferredBatchNorm.convert_deferred_batch_norm(free, chunks=2) is free

    again = DeferredBatchNorm.convert_deferred_batch_norm(converted, chunks=2)
    assert again is converted
    other = DeferredBatchNorm.convert_deferred_batch_norm(converted, chunks=3)
    assert other is not converted and other.chunks == 3

    nested = nn.Sequential(nn.Conv2d(3, 3, 1), nn.Sequential(nn.BatchNorm2d(3)))
    out = DeferredBatchNorm.convert_deferred_batch_norm(nested, chunks=2)
    assert out is nested and isinstance(nested[1][0], DeferredBatchNorm)

    no_affine = DeferredBatchNorm.convert_deferred_batch_norm(nn.BatchNorm2d(3, affine=False), 2)
    assert no_affine.weight is None


def test_state_dict_has_accumulators():
    dbn = DeferredBatchNorm(3, chunks=2)
    assert {'sum', 'sum_squares', 'running_mean', 'running_var', 'num_batches_tracked',
            'weight', 'bias'} == set(dbn.state_dict().keys())


def test_eval_uses_running_stats():
    bn = nn.BatchNorm2d(3)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)
    input = tilt_dist(torch.rand(16, 3, 32, 32))
    bn(input)
    chunked_forward(dbn, input)
    bn.eval()
    dbn.eval()
    assert torch.allclose(bn(input), dbn(input), atol=1e-1, rtol=1e-3)
    assert dbn.tracked == 0


def test_optimize_like_batchnorm():
    bn = nn.BatchNorm2d(3)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)
    opt = optim.SGD(chain(bn.parameters(), dbn.parameters()), lr=1.0)
    for i in range(5):
        input = tilt_dist(torch.rand(16, 3, 32, 32))
        a = bn(input).sum()
        b = chunked_forward(dbn, input).sum()
        (a + b).backward()
        opt.step()
        opt.zero_grad()
        # per-chunk normalisation differs from full-batch: only running stats must agree
        assert torch.allclose(bn.running_mean, dbn.running_mean, atol=1e-1 * (10 ** i))


def test_conv_then_bn():
    bn = nn.Sequential(nn.Conv2d(3, 3, 1), nn.BatchNorm2d(3))
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)
    input = tilt_dist(torch.rand(16, 3, 32, 32))
    opt = optim.SGD(chain(bn.parameters(), dbn.parameters()), lr=0.1)
    a = bn(input)
    a.sum().backward()
    b = chunked_forward(dbn, input)
    b.sum().backward()
    opt.step()
    assert torch.allclose(bn[0].weight, dbn[0].weight, atol=1e-3)


def test_accumulators_do_not_require_grad_and_recompute_does_not_track():
    dbn = DeferredBatchNorm(3, chunks=CHUNKS)
    input = torch.rand(4, 3, 8, 8, requires_grad=True)
    out = checkpoint(dbn, input)
    assert dbn.tracked == 1
    out.sum().backward()
    assert dbn.tracked == 1             # recomputation was not counted
    assert not dbn.sum.requires_grad and dbn.sum.grad_fn is None
    assert not dbn.sum_squares.requires_grad


def test_input_dim_check():
    dbn = DeferredBatchNorm(3)
    with pytest.raises(ValueError, match='expected at least 3D input'):
        dbn(torch.rand(2, 3))
    dbn(torch.rand(2, 3, 5))
