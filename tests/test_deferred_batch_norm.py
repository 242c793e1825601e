from copy import deepcopy
from itertools import chain

import pytest
import torch
from torch import nn, optim

from torchgpipe_b200.batchnorm import DeferredBatchNorm
from torchgpipe_b200.checkpoint import checkpoint

CHUNKS = 4


def drifting_batch(samples=16, hw=32):
    """Channels on very different scales whose mean also drifts from sample to sample, so the
    statistics of a quarter of the batch are nowhere near those of the whole batch."""
    scale = torch.tensor([1.0, 10.0, 100.0]).view(1, 3, 1, 1)
    drift = torch.tensor([2.0, 5.0, 7.0]).view(1, 3, 1, 1)
    position = torch.arange(samples, dtype=torch.float32).view(-1, 1, 1, 1)
    return torch.rand(samples, 3, hw, hw) * scale + drift * position


def run_in_micro_batches(model, input, chunks=CHUNKS):
    return torch.cat([model(piece) for piece in input.chunk(chunks)])


@pytest.mark.parametrize('chunks', [1, 4])
@pytest.mark.parametrize('input_requires_grad', [True, False])
def test_output_and_grad_equal_batchnorm_per_chunk(chunks, input_requires_grad):
    plain = nn.BatchNorm2d(3)
    deferred = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(plain), chunks=chunks)

    data = drifting_batch(hw=64)
    x_plain = data.clone().requires_grad_(input_requires_grad)
    x_deferred = data.clone().requires_grad_(input_requires_grad)

    y_plain = run_in_micro_batches(plain, x_plain, chunks)
    y_deferred = run_in_micro_batches(deferred, x_deferred, chunks)
    assert torch.allclose(y_plain, y_deferred, atol=1e-4)

    y_plain.mean().backward()
    y_deferred.mean().backward()
    assert torch.allclose(plain.weight.grad, deferred.weight.grad, atol=1e-4)
    assert torch.allclose(plain.bias.grad, deferred.bias.grad, atol=1e-4)
    if input_requires_grad:
        assert torch.allclose(x_plain.grad, x_deferred.grad, atol=1e-4)


@pytest.mark.parametrize('momentum', [0.1, None])
def test_running_stats_equal_full_batch(momentum):
    plain = nn.BatchNorm2d(3, momentum=momentum)
    deferred = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(plain), chunks=CHUNKS)

    for step in range(2):       # two mini-batches: the cumulative average has to advance too
        data = drifting_batch(hw=64) + step
        plain(data)
        run_in_micro_batches(deferred, data)
        assert torch.allclose(plain.running_mean, deferred.running_mean, rtol=1e-4, atol=1e-4)
        # nn.BatchNorm commits the unbiased variance, the deferred layer the biased one
        n = data.numel() // 3
        assert torch.allclose(plain.running_var, deferred.running_var, rtol=2.0 / n + 1e-4, atol=1e-3)
        assert int(deferred.num_batches_tracked) == step + 1
        assert deferred.sum.abs().sum() == 0 and deferred.counter == 0 and deferred.tracked == 0


def test_partial_mini_batch_leaves_running_stats_alone():
    deferred = DeferredBatchNorm(3, chunks=CHUNKS)
    before = deferred.running_mean.clone()
    pieces = drifting_batch().chunk(CHUNKS)
    for k, piece in enumerate(pieces[:-1]):
        deferred(piece)
        assert deferred.tracked == k + 1 and deferred.counter == (k + 1) * piece.numel() // 3
        assert torch.equal(deferred.running_mean, before)
    deferred(pieces[-1])
    assert not torch.equal(deferred.running_mean, before) and deferred.tracked == 0


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
    input = drifting_batch()
    bn(input)
    run_in_micro_batches(dbn, input)
    bn.eval()
    dbn.eval()
    assert torch.allclose(bn(input), dbn(input), atol=1e-1, rtol=1e-3)
    assert dbn.tracked == 0


def test_optimize_like_batchnorm():
    bn = nn.BatchNorm2d(3)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)
    opt = optim.SGD(chain(bn.parameters(), dbn.parameters()), lr=1.0)
    for i in range(5):
        input = drifting_batch()
        a = bn(input).sum()
        b = run_in_micro_batches(dbn, input).sum()
        (a + b).backward()
        opt.step()
        opt.zero_grad()
        # per-chunk normalisation differs from full-batch: only running stats must agree
        assert torch.allclose(bn.running_mean, dbn.running_mean, atol=1e-1 * (10 ** i))


def test_conv_then_bn():
    bn = nn.Sequential(nn.Conv2d(3, 3, 1), nn.BatchNorm2d(3))
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(deepcopy(bn), chunks=CHUNKS)
    input = drifting_batch()
    opt = optim.SGD(chain(bn.parameters(), dbn.parameters()), lr=0.1)
    a = bn(input)
    a.sum().backward()
    b = run_in_micro_batches(dbn, input)
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
