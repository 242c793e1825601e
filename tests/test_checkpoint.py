from functools import partial

import pytest
import torch
from torch import nn

from torchgpipe_b200.checkpoint import (Checkpointing, checkpoint, is_checkpointing,
                                        is_recomputing)
from torchgpipe_b200.dependency import fork, join
from torchgpipe_b200.microbatch import Batch

devices = ['cpu', pytest.param('cuda', marks=pytest.mark.gpu)]


@pytest.mark.parametrize('device', devices)
def test_recompute_runs_before_checkpoint_backward(device):
    timeline = []

    class Log(torch.autograd.Function):
        @staticmethod
        def forward(ctx, name, x):
            ctx.name = name
            timeline.append(f'{name}:forward')
            return x.detach()

        @staticmethod
        def backward(ctx, g):
            timeline.append(f'{ctx.name}:backward')
            return None, g

    a = torch.rand(1, device=device, requires_grad=True)
    b = torch.rand(1, device=device, requires_grad=True)

    # Lane a is checkpointed; lane b depends on a via fork/join.
    f = lambda a: Log.apply('a', a)  # noqa: E731
    chk = Checkpointing(f, Batch(a))
    a_out = chk.checkpoint()
    chk.recompute(a_out)
    a1 = a_out.tensor
    a1, phony = fork(a1)
    b = join(b, phony)
    b = Log.apply('b', b)
    c = torch.cat((a1, b))
    out = c.sum()
    out.backward()
    assert timeline == ['a:forward', 'b:forward', 'b:backward', 'a:forward', 'a:backward']


def test_checkpoint_without_grad_inputs():
    model = nn.Linear(1, 1)
    x = torch.rand(2, 1)            # does not require grad
    out = checkpoint(model, x)
    assert out.requires_grad
    out.sum().backward()
    assert model.weight.grad is not None


def test_checkpoint_under_no_grad():
    model = nn.Linear(1, 1)
    x = torch.rand(2, 1, requires_grad=True)
    with torch.no_grad():
        chk = Checkpointing(model, Batch(x))
        out = chk.checkpoint()
        chk.recompute(out)
    assert not out.tensor.requires_grad


def test_checkpoint_matches_plain_gradients():
    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(4, 8), nn.Tanh(), nn.Linear(8, 2))
    x = torch.rand(5, 4, requires_grad=True)
    model(x).sum().backward()
    ref = [p.grad.clone() for p in model.parameters()] + [x.grad.clone()]
    for p in model.parameters():
        p.grad = None
    x.grad = None
    checkpoint(model, x).sum().backward()
    got = [p.grad for p in model.parameters()] + [x.grad]
    for r, g in zip(ref, got):
        assert torch.allclose(r, g)


def test_checkpoint_tuple_io():
    def f(ab):
        a, b = ab
        return a * 2, b * 3
    a = torch.rand(2, requires_grad=True)
    b = torch.rand(2, requires_grad=True)
    c, d = checkpoint(f, (a, b))
    (c.sum() + d.sum()).backward()
    assert torch.allclose(a.grad, torch.full((2,), 2.)) and torch.allclose(b.grad, torch.full((2,), 3.))


@pytest.mark.parametrize('device', devices)
def test_dropout_mask_is_reproduced(device):
    x = torch.ones(1000, device=device, requires_grad=True)
    y = checkpoint(partial(torch.nn.functional.dropout, p=0.5, training=True), x)
    y.norm().backward()
    # grad is non-zero exactly where the first forward kept the element
    assert torch.equal(y.detach() != 0, x.grad != 0)


def test_rng_state_outside_is_untouched_by_recompute():
    x = torch.ones(10, requires_grad=True)
    y = checkpoint(partial(torch.nn.functional.dropout, p=0.5, training=True), x)
    before = torch.get_rng_state()
    y.sum().backward()
    assert torch.equal(before, torch.get_rng_state())


def test_phase_flags():
    log = []

    class Probe(nn.Module):
        def forward(self, x):
            log.append((is_checkpointing(), is_recomputing()))
            return x * 1

    m = Probe()
    x = torch.rand(1, requires_grad=True)
    m(x)
    assert log == [(False, False)]
    log.clear()
    y = checkpoint(m, x)
    assert log == [(True, False)]
    y.backward()
    assert log == [(True, False), (False, True)]
    assert (is_checkpointing(), is_recomputing()) == (False, False)
