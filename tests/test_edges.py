import weakref

import pytest
import torch

from torchgpipe_b200.dependency import Fork, Join, fork, join
from torchgpipe_b200.phony import get_phony


def test_phony_is_empty_and_cached():
    p = get_phony(torch.device('cpu'), requires_grad=False)
    assert p.size() == (0,)
    assert get_phony(torch.device('cpu'), requires_grad=False) is p
    q = get_phony(torch.device('cpu'), requires_grad=True)
    assert q is not p and q.requires_grad and not p.requires_grad


def test_phony_must_be_detached_inside_functions():
    p = get_phony(torch.device('cpu'), requires_grad=False)

    class Leaky(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return p

        @staticmethod
        def backward(ctx, g):
            return g

    class Clean(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return p.detach()

        @staticmethod
        def backward(ctx, g):
            return g

    x = torch.rand(1, requires_grad=True)
    out = Clean.apply(x)
    assert out is not p and p.grad_fn is None
    assert not p.requires_grad


def test_fork_join_backward_order_across_lanes():
    log = []

    class Log(torch.autograd.Function):
        @staticmethod
        def forward(ctx, tag, x):
            ctx.tag = tag
            return x.detach()

        @staticmethod
        def backward(ctx, g):
            log.append(ctx.tag)
            return None, g

    a = torch.rand(1, requires_grad=True)
    b = torch.rand(1, requires_grad=True)
    a = Log.apply('a', a)
    a, phony = fork(a)
    b = join(b, phony)
    b = Log.apply('b', b)
    (a + b).backward()
    assert log == ['b', 'a']


def test_fork_without_grad_creates_no_node(monkeypatch):
    def deny(*args, **kwargs):
        raise AssertionError('Function.apply must not be called')
    monkeypatch.setattr(Fork, 'apply', deny)
    monkeypatch.setattr(Join, 'apply', deny)

    x = torch.rand(1)
    y, phony = fork(x)
    assert y is x and not phony.requires_grad
    assert join(x, phony) is x

    z = torch.rand(1, requires_grad=True)
    with torch.no_grad():
        w, phony = fork(z)
        assert w is z
        assert join(z, phony) is z


def test_join_when_only_phony_requires_grad():
    x = torch.rand(1)
    p = torch.empty(0, requires_grad=True)
    y = join(x, p)
    assert y.requires_grad and y is not x


def test_fork_join_leak():
    leak = None

    class F(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return x

        @staticmethod
        def backward(ctx, g):
            nonlocal leak
            leak = weakref.ref(ctx)
            return g

    a = torch.rand(1, requires_grad=True)
    b = torch.rand(1, requires_grad=True)
    a = F.apply(a)
    a, phony = fork(a)
    b = join(b, phony)
    b.backward()
    del a, b, phony
    assert leak is None or leak() is None
