import copy

import pytest
import torch
from torch import nn

from torchgpipe_b200.skip import Namespace, pop, skippable, stash, verify_skippables
from torchgpipe_b200.skip.tracker import (SkipTracker, SkipTrackerThroughPotals,
                                          current_skip_tracker, use_skip_tracker)
from torchgpipe_b200.skip.layout import SkipLayout, inspect_skip_layout


def test_namespace_semantics():
    a, b = Namespace(), Namespace()
    assert a != b and a == a and a == copy.copy(a)
    assert hash(a) == hash(copy.copy(a))
    assert (a < b) != (b < a)
    assert isinstance(None, Namespace)
    assert sorted([b, a]) == sorted([a, b])
    assert "Namespace" in repr(a)


def test_skippable_repr_and_name():
    @skippable(stash=['hello'])
    class Hello(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(1, 1, 1)

        def forward(self, x):
            yield stash('hello', x)
            return self.conv(x)

    m = Hello()
    assert Hello.__name__ == 'Hello'
    assert repr(m).startswith('@skippable(Hello(')
    assert 'module.conv.weight' in m.state_dict()


def test_1to3_without_gpipe():
    @skippable(stash=['1to3'])
    class L1(nn.Module):
        def forward(self, x):
            yield stash('1to3', x)
            return x * 2

    class L2(nn.Module):
        def forward(self, x):
            return x + 1

    @skippable(pop=['1to3'])
    class L3(nn.Module):
        def forward(self, x):
            skip = yield pop('1to3')
            return x + skip

    model = nn.Sequential(L1(), L2(), L3())
    x = torch.rand(4, requires_grad=True)
    y = model(x)
    assert torch.allclose(y, x * 3 + 1)
    y.sum().backward()
    assert torch.allclose(x.grad, torch.full((4,), 3.))


def test_plain_forward_and_none_skip():
    @skippable(stash=['none'])
    class S(nn.Module):
        def forward(self, x):
            yield stash('none', None)
            return x

    @skippable(pop=['none'])
    class P(nn.Module):
        def forward(self, x):
            n = yield pop('none')
            assert n is None
            return x

    @skippable()
    class Plain(nn.Module):
        def forward(self, x):
            return x + 1

    x = torch.rand(1)
    assert torch.equal(nn.Sequential(S(), Plain(), P())(x), x + 1)


def test_stash_pop_runtime_errors():
    @skippable(stash=['foo'])
    class StashWrong(nn.Module):
        def forward(self, x):
            yield stash('bar', x)
            return x

    @skippable(stash=['foo'])
    class StashNothing(nn.Module):
        def forward(self, x):
            return x

    @skippable(pop=['foo'])
    class PopNothing(nn.Module):
        def forward(self, x):
            return x

    @skippable(pop=['foo'])
    class PopWrong(nn.Module):
        def forward(self, x):
            y = yield pop('bar')
            return x

    @skippable(stash=['foo'])
    class StashFoo(nn.Module):
        def forward(self, x):
            yield stash('foo', x)
            return x

    @skippable()
    class BadCommand(nn.Module):
        def forward(self, x):
            yield 'not a command'
            return x

    x = torch.rand(1)
    with pytest.raises(RuntimeError, match="'bar' has not been declared as stashable"):
        StashWrong()(x)
    with pytest.raises(RuntimeError, match="'foo' must be stashed but have not"):
        StashNothing()(x)
    with pytest.raises(RuntimeError, match="'foo' has not been stashed"):
        PopNothing()(x)
    with pytest.raises(RuntimeError, match="'foo' must be popped but have not"):
        nn.Sequential(StashFoo(), PopNothing())(x)
    with use_skip_tracker(SkipTracker()):
        with pytest.raises(RuntimeError, match="'bar' has not been declared as poppable"):
            nn.Sequential(StashFoo(), PopWrong())(x)
    with pytest.raises(TypeError, match='not a command from @skippable'):
        BadCommand()(x)


def test_isolate_and_stash_many():
    @skippable(stash=['a', 'b'])
    class SS(nn.Module):
        def forward(self, x):
            yield stash('a', x)
            yield stash('b', x * 2)
            return x

    @skippable(pop=['a', 'b'])
    class PP(nn.Module):
        def forward(self, x):
            a = yield pop('a')
            b = yield pop('b')
            return a + b

    ns1, ns2 = Namespace(), Namespace()
    s = SS().isolate(ns1, only=['a']).isolate(ns2, only=['b'])
    assert s.isolate(ns1, only=['a']) is s
    assert set(s.stashable()) == {(ns1, 'a'), (ns2, 'b')}
    p = PP().isolate(ns1, only=['a']).isolate(ns2, only=['b'])
    x = torch.rand(2)
    assert torch.allclose(nn.Sequential(s, p)(x), x * 3)


def _layers():
    @skippable(stash=['foo'])
    class Stash(nn.Module):
        def forward(self, x):
            yield stash('foo', x)
            return x

    @skippable(pop=['foo'])
    class Pop(nn.Module):
        def forward(self, x):
            foo = yield pop('foo')
            return x + foo

    return Stash, Pop


def test_verify_skippables_messages():
    Stash, Pop = _layers()
    verify_skippables(nn.Sequential(Stash(), Pop()))

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Stash()))
    assert "no module declared 'foo' as poppable but stashed" in str(e.value)

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Pop()))
    assert "'0' declared 'foo' as poppable but it was not stashed" in str(e.value)

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Stash(), Stash(), Pop()))
    assert "'1' redeclared 'foo' as stashable but not isolated by namespace" in str(e.value)

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Stash(), Pop(), Pop()))
    assert "'2' redeclared 'foo' as poppable but not isolated by namespace" in str(e.value)

    @skippable(stash=['foo'], pop=['foo'])
    class Both(nn.Module):
        pass

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Both()))
    assert "'0' declared 'foo' both as stashable and as poppable" in str(e.value)

    with pytest.raises(TypeError) as e:
        verify_skippables(nn.Sequential(Stash(), Stash(), Pop(), Pop()))
    assert str(e.value).count('*') == 2 and 'one or more pairs of stash and pop do not match' in str(e.value)

    ns1, ns2 = Namespace(), Namespace()
    verify_skippables(nn.Sequential(Stash().isolate(ns1), Stash().isolate(ns2),
                                    Pop().isolate(ns2), Pop().isolate(ns1)))


def test_inspect_skip_layout():
    Stash, Pop = _layers()

    class Pass(nn.Module):
        def forward(self, x):
            return x

    layout = inspect_skip_layout([nn.Sequential(Pass()), nn.Sequential(Pass())])
    assert list(layout.copy_policy(0)) == [] and list(layout.copy_policy(1)) == []

    layout = inspect_skip_layout([nn.Sequential(Stash(), Pop())])
    assert list(layout.copy_policy(0)) == []
    assert not layout.requires_copy(None, 'foo')

    layout = inspect_skip_layout([nn.Sequential(Stash()), nn.Sequential(Pass()),
                                  nn.Sequential(Pop())])
    assert list(layout.copy_policy(2)) == [(0, None, 'foo')]
    assert list(layout.copy_policy(1)) == []
    assert layout.requires_copy(None, 'foo')
    assert not layout.requires_copy(None, 'unknown')

    ns1, ns2 = Namespace(), Namespace()
    layout = inspect_skip_layout([
        nn.Sequential(Stash().isolate(ns1)), nn.Sequential(Stash().isolate(ns2)),
        nn.Sequential(Pop().isolate(ns2), Pop().isolate(ns1))])
    assert sorted((src, name) for src, ns, name in layout.copy_policy(2)) == [(0, 'foo'), (1, 'foo')]
    assert [src for src, _, _ in layout.copy_policy(2)] == [0, 1]     # ascending source
    assert [(s, d) for s, d, _, _ in layout.cross_routes()] == [(0, 2), (1, 2)]


def test_default_tracker_is_thread_local_and_lazy():
    import threading
    main = current_skip_tracker()
    assert type(main) is SkipTracker and current_skip_tracker() is main
    seen = []
    t = threading.Thread(target=lambda: seen.append(current_skip_tracker()))
    t.start()
    t.join()
    assert seen[0] is not main
    with pytest.raises(TypeError, match='copy is not supported'):
        main.copy(None, None, None, None, 'x')


def test_use_skip_tracker_restores():
    a = SkipTracker()
    before = current_skip_tracker()
    with use_skip_tracker(a):
        assert current_skip_tracker() is a
    assert current_skip_tracker() is before
