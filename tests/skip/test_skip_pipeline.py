import pytest
import torch
from torch import nn

from torchgpipe_b200 import GPipe, is_checkpointing, is_recomputing
from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.skip import Namespace, pop, skippable, stash
from torchgpipe_b200.skip.layout import SkipLayout
from torchgpipe_b200.skip.portal import Portal, PortalBlue, PortalCopy, PortalOrange
from torchgpipe_b200.skip.tracker import SkipTrackerThroughPotals, current_skip_tracker
from torchgpipe_b200.stream import CPUStream


def _model_1to3():
    @skippable(stash=['1to3'])
    class Layer1(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(3, 3, 1)

        def forward(self, x):
            yield stash('1to3', x)
            return self.conv(x)

    class Layer2(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(3, 3, 1)

        def forward(self, x):
            return self.conv(x)

    @skippable(pop=['1to3'])
    class Layer3(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(3, 3, 1)

        def forward(self, x):
            skip = yield pop('1to3')
            return self.conv(x) + skip

    return nn.Sequential(Layer1(), Layer2(), Layer3())


@pytest.mark.parametrize('balance', [[3], [1, 2], [2, 1], [1, 1, 1]],
                         ids=lambda b: ':'.join(map(str, b)))
@pytest.mark.parametrize('checkpoint', ['never', 'always', 'except_last'])
def test_1to3_matches_plain_model(balance, checkpoint):
    torch.manual_seed(0)
    model = _model_1to3()
    x = torch.rand(6, 3, 8, 8, requires_grad=True)
    ref = model(x)
    ref.mean().backward()
    ref_grads = [p.grad.clone() for p in model.parameters()]
    ref_xgrad = x.grad.clone()
    model.zero_grad()
    x.grad = None

    pipe = GPipe(model, balance, devices=['cpu'] * len(balance), chunks=3, checkpoint=checkpoint)
    out = pipe(x)
    assert torch.allclose(out, ref, atol=1e-6)
    out.mean().backward()
    for p, g in zip(pipe.parameters(), ref_grads):
        assert torch.allclose(p.grad, g, atol=1e-6)
    assert torch.allclose(x.grad, ref_xgrad, atol=1e-6)


@pytest.mark.parametrize('checkpoint', ['never', 'always', 'except_last'])
def test_none_skip_builds_no_portal_nodes(checkpoint):
    @skippable(stash=['none'])
    class Stash(nn.Module):
        def forward(self, x):
            yield stash('none', None)
            return x

    @skippable(pop=['none'])
    class Pop(nn.Module):
        def forward(self, x):
            n = yield pop('none')
            assert n is None
            return x

    pipe = GPipe(nn.Sequential(Stash(), Pop()), [1, 1], devices=['cpu', 'cpu'], chunks=5,
                 checkpoint=checkpoint)
    x = torch.rand(10, requires_grad=True)
    out = pipe(x)

    def names(fn, seen=None):
        seen = set() if seen is None else seen
        if fn is None or fn in seen:
            return seen
        seen.add(fn)
        for nxt, _ in fn.next_functions:
            names(nxt, seen)
        return seen

    kinds = {f.__class__.__name__ for f in names(out.grad_fn)}
    assert not any('Portal' in k for k in kinds)
    out.sum().backward()
    assert torch.allclose(x.grad, torch.ones(10))


def test_same_partition_skip_uses_no_portal(monkeypatch):
    def deny(*a, **k):
        raise AssertionError('Portal must not be created')
    monkeypatch.setattr(Portal, '__init__', deny)
    model = _model_1to3()
    pipe = GPipe(model, [3], devices=['cpu'], chunks=2)
    pipe(torch.rand(4, 3, 4, 4)).mean().backward()
    # and outside GPipe
    model(torch.rand(2, 3, 4, 4))


def test_skip_tensor_travels_directly_between_stages(monkeypatch):
    hops = []
    from torchgpipe_b200.skip import portal as portal_mod
    real = portal_mod.move_tensors

    def spy(tensors, src, dst):
        hops.append('skip')
        return real(tensors, src, dst)
    monkeypatch.setattr(portal_mod, 'move_tensors', spy)

    model = _model_1to3()
    pipe = GPipe(model, [1, 1, 1], devices=['cpu'] * 3, chunks=2, checkpoint='never')
    pipe(torch.rand(2, 3, 4, 4))
    assert len(hops) == 2           # one direct copy per micro-batch, not one per hop


def test_portal_tensor_life_cycle():
    t = torch.rand(1)
    portal = Portal(t, tensor_life=2)
    assert portal.use_tensor() is t and portal.tensor is t
    assert portal.use_tensor() is t and portal.tensor is None
    with pytest.raises(RuntimeError, match='tensor in portal has been removed'):
        portal.use_tensor()
    portal.put_tensor(t, 1)
    assert portal.tensor is t
    portal.put_tensor(t, 0)
    assert portal.tensor is None


def test_portal_grad_is_ephemeral():
    portal = Portal(torch.rand(1), 1)
    with pytest.raises(RuntimeError, match='grad in portal has been removed or never set'):
        portal.use_grad()
    g = torch.rand(1)
    portal.put_grad(g)
    assert portal.use_grad() is g
    with pytest.raises(RuntimeError):
        portal.use_grad()


def test_portal_blue_orange_roundtrip():
    t = torch.rand(3, requires_grad=True)
    main = torch.rand(1, requires_grad=True)
    portal = Portal(t, 2)
    phony = portal.blue()
    assert phony.numel() == 0
    from torchgpipe_b200.dependency import fork, join
    lane = join(main * 1, phony)
    lane, phony2 = fork(lane)
    out = portal.orange(phony2)
    assert torch.equal(out, t) and portal.tensor is None
    (lane.sum() + (out * 2).sum()).backward()
    assert torch.allclose(t.grad, torch.full((3,), 2.))
    assert portal.grad is None


def test_portal_copy_moves_tensor():
    t = torch.rand(3)
    portal = Portal(t, 2)
    phony = torch.empty(0)
    out = portal.copy(CPUStream, CPUStream, phony)
    assert out.numel() == 0
    assert torch.equal(portal.tensor, t)
    none_portal = Portal(None, 2)
    assert none_portal.copy(CPUStream, CPUStream, phony).numel() == 0


def test_tracker_tensor_life_with_checkpointing():
    from torchgpipe_b200.checkpoint import enable_checkpointing, enable_recomputing
    layout = SkipLayout(2, {(None, 'x'): (0, 1)})
    tracker = SkipTrackerThroughPotals(layout)
    batch = Batch(torch.rand(1, requires_grad=True) * 1)
    t = torch.rand(1)
    with enable_checkpointing():
        tracker.save(batch, None, 'x', t)
    portal = tracker.portals[(None, 'x')]
    assert portal.tensor_life == 2      # 3 at creation, blue() spent one
    tracker2 = SkipTrackerThroughPotals(layout)
    tracker2.save(batch, None, 'x', t)
    assert tracker2.portals[(None, 'x')].tensor_life == 1
    with enable_recomputing():
        tracker2.save(batch, None, 'x', t)      # re-stash during recompute: life 1 -> blue -> 0
    assert tracker2.portals[(None, 'x')].tensor is None


def test_no_portal_without_pipeline(monkeypatch):
    def deny(*a, **k):
        raise AssertionError('Portal must not be created')
    monkeypatch.setattr(Portal, '__init__', deny)
    model = _model_1to3()
    model(torch.rand(2, 3, 4, 4)).mean().backward()


def test_skip_leak_after_backward():
    model = _model_1to3()
    pipe = GPipe(model, [1, 1, 1], devices=['cpu'] * 3, chunks=2, checkpoint='always')
    trackers = []
    import torchgpipe_b200.runtime.pipeline as rp
    real = rp.SkipTrackerThroughPotals

    class Spy(real):
        def __init__(self, layout):
            super().__init__(layout)
            trackers.append(self)
    rp.SkipTrackerThroughPotals = Spy
    try:
        out = pipe(torch.rand(2, 3, 4, 4, requires_grad=True))
        out.mean().backward()
    finally:
        rp.SkipTrackerThroughPotals = real
    for tr in trackers:
        for portal in tr.portals.values():
            assert portal.tensor is None and portal.grad is None


def test_tracker_under_data_parallel_threads():
    import threading

    @skippable(stash=['x'])
    class S(nn.Module):
        def forward(self, x):
            yield stash('x', x)
            return x * 2

    @skippable(pop=['x'])
    class P(nn.Module):
        def forward(self, x):
            s = yield pop('x')
            return x + s

    model = nn.Sequential(S(), P())
    results = {}

    def run(k):
        x = torch.full((2,), float(k))
        results[k] = model(x)

    threads = [threading.Thread(target=run, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for k in range(4):
        assert torch.equal(results[k], torch.full((2,), 3. * k))


@pytest.mark.multigpu
@pytest.mark.gpu
@pytest.mark.parametrize('checkpoint', ['never', 'always', 'except_last'])
def test_1to3_on_two_gpus(checkpoint):
    torch.manual_seed(0)
    model = _model_1to3()
    x = torch.rand(6, 3, 8, 8)
    ref = model(x)
    pipe = GPipe(model, [1, 2], devices=['cuda:0', 'cuda:1'], chunks=3, checkpoint=checkpoint)
    out = pipe(x.cuda(0))
    assert torch.allclose(out.cpu(), ref, atol=1e-5)
    out.mean().backward()
