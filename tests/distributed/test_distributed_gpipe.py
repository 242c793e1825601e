import copy
import os
import threading

import pytest
import torch
from torch import nn
import torch.multiprocessing as mp

from torchgpipe_b200.distributed import context
from torchgpipe_b200.distributed.gpipe import (DistributedGPipe, DistributedGPipeDataLoader,
                                               get_module_partition)
from torchgpipe_b200.distributed.utils import to
from torchgpipe_b200.skip import Namespace, pop, skippable, stash


def test_get_module_partition():
    a, b, c = nn.Linear(1, 1), nn.ReLU(), nn.Linear(1, 1)
    model = nn.Sequential(a, b, c)
    p0 = get_module_partition(model, 0, [2, 1], None)
    p1 = get_module_partition(model, 1, [2, 1], torch.device('cpu'))
    assert list(p0) == [a, b] and list(p1) == [c]
    assert list(p1.state_dict().keys()) == ['2.weight', '2.bias']
    with pytest.raises(ValueError):
        get_module_partition(model, 0, [1, 1], None)
    with pytest.raises(ValueError):
        get_module_partition(model, 0, [3, 0], None)
    with pytest.raises(RuntimeError, match='mismatch'):
        get_module_partition(model, 2, [2, 1], None)


def test_to_helper():
    x = torch.rand(2, requires_grad=True)
    y = to(torch.device('cpu'), x * 2)
    assert y.requires_grad and y.grad_fn is None
    a, n, b = to(None, (x.detach(), None, x * 3))
    assert n is None and not a.requires_grad and b.requires_grad and b.is_leaf
    assert to(torch.device('cpu'), None) is None


def test_ctor_validation():
    model = nn.Sequential(nn.Linear(1, 1))
    with pytest.raises(ValueError, match='balance is required'):
        DistributedGPipe(model, 0, {0: 'a'})
    with pytest.raises(ValueError, match='chunks'):
        DistributedGPipe(model, 0, {0: 'a'}, [1], 0)
    with pytest.raises(ValueError, match='checkpoint'):
        DistributedGPipe(model, 0, {0: 'a'}, [1], 1, checkpoint='bad')
    with pytest.raises(TypeError):
        DistributedGPipe(nn.Linear(1, 1), 0, {0: 'a'}, [1], 1)
    g = DistributedGPipe(model, 0, {0: 'a'}, [1], 2)
    assert (g.rank, g.world_size, g.chunks, g.name) == (0, 1, 2, 'a')
    assert g.model() is g.module and isinstance(g.module, nn.Sequential)


def _full_model(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(4, 8), nn.Tanh(), nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 2))


def _reference_grads(model, x, chunks):
    model = copy.deepcopy(model)
    losses = [model(c).pow(2).mean() for c in x.chunk(chunks)]
    sum(losses).backward()
    return [p.grad.clone() for p in model.parameters()], [loss.item() for loss in losses]


def _run_stage_threads(model, balance, chunks, x, checkpoint, dbn=False):
    """Simulates the ranks as threads in one process over the channel backend."""
    n = len(balance)
    workers = {r: f'stage{r}-{id(model)}' for r in range(n)}
    results = {}
    errors = []

    def run(rank):
        try:
            pipe = DistributedGPipe(model, rank, workers, balance, chunks,
                                    device=torch.device('cpu'), checkpoint=checkpoint,
                                    deferred_batch_norm=dbn)
            outs = pipe.forward(x if rank == 0 else None)
            if rank == n - 1:
                losses = [o.pow(2).mean() for o in outs]
                results['losses'] = [loss.item() for loss in losses]
                pipe.backward(losses)
            else:
                pipe.backward(None)
            results[rank] = [p.grad.clone() for p in pipe.model().parameters()]
        except Exception as e:      # pragma: no cover
            import traceback
            traceback.print_exc()
            errors.append(e)

    from contextlib import ExitStack
    with ExitStack() as stack:
        for name in workers.values():
            stack.enter_context(context.worker(name, chunks))
        threads = [threading.Thread(target=run, args=(r,)) for r in range(n)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=60)
            assert not t.is_alive(), 'stage thread hung'
    assert not errors
    return results


@pytest.mark.timeout(120)
@pytest.mark.parametrize('checkpoint', ['never', 'except_last', 'always'])
@pytest.mark.parametrize('balance', [[5], [2, 3], [1, 2, 2]])
def test_channel_backend_matches_plain_model(balance, checkpoint):
    model = _full_model()
    x = torch.rand(8, 4)
    ref_grads, ref_losses = _reference_grads(model, x, 4)
    results = _run_stage_threads(copy.deepcopy(model), balance, 4, x, checkpoint)
    got = sum((results[r] for r in range(len(balance))), [])
    assert len(got) == len(ref_grads)
    for g, r in zip(got, ref_grads):
        assert torch.allclose(g, r, atol=1e-6)
    assert results['losses'] == pytest.approx(ref_losses, abs=1e-6)


def _skip_model(seed=0):
    torch.manual_seed(seed)

    @skippable(stash=['a'])
    class S(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = nn.Linear(4, 4)

        def forward(self, x):
            y = self.fc(x)
            yield stash('a', y)
            return y

    @skippable(pop=['a'])
    class P(nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = nn.Linear(4, 2)

        def forward(self, x):
            a = yield pop('a')
            return self.fc(x + 2 * a)

    return nn.Sequential(S(), nn.Linear(4, 4), nn.Tanh(), P())


@pytest.mark.timeout(120)
@pytest.mark.parametrize('checkpoint', ['never', 'always'])
@pytest.mark.parametrize('balance', [[1, 1, 2], [2, 2], [1, 3]])
def test_cross_stage_skip_in_distributed_mode(balance, checkpoint):
    model = _skip_model()
    x = torch.rand(6, 4)
    ref_grads, ref_losses = _reference_grads(model, x, 3)
    results = _run_stage_threads(model, balance, 3, x, checkpoint)
    got = sum((results[r] for r in range(len(balance))), [])
    for g, r in zip(got, ref_grads):
        assert torch.allclose(g, r, atol=1e-6)
    assert results['losses'] == pytest.approx(ref_losses, abs=1e-6)


@pytest.mark.timeout(60)
def test_patched_put_get_hooks():
    """The boundary traffic goes through DistributedGPipe._put/_get (the reference's patch points)."""
    from unittest import mock
    mailbox = {}
    log = []

    def fake_put(name, id, values, backward=False):
        log.append(('put', name, id, backward))
        mailbox[(name, id, backward)] = values

    def fake_get(name, id, backward=False):
        log.append(('get', name, id, backward))
        return mailbox.pop((name, id, backward))

    model = _full_model()
    workers = {0: 'h0', 1: 'h1'}
    with mock.patch.object(DistributedGPipe, '_put', staticmethod(fake_put)), \
            mock.patch.object(DistributedGPipe, '_get', staticmethod(fake_get)):
        s0 = DistributedGPipe(model, 0, workers, [2, 3], 2)
        s1 = DistributedGPipe(model, 1, workers, [2, 3], 2)
        x = torch.rand(4, 4)
        s0.forward(x)
        outs = s1.forward(None)
        assert len(outs) == 2 and outs[0].shape == (2, 2)
        s1.backward([o.sum() for o in outs])
        s0.backward(None)
    assert ('put', 'h1', 0, False) in log and ('get', 'h1', 1, False) in log
    assert ('put', 'h0', 1, True) in log and ('get', 'h0', 0, True) in log
    assert all(p.grad is not None for p in s0.model().parameters())


@pytest.mark.timeout(60)
def test_deferred_batch_norm_is_effective():
    torch.manual_seed(0)
    model = nn.Sequential(nn.Conv2d(3, 3, 1), nn.BatchNorm2d(3))
    ref = copy.deepcopy(model)
    x = torch.rand(8, 3, 6, 6) * 5
    ref(x)
    workers = {0: 'dbn0'}
    with context.worker('dbn0', 4):
        pipe = DistributedGPipe(model, 0, workers, [2], 4, deferred_batch_norm=True)
        pipe.forward(x)
    assert torch.allclose(pipe.model()[1].running_mean, ref[1].running_mean, atol=1e-5)
    assert torch.allclose(pipe.model()[1].running_var, ref[1].running_var, atol=1e-3)


def test_data_loader_roles():
    data = [(torch.full((2,), float(i)), torch.tensor([i, i])) for i in range(5)]
    with context.worker('last', 1):
        first = DistributedGPipeDataLoader(data, 0, 1, 3, False, 'last')
        middle = DistributedGPipeDataLoader(None, 1, 1, 3, False, 'last')
        last = DistributedGPipeDataLoader(None, 2, 1, 3, True, 'last')
        assert len(first) == len(middle) == len(last) == 3
        got_first = list(first)
        assert [d[0].item() for d, t in got_first] == [0., 1., 2.]
        assert all(t is None for _, t in got_first)
        assert list(middle) == [(None, None)] * 3
        got_last = list(last)
        assert all(d is None for d, _ in got_last)
        assert [t[0].item() for _, t in got_last] == [0, 1, 2]


def _gloo_worker(rank, world, port, balance, chunks, checkpoint, ret):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torchgpipe_b200.distributed.transport import ProcessGroupTransport
        model = _skip_model()
        x = torch.rand(6, 4, generator=torch.Generator().manual_seed(1))
        tp = ProcessGroupTransport(torch.device('cpu'), pair_groups=True)
        pipe = DistributedGPipe(model, rank, {r: f'r{r}' for r in range(world)}, balance, chunks,
                                device=torch.device('cpu'), checkpoint=checkpoint, transport=tp)
        for step in range(2):       # second step exercises the cached-shape path
            outs = pipe.forward(x if rank == 0 else None)
            if rank == world - 1:
                pipe.backward([o.pow(2).mean() for o in outs])
            else:
                pipe.backward(None)
        tp.flush()
        ret[rank] = [p.grad.clone() for p in pipe.model().parameters()]
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize('checkpoint,world,balance', [('never', 2, [1, 3]),
                                                      ('except_last', 2, [1, 3]),
                                                      ('except_last', 3, [1, 1, 2])])
def test_gloo_multiprocess_matches_plain_model(checkpoint, world, balance):
    # world 3 has a skip route 0 -> 2 that jumps over stage 1 (per-pair communicators)
    chunks = 3
    port = 29500 + os.getpid() % 2000
    manager = mp.Manager()
    ret = manager.dict()
    mp.spawn(_gloo_worker, args=(world, port, balance, chunks, checkpoint, ret), nprocs=world,
             join=True)
    model = _skip_model()
    x = torch.rand(6, 4, generator=torch.Generator().manual_seed(1))
    ref_grads, _ = _reference_grads(model, x, chunks)
    got = sum((ret[r] for r in range(world)), [])
    for g, r in zip(got, ref_grads):
        assert torch.allclose(g, 2 * r, atol=1e-5)     # two accumulated steps


def _state_worker(rank, world, port, balance, ret):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        model = _full_model(seed=3)
        pipe = DistributedGPipe(model, rank, {r: f's{r}' for r in range(world)}, balance, 2,
                                device=torch.device('cpu'))
        with torch.no_grad():
            for p in pipe.model().parameters():
                p.add_(rank + 1.0)                       # every stage changes its own part
        merged = pipe.gather_sequential_state_dict(dst=0)
        if rank == 0:
            ret['merged'] = merged
        else:
            assert merged is None
        # restore a whole-model checkpoint: each rank takes only its own entries
        fresh = _full_model(seed=4).state_dict()
        pipe.load_sequential_state_dict(fresh)
        ret[rank] = {k: v.clone() for k, v in pipe.sequential_state_dict().items()}
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_state_dict_gather_and_load_across_ranks():
    world, balance = 2, [2, 3]
    port = 31500 + os.getpid() % 2000
    manager = mp.Manager()
    ret = manager.dict()
    mp.spawn(_state_worker, args=(world, port, balance, ret), nprocs=world, join=True)
    plain = _full_model(seed=3)
    want = plain.state_dict()
    merged = ret['merged']
    assert list(merged.keys()) == list(want.keys())          # plain nn.Sequential keys, in order
    for k, v in want.items():
        bump = 1.0 if int(k.split('.')[0]) < 2 else 2.0
        assert torch.allclose(merged[k], v + bump), k
    fresh = _full_model(seed=4).state_dict()
    loaded = {}
    for r in range(world):
        loaded.update(ret[r])
    assert set(loaded) == set(fresh)
    for k, v in fresh.items():
        assert torch.equal(loaded[k], v), k


def test_load_sequential_state_dict_is_strict_about_own_entries():
    model = _full_model()
    pipe = DistributedGPipe(model, 0, {0: 'x0', 1: 'x1'}, [2, 3], 1)
    state = _full_model(seed=9).state_dict()
    del state['0.bias']
    with pytest.raises(RuntimeError, match='0.bias'):
        pipe.load_sequential_state_dict(state)
    pipe.load_sequential_state_dict(state, strict=False)
    assert torch.equal(pipe.model()[0].weight, state['0.weight'])


def _shape_worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torchgpipe_b200.distributed.transport import ProcessGroupTransport
        static = ProcessGroupTransport(torch.device('cpu'))
        dynamic = ProcessGroupTransport(torch.device('cpu'), static_shapes=False)
        if rank == 0:
            static.send(1, ('fwd', 0, 0), torch.ones(2, 3))
            static.send(1, ('fwd', 0, 0), torch.ones(2, 3) * 2)        # same shape: no header
            try:
                static.send(1, ('fwd', 0, 0), torch.ones(4, 3))
                ret['error'] = None
            except ValueError as exc:
                ret['error'] = str(exc)
            dynamic.send(1, ('fwd', 1, 0), torch.ones(2, 3))
            dynamic.send(1, ('fwd', 1, 0), (torch.ones(5), torch.zeros(1, 2)))
            static.flush()
            dynamic.flush()
        else:
            a = static.recv(0, ('fwd', 0, 0))
            b = static.recv(0, ('fwd', 0, 0))
            c = dynamic.recv(0, ('fwd', 1, 0))
            d = dynamic.recv(0, ('fwd', 1, 0))
            ret['got'] = (a.sum().item(), b.sum().item(), tuple(c.shape),
                          tuple(tuple(t.shape) for t in d))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_static_shape_transport_refuses_a_changed_shape_and_dynamic_one_follows_it():
    port = 33500 + os.getpid() % 2000
    manager = mp.Manager()
    ret = manager.dict()
    mp.spawn(_shape_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret['error'] is not None and 'static_shapes' in ret['error']
    assert ret['got'] == (6.0, 12.0, (2, 3), ((5,), (1, 2)))


def _silent_peer_worker(rank, world, port, ret):
    import time
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torchgpipe_b200.distributed.transport import ProcessGroupTransport
        tp = ProcessGroupTransport(torch.device('cpu'), timeout=1.5)
        if rank == 1:
            t0 = time.monotonic()
            try:
                # `like` given: no header round-trip, the payload itself never comes
                tp.recv(0, ('bwd', 0, 0), like=torch.zeros(3))
                ret['error'] = None
            except TimeoutError as exc:
                ret['error'] = str(exc)
            ret['waited'] = time.monotonic() - t0
        else:
            time.sleep(4.0)         # the peer is alive but never sends
    finally:
        try:
            dist.destroy_process_group()
        except Exception:
            pass


@pytest.mark.timeout(120)
def test_receive_from_a_silent_peer_times_out_instead_of_hanging():
    # the reference's Queue.get / rpc wait has no deadline (distributed/context.py:141)
    port = 35500 + os.getpid() % 2000
    manager = mp.Manager()
    ret = manager.dict()
    mp.spawn(_silent_peer_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret['error'] is not None and 'did not arrive' in ret['error']
    assert 1.0 < ret['waited'] < 30.0


def test_channel_get_times_out_with_a_message_naming_the_waiter():
    import queue
    with context.worker('lonely', 1):
        with pytest.raises(context.ChannelTimeout, match="worker 'lonely'"):
            context.get_forward('lonely', 0, timeout=0.05)
        with pytest.raises(queue.Empty):            # the documented older catch still works
            context.get_backward('lonely', 0, timeout=0.05)
        with pytest.raises(TimeoutError):
            context.get_target('lonely', timeout=0.05)


def test_default_channel_timeout_comes_from_the_environment(monkeypatch):
    monkeypatch.setenv('TGB_CHANNEL_TIMEOUT', '12.5')
    assert context.default_timeout() == 12.5
    monkeypatch.setenv('TGB_CHANNEL_TIMEOUT', '0')
    assert context.default_timeout() is None        # wait forever, like the reference
    monkeypatch.delenv('TGB_CHANNEL_TIMEOUT')
    assert context.default_timeout() == 1800.0
