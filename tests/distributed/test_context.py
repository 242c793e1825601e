import queue
import threading

import pytest
import torch

from torchgpipe_b200.distributed import context


def test_worker_registers_and_unregisters():
    with context.worker('w0', 2):
        assert 'w0' in context.GlobalContext.ctxs
        ctx = context.GlobalContext.get_context('w0')
        assert len(ctx.forward_channels) == 2 and len(ctx.backward_channels) == 2
    assert 'w0' not in context.GlobalContext.ctxs


def test_duplicate_worker():
    with context.worker('dup', 1):
        with pytest.raises(RuntimeError, match='already exists'):
            with context.worker('dup', 1):
                pass
    # the failed inner registration must not have removed or corrupted anything
    with context.worker('dup', 1):
        pass


def test_unregisters_on_exception():
    with pytest.raises(ValueError):
        with context.worker('boom', 1):
            raise ValueError
    assert 'boom' not in context.GlobalContext.ctxs


def test_put_get_roundtrip():
    with context.worker('w', 3):
        t = torch.rand(2)
        context.put_forward('w', 1, t)
        assert context.get_forward('w', 1) is t
        context.put_backward('w', 2, (t, t))
        assert context.get_backward('w', 2) == (t, t)
        context.put_target('w', t)
        assert context.get_target('w') is t


def test_unknown_worker_is_key_error():
    with pytest.raises(KeyError):
        context.put_forward('nobody', 0, torch.rand(1))
    with pytest.raises(KeyError):
        context.get_target('nobody')


def test_get_blocks_until_put_and_times_out():
    with context.worker('w', 1):
        with pytest.raises(queue.Empty):
            context.get_forward('w', 0, timeout=0.05)
        t = torch.rand(1)
        threading.Timer(0.05, context.put_forward, args=('w', 0, t)).start()
        assert context.get_forward('w', 0, timeout=2) is t


def test_decorator_nests():
    seen = []

    @context.distributed('outer', 1)
    @context.distributed('inner', 2)
    def train(x):
        seen.append(set(context.GlobalContext.ctxs) >= {'outer', 'inner'})
        return x + 1

    assert train(1) == 2
    assert seen == [True]
    assert 'outer' not in context.GlobalContext.ctxs and 'inner' not in context.GlobalContext.ctxs
