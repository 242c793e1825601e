import threading
import time

import pytest
import torch

from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.runtime.executor import ExecutorPool, StageExecutor, Task
from torchgpipe_b200.runtime.streams import CPUStream
from torchgpipe_b200.worker import spawn_workers


class fake_device:
    """Stands in for a device; every instance is distinct, so each gets its own worker."""
    type = 'fake'
    index = None


def test_pool_shares_thread_per_device():
    with ExecutorPool([torch.device('cpu'), torch.device('cpu', 0), fake_device()]) as pool:
        assert pool[0] is pool[1]
        assert pool[0] is not pool[2]
        assert pool.num_threads == 2


def test_threads_persist_across_submissions_and_join_on_close():
    pool = ExecutorPool([fake_device(), fake_device()])
    seen = []

    def compute():
        seen.append(threading.current_thread().ident)
        return Batch(torch.tensor(1))

    for _ in range(3):
        tickets = [pool[j].submit(Task(CPUStream, compute=compute, finalize=None))
                   for j in range(2)]
        for t in tickets:
            ok, _ = t.result()
            assert ok
    assert len(set(seen)) == 2          # two long-lived threads served all 6 tasks
    pool.close()
    assert not pool[0].alive and not pool[1].alive
    with pytest.raises(RuntimeError):
        pool[0].submit(Task(CPUStream, compute=compute, finalize=None))


def test_tasks_run_in_parallel_across_devices():
    with ExecutorPool([fake_device(), fake_device()]) as pool:
        def slow():
            time.sleep(0.2)
            return Batch(torch.tensor(1))
        tick = time.time()
        tickets = [pool[j].submit(Task(CPUStream, compute=slow, finalize=None)) for j in range(2)]
        for t in tickets:
            t.result()
        assert time.time() - tick < 0.35


def test_exception_is_captured_and_worker_survives():
    class Boom(Exception):
        pass

    def bad():
        raise Boom('x')

    def good():
        return Batch(torch.tensor(7))

    with ExecutorPool([fake_device()]) as pool:
        ok, payload = pool[0].submit(Task(CPUStream, compute=bad, finalize=None)).result()
        assert not ok
        assert payload[0] is Boom and payload[2] is not None
        ok, (task, batch) = pool[0].submit(Task(CPUStream, compute=good, finalize=None)).result()
        assert ok and batch.tensor.item() == 7


@pytest.mark.parametrize('grad_mode', [True, False])
def test_grad_mode_is_carried_to_worker(grad_mode):
    observed = []

    def compute():
        observed.append(torch.is_grad_enabled())
        return Batch(torch.tensor(1))

    with ExecutorPool([fake_device()]) as pool:
        with torch.set_grad_enabled(grad_mode):
            task = Task(CPUStream, compute=compute, finalize=None)
        pool[0].submit(task).result()
    assert observed == [grad_mode]


def test_finalize_runs_only_if_given():
    calls = []
    t = Task(CPUStream, compute=lambda: Batch(torch.tensor(1)), finalize=lambda b: calls.append(b))
    b = t.compute()
    t.finalize(b)
    assert calls == [b]
    Task(CPUStream, compute=lambda: b, finalize=None).finalize(b)


def test_spawn_workers_facade():
    with spawn_workers([fake_device(), fake_device()]) as (in_queues, out_queues):
        assert len(in_queues) == len(out_queues) == 2
        for q in in_queues:
            q.put(Task(CPUStream, compute=lambda: Batch(torch.tensor(3)), finalize=None))
        for q in out_queues:
            ok, (task, batch) = q.get(timeout=5)
            assert ok and batch.tensor.item() == 3


def test_spawn_workers_joins_on_exception():
    before = threading.active_count()
    with pytest.raises(RuntimeError):
        with spawn_workers([fake_device() for _ in range(4)]):
            assert threading.active_count() == before + 4
            raise RuntimeError('boom')
    time.sleep(0.05)
    assert threading.active_count() == before


def test_graph_engine_mailboxes_wake_up_when_another_stage_fails():
    """The in-process transport of GPipe(cuda_graphs=True): a stage blocked on a mailbox must
    stop as soon as some other stage has failed (no GPU needed for this part)."""
    import queue
    import threading
    import time
    from torchgpipe_b200.runtime.graph_pipeline import PipelineAborted, _Mailboxes, _StageThread
    boxes, lock, failed = {}, threading.Lock(), threading.Event()
    a, b = _Mailboxes(0, boxes, lock, failed), _Mailboxes(1, boxes, lock, failed)
    a.send(1, ('fwd', 0, 0), 'payload')
    assert b.recv(0, ('fwd', 0, 0)) == 'payload'
    # messages are keyed by (destination, source, key): nothing crosses over
    a.send(1, ('fwd', 1, 0), 'x')
    b.send(0, ('bwd', 1, 0), 'y')
    assert a.recv(1, ('bwd', 1, 0)) == 'y' and b.recv(0, ('fwd', 1, 0)) == 'x'

    got = []

    def waiter():
        try:
            b.recv(0, ('fwd', 7, 0))
        except PipelineAborted as exc:
            got.append(str(exc))

    t = threading.Thread(target=waiter)
    t.start()
    time.sleep(0.2)
    failed.set()
    t.join(timeout=5)
    assert not t.is_alive() and got and 'another stage failed' in got[0]

    # the stage thread delivers results and exceptions to the submitter and ends on None
    import torch
    worker = _StageThread.__new__(_StageThread)
    threading.Thread.__init__(worker, daemon=True)
    worker.device, worker.jobs = torch.device('cpu'), queue.Queue()
    worker.run = lambda: _run_without_cuda(worker)     # no torch.cuda.set_device on this box
    worker.start()
    ok = worker.submit(lambda: 41 + 1)
    bad = worker.submit(lambda: 1 / 0)
    ok[0].wait(5), bad[0].wait(5)
    assert ok[1] == (True, 42) and bad[1][0] is False and isinstance(bad[1][1], ZeroDivisionError)
    worker.jobs.put(None)
    worker.join(timeout=5)
    assert not worker.is_alive()


def _run_without_cuda(worker):
    while True:
        job = worker.jobs.get()
        if job is None:
            return
        fn, box = job
        try:
            box.append((True, fn()))
        except BaseException as exc:
            box.append((False, exc))
        finally:
            box[0].set()
