"""Worker API under the reference's module path.

The runtime uses persistent :class:`~torchgpipe_b200.runtime.executor.StageExecutor`
threads. ``spawn_workers`` is kept as a context manager with the reference's
shape (``torchgpipe/worker.py:94-150``): it yields per-stage in/out queues
backed by a temporary executor pool and joins the threads on exit.
"""
from contextlib import contextmanager
from queue import Queue
from typing import Generator, List, Tuple

import torch

from torchgpipe_b200.runtime.executor import (ExecutorPool, StageExecutor, Task,  # noqa: F401
                                              Ticket)

__all__ = ['Task', 'spawn_workers', 'StageExecutor', 'ExecutorPool']


class _InQueue:
    """Queue facade: ``put(task)`` submits to an executor and forwards the result."""

    def __init__(self, executor: StageExecutor, out_queue: Queue) -> None:
        self._executor = executor
        self._out = out_queue

    def put(self, task: Task) -> None:
        ticket = self._executor.submit(task)

        def relay() -> None:
            self._out.put(ticket.result())

        # Results must appear in submission order: the executor is FIFO, so a
        # chained relay on the same executor preserves it.
        self._executor.submit(_Relay(relay))


class _Relay(Task):
    def __init__(self, fn) -> None:  # type: ignore[no-untyped-def]
        self._fn = fn
        self.grad_mode = False

    def compute(self):  # type: ignore[no-untyped-def]
        self._fn()
        return None


@contextmanager
def spawn_workers(devices: List[torch.device],
                  ) -> Generator[Tuple[List[_InQueue], List[Queue]], None, None]:
    """Yields ``(in_queues, out_queues)``; one worker thread per distinct device."""
    pool = ExecutorPool(devices)
    outs = {}
    in_queues: List[_InQueue] = []
    out_queues: List[Queue] = []
    for executor in pool.executors:
        if executor not in outs:
            q: Queue = Queue()
            outs[executor] = (_InQueue(executor, q), q)
        in_queues.append(outs[executor][0])
        out_queues.append(outs[executor][1])
    try:
        yield in_queues, out_queues
    finally:
        pool.close()
