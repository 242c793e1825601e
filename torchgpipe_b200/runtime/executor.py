"""Persistent per-device stage executors.

The reference creates one thread and two queues per device on *every*
``forward`` and joins them at the end (``torchgpipe/pipeline.py:112``,
``torchgpipe/worker.py:94-150``). On B200 a stage's kernels are microseconds
long, so the host launch path is the bottleneck and thread churn is pure
overhead. Here each :class:`GPipe` owns an :class:`ExecutorPool` whose worker
threads live as long as the module: one thread per *distinct* device
(partitions placed on the same device share a thread), each pinned to its
device, each launching onto that device's compute stream. CUDA launches drop
the GIL, so the threads overlap host-side launch work across GPUs.

A :class:`Task` has two halves, like the reference's (``worker.py:30-63``):
``compute`` runs on the worker thread; ``finalize`` runs on the scheduling
thread once the result has been collected (it attaches recomputation).

Failure contract (reference ``worker.py:81-86``, ``pipeline.py:222-249``):
an exception inside ``compute`` is captured with its traceback and returned as
the task's result; the worker thread survives and keeps serving.
"""
import sys
import threading
from queue import SimpleQueue
from types import TracebackType
from typing import Callable, Dict, List, Optional, Tuple, Type, Union
import weakref

import torch

from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.runtime.streams import AbstractStream, use_device, use_stream

__all__ = ['Task', 'StageExecutor', 'ExecutorPool', 'Ticket', 'normalize_device']

ExcInfo = Tuple[Type[BaseException], BaseException, TracebackType]


class Task:
    """A unit of stage work: ``compute`` on the worker, ``finalize`` on the scheduler."""

    __slots__ = ('stream', '_compute', '_finalize', 'grad_mode')

    def __init__(self,
                 stream: AbstractStream,
                 *,
                 compute: Callable[[], Batch],
                 finalize: Optional[Callable[[Batch], None]],
                 ) -> None:
        self.stream = stream
        self._compute = compute
        self._finalize = finalize
        # Grad mode is thread-local in PyTorch: capture it where the task is
        # created and re-establish it on the worker thread.
        self.grad_mode = torch.is_grad_enabled()

    def compute(self) -> Batch:
        with use_stream(self.stream):
            return self._compute()

    def finalize(self, batch: Batch) -> None:
        if self._finalize is None:
            return
        with use_stream(self.stream):
            self._finalize(batch)


class Ticket:
    """Completion handle for a submitted task (a one-shot future)."""

    __slots__ = ('_done', 'ok', 'payload')

    def __init__(self) -> None:
        self._done = threading.Event()
        self.ok = False
        self.payload: Union[Tuple[Task, Batch], ExcInfo, None] = None

    def _resolve(self, ok: bool, payload: Union[Tuple[Task, Batch], ExcInfo, None]) -> None:
        self.ok = ok
        self.payload = payload
        self._done.set()

    def result(self) -> Tuple[bool, Union[Tuple[Task, Batch], ExcInfo, None]]:
        self._done.wait()
        return self.ok, self.payload


def normalize_device(device: torch.device) -> torch.device:
    """Maps aliases of one physical device to one key (``cuda`` -> ``cuda:N``, ``cpu:N`` -> ``cpu``)."""
    if getattr(device, 'type', None) == 'cuda' and device.index is None:
        return torch.device('cuda', torch.cuda.current_device())
    if getattr(device, 'type', None) == 'cpu' and device.index is not None:
        return torch.device('cpu')
    return device


class StageExecutor:
    """One long-lived daemon thread bound to one device."""

    def __init__(self, device: torch.device) -> None:
        self.device = device
        self._inbox: 'SimpleQueue[Optional[Tuple[Task, Ticket]]]' = SimpleQueue()
        self._thread = threading.Thread(target=self._serve, name=f'gpipe-stage-{device}',
                                        daemon=True)
        self._closed = False
        self._thread.start()

    def submit(self, task: Task) -> Ticket:
        if self._closed:
            raise RuntimeError('stage executor has been shut down')
        ticket = Ticket()
        self._inbox.put((task, ticket))
        return ticket

    def _serve(self) -> None:
        with use_device(self.device):
            while True:
                item = self._inbox.get()
                if item is None:
                    return
                task, ticket = item
                try:
                    with torch.set_grad_enabled(task.grad_mode):
                        batch = task.compute()
                except Exception:
                    ticket._resolve(False, sys.exc_info())  # type: ignore[arg-type]
                    continue
                ticket._resolve(True, (task, batch))
                del task, ticket, batch, item

    def shutdown(self, wait: bool = True) -> None:
        if self._closed:
            return
        self._closed = True
        self._inbox.put(None)
        if wait and self._thread is not threading.current_thread():
            self._thread.join()

    @property
    def alive(self) -> bool:
        return self._thread.is_alive()


class ExecutorPool:
    """Executors for a list of stage devices; same device -> same executor."""

    def __init__(self, devices: List[torch.device]) -> None:
        self._by_device: Dict[torch.device, StageExecutor] = {}
        self.executors: List[StageExecutor] = []
        for device in devices:
            key = normalize_device(device)
            executor = self._by_device.get(key)
            if executor is None:
                executor = StageExecutor(key)
                self._by_device[key] = executor
            self.executors.append(executor)
        # Stop the threads when the pool is garbage collected.
        self._finalizer = weakref.finalize(self, ExecutorPool._shutdown_all,
                                           list(self._by_device.values()))

    @staticmethod
    def _shutdown_all(executors: List[StageExecutor]) -> None:
        for e in executors:
            e.shutdown(wait=False)

    def __getitem__(self, j: int) -> StageExecutor:
        return self.executors[j]

    def __len__(self) -> int:
        return len(self.executors)

    @property
    def num_threads(self) -> int:
        return len(self._by_device)

    def close(self) -> None:
        for e in self._by_device.values():
            e.shutdown(wait=True)

    def __enter__(self) -> 'ExecutorPool':
        return self

    def __exit__(self, *exc: object) -> None:
        self.close()
