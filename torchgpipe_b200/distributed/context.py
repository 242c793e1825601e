"""In-process message channels between pipeline workers.

Every worker (stage) registers a named :class:`TrainingContext` that owns one
forward queue and one backward queue per micro-batch plus a target queue.
``put_*`` functions are what a *remote* peer executes to deliver a value;
``get_*`` block until the value is there. This is the CPU / single-process
backend (tests, debugging, or an RPC layer calling ``put_*`` remotely); the
GPU path uses :class:`~torchgpipe_b200.distributed.transport.ProcessGroupTransport`.

API contract: reference ``torchgpipe/distributed/context.py:19-193``. One
addition: ``get_*`` give up after ``timeout`` seconds (default
:func:`default_timeout`: ``TGB_CHANNEL_TIMEOUT``, else 1800 s; ``0`` or ``inf``
= wait forever like the reference, ``context.py:141``) and raise
:class:`ChannelTimeout`, so a dead peer ends the job instead of hanging it.
"""
import math
import os
from contextlib import contextmanager
from functools import wraps
from queue import Empty, Queue
from typing import Any, Callable, Dict, Generator, List, Optional, Tuple, Union

from torch import Tensor

__all__ = ['TrainingContext', 'GlobalContext', 'worker', 'distributed', 'put_forward',
           'get_forward', 'put_backward', 'get_backward', 'put_target', 'get_target',
           'ChannelTimeout', 'default_timeout']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


class ChannelTimeout(TimeoutError, Empty):
    """Nothing arrived on a channel in time (also catchable as ``queue.Empty``)."""


def default_timeout() -> Optional[float]:
    """Seconds a ``get_*`` waits when no ``timeout`` is given; ``None`` = forever."""
    raw = os.environ.get('TGB_CHANNEL_TIMEOUT', '1800')
    try:
        value = float(raw)
    except ValueError:
        return 1800.0
    return None if value <= 0 or math.isinf(value) else value


def _take(queue: Queue, timeout: Optional[float], what: str) -> Any:
    if timeout is None:
        timeout = default_timeout()
    try:
        return queue.get(timeout=timeout)
    except Empty:
        raise ChannelTimeout(f'no {what} within {timeout:g} s: the peer stage died or never '
                             'sent it (TGB_CHANNEL_TIMEOUT changes the limit)') from None


class TrainingContext:
    """Mailboxes of one worker."""

    def __init__(self, context_name: str, microbatch_chunks: int) -> None:
        self.name = context_name
        self.forward_channels: List[Queue] = [Queue() for _ in range(microbatch_chunks)]
        self.backward_channels: List[Queue] = [Queue() for _ in range(microbatch_chunks)]
        self.target_channel: Queue = Queue()


class GlobalContext:
    """Process-wide registry ``worker name -> TrainingContext``."""

    ctxs: Dict[str, TrainingContext] = {}

    @staticmethod
    def get_context(context_name: str) -> TrainingContext:
        return GlobalContext.ctxs[context_name]     # KeyError for unknown workers


@contextmanager
def worker(context_name: str, microbatch_chunks: int) -> Generator[None, None, None]:
    """Registers a worker for the duration of the ``with`` block.

    ::

        with worker('worker0', chunks):
            train()
    """
    if context_name in GlobalContext.ctxs:
        raise RuntimeError(f'worker {context_name} already exists')
    GlobalContext.ctxs[context_name] = TrainingContext(context_name, microbatch_chunks)
    try:
        yield
    finally:
        GlobalContext.ctxs.pop(context_name, None)


def distributed(context_name: str, microbatch_chunks: int) -> Callable[[Callable], Callable]:
    """Decorator form of :func:`worker`; decorators for different names nest."""
    def decorator(func: Callable) -> Callable:
        @wraps(func)
        def forwarder(*args: Any, **kwargs: Any) -> Any:
            with worker(context_name, microbatch_chunks):
                return func(*args, **kwargs)
        return forwarder
    return decorator


def put_forward(context_name: str, id: int, value: TensorOrTensors) -> None:
    """Delivers the input of micro-batch ``id`` to worker ``context_name``."""
    GlobalContext.get_context(context_name).forward_channels[id].put(value)


def get_forward(name: str, id: int, timeout: Optional[float] = None) -> TensorOrTensors:
    """Blocks until the input of micro-batch ``id`` has been delivered to ``name``."""
    return _take(GlobalContext.get_context(name).forward_channels[id], timeout,
                 f'input of micro-batch {id} for worker {name!r}')


def put_backward(name: str, id: int, value: TensorOrTensors) -> None:
    """Delivers the output-gradient of micro-batch ``id`` to worker ``name``."""
    GlobalContext.get_context(name).backward_channels[id].put(value)


def get_backward(name: str, id: int, timeout: Optional[float] = None) -> TensorOrTensors:
    return _take(GlobalContext.get_context(name).backward_channels[id], timeout,
                 f'gradient of micro-batch {id} for worker {name!r}')


def put_target(name: str, value: TensorOrTensors) -> None:
    """Delivers a mini-batch of targets to worker ``name`` (the last stage)."""
    GlobalContext.get_context(name).target_channel.put(value)


def get_target(name: str, timeout: Optional[float] = None) -> TensorOrTensors:
    return _take(GlobalContext.get_context(name).target_channel, timeout,
                 f'target for worker {name!r}')
