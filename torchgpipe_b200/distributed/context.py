"""In-process message channels between pipeline workers.

Every worker (stage) registers a named :class:`TrainingContext` that owns one
forward queue and one backward queue per micro-batch plus a target queue.
``put_*`` functions are what a *remote* peer executes to deliver a value;
``get_*`` block until the value is there. This is the CPU / single-process
backend (tests, debugging, or an RPC layer calling ``put_*`` remotely); the
GPU path uses :class:`~torchgpipe_b200.distributed.transport.ProcessGroupTransport`.

API contract: reference ``torchgpipe/distributed/context.py:19-193``. One
addition: ``get_*`` take an optional ``timeout`` (seconds) so a dead peer
raises ``queue.Empty`` instead of blocking forever.
"""
from contextlib import contextmanager
from functools import wraps
from queue import Queue
from typing import Any, Callable, Dict, Generator, List, Optional, Tuple, Union

from torch import Tensor

__all__ = ['TrainingContext', 'GlobalContext', 'worker', 'distributed', 'put_forward',
           'get_forward', 'put_backward', 'get_backward', 'put_target', 'get_target']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


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
    return GlobalContext.get_context(name).forward_channels[id].get(timeout=timeout)


def put_backward(name: str, id: int, value: TensorOrTensors) -> None:
    """Delivers the output-gradient of micro-batch ``id`` to worker ``name``."""
    GlobalContext.get_context(name).backward_channels[id].put(value)


def get_backward(name: str, id: int, timeout: Optional[float] = None) -> TensorOrTensors:
    return GlobalContext.get_context(name).backward_channels[id].get(timeout=timeout)


def put_target(name: str, value: TensorOrTensors) -> None:
    """Delivers a mini-batch of targets to worker ``name`` (the last stage)."""
    GlobalContext.get_context(name).target_channel.put(value)


def get_target(name: str, timeout: Optional[float] = None) -> TensorOrTensors:
    return GlobalContext.get_context(name).target_channel.get(timeout=timeout)
