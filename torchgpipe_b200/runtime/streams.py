"""Device/stream helpers that treat CPU as a degenerate "stream".

Every scheduler and transfer primitive in this package talks to streams only
through the functions below, so the whole pipeline also runs on
``devices=['cpu', 'cpu', ...]`` (the property the reference gets from its
``CPUStream`` placeholder, reference ``torchgpipe/stream.py:12-101``).

B200 notes: compute runs on each device's *current* stream; stage-boundary and
skip transfers run on a small fixed pool of side streams per device (see
:class:`StreamPool`) instead of one stream per micro-batch
(reference ``torchgpipe/gpipe.py:316-328`` creates ``chunks`` streams/device).
"""
from contextlib import contextmanager
from typing import Generator, List, Union

import torch

__all__ = ['CPUStream', 'CPUStreamType', 'AbstractStream', 'new_stream', 'current_stream',
           'default_stream', 'use_device', 'use_stream', 'get_device', 'wait_stream',
           'record_stream', 'is_cuda', 'as_cuda', 'StreamPool']


class CPUStreamType:
    """Singleton type standing in for "the stream of a CPU device"."""

    _instance = None

    def __new__(cls) -> 'CPUStreamType':
        if cls._instance is None:
            cls._instance = super().__new__(cls)
        return cls._instance

    def __repr__(self) -> str:
        return '<CPUStream>'


CPUStream = CPUStreamType()

AbstractStream = Union[torch.cuda.Stream, CPUStreamType]


def is_cuda(stream: AbstractStream) -> bool:
    """True when ``stream`` is a real CUDA stream."""
    return stream is not CPUStream


def as_cuda(stream: AbstractStream) -> torch.cuda.Stream:
    """Narrows an abstract stream to ``torch.cuda.Stream`` (no runtime check)."""
    return stream  # type: ignore[return-value]


def new_stream(device: torch.device) -> AbstractStream:
    """A fresh side stream on ``device`` (the CPU placeholder for CPU devices)."""
    if device.type != 'cuda':
        return CPUStream
    return torch.cuda.Stream(device)


def current_stream(device: torch.device) -> AbstractStream:
    if device.type != 'cuda':
        return CPUStream
    return torch.cuda.current_stream(device)


def default_stream(device: torch.device) -> AbstractStream:
    if device.type != 'cuda':
        return CPUStream
    return torch.cuda.default_stream(device)


@contextmanager
def use_device(device: torch.device) -> Generator[None, None, None]:
    """``torch.cuda.device`` that is a no-op for CPU devices."""
    if device.type != 'cuda':
        yield
        return
    with torch.cuda.device(device):
        yield


@contextmanager
def use_stream(stream: AbstractStream) -> Generator[None, None, None]:
    """``torch.cuda.stream`` that is a no-op for the CPU placeholder."""
    if not is_cuda(stream):
        yield
        return
    with torch.cuda.stream(as_cuda(stream)):
        yield


def get_device(stream: AbstractStream) -> torch.device:
    if is_cuda(stream):
        return as_cuda(stream).device
    return torch.device('cpu')


def wait_stream(source: AbstractStream, target: AbstractStream) -> None:
    """Makes ``source`` wait for everything queued on ``target`` so far.

    cuda<-cuda enqueues an event wait; cpu<-cuda blocks the host; anything
    waiting on the CPU placeholder is already ordered (host code is serial).
    """
    if not is_cuda(target):
        return
    if is_cuda(source):
        as_cuda(source).wait_stream(as_cuda(target))
    else:
        as_cuda(target).synchronize()


def record_stream(tensor: torch.Tensor, stream: AbstractStream) -> None:
    """Tells the caching allocator that ``tensor`` is in use on ``stream``.

    ``Tensor.record_stream`` marks the whole storage block; that is what we
    want for views as well (reference regression ``tests/test_stream.py:151-172``).
    """
    if is_cuda(stream):
        tensor.record_stream(as_cuda(stream))


class StreamPool:
    """A fixed, small set of side streams per device, indexed by micro-batch.

    ``pool[j][i]`` is the transfer stream used for micro-batch ``i`` at stage
    ``j``. Micro-batches share ``width`` physical streams round-robin, which
    is enough to overlap the transfer of micro-batch ``i+1`` with ``i`` while
    keeping the number of live CUDA streams independent of ``chunks``.
    """

    def __init__(self, devices: List[torch.device], chunks: int, width: int = 4) -> None:
        self.width = max(1, min(width, chunks))
        self._physical: List[List[AbstractStream]] = [
            [new_stream(d) for _ in range(self.width)] for d in devices
        ]
        self.chunks = chunks

    def __len__(self) -> int:
        return len(self._physical)

    def __getitem__(self, j: int) -> '_StreamRow':
        return _StreamRow(self._physical[j])


class _StreamRow:
    __slots__ = ('_streams',)

    def __init__(self, streams: List[AbstractStream]) -> None:
        self._streams = streams

    def __getitem__(self, i: int) -> AbstractStream:
        return self._streams[i % len(self._streams)]

    def __len__(self) -> int:
        return len(self._streams)
