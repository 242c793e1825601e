"""Per-thread registry of live skip tensors.

Outside a pipeline a plain dictionary is enough (``SkipTracker``): ``stash``
stores, ``pop`` removes. Inside :class:`~torchgpipe_b200.GPipe` every
micro-batch gets a ``SkipTrackerThroughPotals`` [sic - name kept for API
compatibility with the reference], which routes *cross-stage* skips through
:class:`~torchgpipe_b200.skip.portal.Portal` objects and leaves same-stage skips
in the dictionary. Contract: reference ``torchgpipe/skip/tracker.py:19-179``.
"""
from contextlib import contextmanager
import threading
from typing import Dict, Generator, Optional, Tuple

from torch import Tensor

from torchgpipe_b200.autograd.edges import fork, join
from torchgpipe_b200.checkpoint import is_checkpointing
from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.runtime.streams import AbstractStream
from torchgpipe_b200.skip.layout import SkipLayout
from torchgpipe_b200.skip.namespace import Namespace
from torchgpipe_b200.skip.portal import Portal

__all__ = ['SkipTracker', 'SkipTrackerThroughPotals', 'use_skip_tracker',
           'current_skip_tracker']

SkipKey = Tuple[Namespace, str]


class SkipTracker:
    """Dictionary-backed tracker used when no pipeline is involved."""

    def __init__(self) -> None:
        self.tensors: Dict[SkipKey, Optional[Tensor]] = {}

    def save(self, batch: Batch, ns: Namespace, name: str, tensor: Optional[Tensor]) -> None:
        self.tensors[(ns, name)] = tensor

    def load(self, batch: Batch, ns: Namespace, name: str) -> Optional[Tensor]:
        return self.tensors.pop((ns, name))

    def copy(self, batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream,
             ns: Namespace, name: str) -> None:
        raise TypeError('copy is not supported for non-portal skip tensors')


class SkipTrackerThroughPotals(SkipTracker):
    """Tracker of one micro-batch inside a pipeline.

    A skip whose stash and pop sit on the same stage stays in the dictionary. One that crosses
    stages gets a :class:`Portal`, and every operation on it threads a phony edge through the
    first tensor of the batch so that autograd orders the portal's nodes with the stage's own
    computation (``fork`` takes the edge out of the batch, ``join`` puts one back in).
    """

    def __init__(self, skip_layout: SkipLayout) -> None:
        super().__init__()
        self.skip_layout = skip_layout
        self.portals: Dict[SkipKey, Portal] = {}

    def _crosses_stages(self, ns: Namespace, name: str) -> bool:
        return self.skip_layout.requires_copy(ns, name)

    def save(self, batch: Batch, ns: Namespace, name: str, tensor: Optional[Tensor]) -> None:
        if not self._crosses_stages(ns, name):
            super().save(batch, ns, name, tensor)
            return
        key = (ns, name)
        if key in self.portals:
            # the stashing stage is being recomputed: the portal exists, only its blue node
            # will look at the tensor again
            portal = self.portals[key]
            portal.put_tensor(tensor, 1)
        else:
            # first forward: blue and orange each use the tensor once, and orange a second
            # time if the popping stage is going to be recomputed
            uses = 2 + int(is_checkpointing())
            portal = self.portals[key] = Portal(tensor, uses)
        batch[0] = join(batch[0], portal.blue())

    def load(self, batch: Batch, ns: Namespace, name: str) -> Optional[Tensor]:
        if not self._crosses_stages(ns, name):
            return super().load(batch, ns, name)
        batch[0], edge = fork(batch[0])
        return self.portals[(ns, name)].orange(edge)

    def copy(self, batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream,
             ns: Namespace, name: str) -> None:
        """Moves the portal's tensor to the popping stage's device, on the copy streams."""
        assert self._crosses_stages(ns, name)
        batch[0], edge = fork(batch[0])
        arrived = self.portals[(ns, name)].copy(prev_stream, next_stream, edge)
        batch[0] = join(batch[0], arrived)


class _Current(threading.local):
    def __init__(self) -> None:
        self.tracker: Optional[SkipTracker] = None


_current = _Current()


@contextmanager
def use_skip_tracker(skip_tracker: SkipTracker) -> Generator[None, None, None]:
    """Installs ``skip_tracker`` as this thread's tracker for the duration."""
    previous, _current.tracker = _current.tracker, skip_tracker
    try:
        yield
    finally:
        _current.tracker = previous


def current_skip_tracker() -> SkipTracker:
    """This thread's tracker (a plain :class:`SkipTracker` is created on first use)."""
    if _current.tracker is None:
        _current.tracker = SkipTracker()
    return _current.tracker
