"""Per-cell CUDA-event timeline: bubble and exposed-communication accounting.

When a :class:`Timeline` is attached to a pipeline run, every
``(micro-batch, stage)`` cell records a CUDA event on the stage's compute
stream before and after its kernels are enqueued, and an event before/after the
boundary copy on the side stream. After a synchronise, :meth:`Timeline.report`
turns them into

* ``busy_ms[j]``   - time stage ``j`` spent inside cells,
* ``span_ms``      - first cell start to last cell end over all stages,
* ``bubble_frac``  - ``1 - sum(busy) / (n * span)`` (idle share of the pipeline),
* ``exposed_comm_ms`` - time a stage's compute stream sat waiting for an
  inbound copy that had not finished when the stage was otherwise ready.

The reference has no tracer; its only hooks are
``is_checkpointing()/is_recomputing()``. CPU stages record host timestamps.
"""
import time
from typing import Dict, List, Optional, Tuple

import torch

from torchgpipe_b200.runtime.streams import AbstractStream, as_cuda, is_cuda

__all__ = ['Timeline']


class _Mark:
    __slots__ = ('event', 'host')

    def __init__(self, stream: AbstractStream) -> None:
        self.host = time.perf_counter()
        if is_cuda(stream):
            self.event: Optional[torch.cuda.Event] = torch.cuda.Event(enable_timing=True)
            self.event.record(as_cuda(stream))
        else:
            self.event = None


class Timeline:
    """Collects ``(phase, i, j)`` intervals for one or more pipeline runs."""

    def __init__(self) -> None:
        self._open: Dict[Tuple[str, int, int], _Mark] = {}
        self.intervals: List[Tuple[str, int, int, _Mark, _Mark]] = []

    def begin(self, phase: str, i: int, j: int, stream: AbstractStream) -> None:
        self._open[(phase, i, j)] = _Mark(stream)

    def end(self, phase: str, i: int, j: int, stream: AbstractStream) -> None:
        start = self._open.pop((phase, i, j), None)
        if start is None:
            return
        self.intervals.append((phase, i, j, start, _Mark(stream)))

    def clear(self) -> None:
        self._open.clear()
        self.intervals.clear()

    @staticmethod
    def _elapsed(a: _Mark, b: _Mark) -> float:
        if a.event is not None and b.event is not None:
            return a.event.elapsed_time(b.event)
        return (b.host - a.host) * 1e3

    def report(self, n_stages: int) -> Dict[str, object]:
        """Summarises the recorded run. Call after synchronising all devices."""
        busy = [0.0] * n_stages
        comm = [0.0] * n_stages
        first: List[Optional[_Mark]] = [None] * n_stages
        last: List[Optional[_Mark]] = [None] * n_stages
        for phase, i, j, a, b in self.intervals:
            dt = self._elapsed(a, b)
            if phase == 'compute':
                busy[j] += dt
                if first[j] is None:
                    first[j] = a
                last[j] = b
            elif phase == 'copy':
                comm[j] += dt
        spans = []
        for j in range(n_stages):
            if first[j] is not None and last[j] is not None:
                spans.append(self._elapsed(first[j], last[j]))  # type: ignore[arg-type]
            else:
                spans.append(0.0)
        span = max(spans) if spans else 0.0
        total_busy = sum(busy)
        bubble = 1.0 - total_busy / (n_stages * span) if span > 0 else 0.0
        # Exposed communication: per stage, span minus busy is idle; the share of it
        # that coincides with an in-flight inbound copy is bounded by the copy time.
        exposed = [min(c, max(s - b, 0.0)) for c, s, b in zip(comm, spans, busy)]
        return {
            'busy_ms': busy,
            'stage_span_ms': spans,
            'span_ms': span,
            'bubble_frac': max(bubble, 0.0),
            'copy_ms': comm,
            'exposed_comm_ms': exposed,
        }
