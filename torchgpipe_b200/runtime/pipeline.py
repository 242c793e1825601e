"""The clock-cycle driver: fence (transfers + ordering edges) then compute.

One ``Pipeline.run()`` walks the static :class:`StagePlan`. Per tick it

1. *fences* on the scheduling thread: adds the ``i-1 <- i`` backward-ordering
   edge, moves cross-stage skip tensors into the destination stage, and moves
   the micro-batch across the stage boundary - all on side streams, so the
   copy for cell ``(i, j)`` overlaps the kernels of cell ``(i-1, j)``;
2. *computes*: hands one :class:`Task` per cell to that stage's persistent
   executor, then collects results in schedule order, orders the side stream
   behind the compute stream, and attaches recomputation.

The autograd graph this builds per cell (with checkpointing) is::

    Copy(j-1 -> j) -> Wait(copy_j -> compute_j) -> Checkpoint -> Wait(compute_j -> copy_j)
                   -> Fork -+----------------> Join -> [next fence]
                            +--> Recompute ----^

so the backward pipeline (including early recomputation) is executed by
PyTorch's autograd engine with no scheduler code of our own.

Behavioural contract: reference ``torchgpipe/pipeline.py:96-249`` (lock-step
order ``tests/test_pipeline.py:32-62``, first-exception-wins and early stop
``tests/test_gpipe.py:227-275``). Differences by design: executors are
persistent (reference spawns threads per call, ``pipeline.py:112``) and side
streams come from a fixed pool.
"""
from typing import List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor, nn

from torchgpipe_b200.autograd.edges import fork, join
from torchgpipe_b200.autograd.transfer import Copy, Wait
from torchgpipe_b200.checkpoint import Checkpointing
from torchgpipe_b200.microbatch import Batch, check
from torchgpipe_b200.runtime.executor import ExcInfo, ExecutorPool, Task, Ticket
from torchgpipe_b200.runtime.schedule import StagePlan, clock_cycles
from torchgpipe_b200.runtime.streams import AbstractStream, current_stream, use_device
from torchgpipe_b200.runtime.timeline import Timeline
from torchgpipe_b200.skip.layout import SkipLayout, inspect_skip_layout
from torchgpipe_b200.skip.tracker import SkipTrackerThroughPotals, use_skip_tracker
from torchgpipe_b200.utils import nvtx

__all__ = ['Pipeline', 'clock_cycles', 'depend', 'copy', 'wait']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


def depend(fork_from: Batch, join_to: Batch) -> None:
    """Makes the backward of ``join_to`` precede the backward of ``fork_from``."""
    fork_from[0], phony = fork(fork_from[0])
    join_to[0] = join(join_to[0], phony)


def copy(batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream) -> None:
    batch[:] = Copy.apply(prev_stream, next_stream, *batch)


def wait(batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream) -> None:
    batch[:] = Wait.apply(prev_stream, next_stream, *batch)


class Pipeline:
    """Runs ``batches`` through ``partitions`` in fill/drain order, in place."""

    def __init__(self,
                 batches: List[Batch],
                 partitions: Sequence[nn.Sequential],
                 devices: Optional[List[torch.device]] = None,
                 copy_streams: Optional[Sequence[Sequence[AbstractStream]]] = None,
                 skip_layout: Optional[SkipLayout] = None,
                 checkpoint_stop: int = 0,
                 *,
                 executors: Optional[ExecutorPool] = None,
                 timeline: Optional[Timeline] = None,
                 ) -> None:
        self.batches = batches
        self.partitions = partitions
        if devices is None:
            devices = [torch.device('cpu') for _ in partitions]
        self.devices = devices
        if copy_streams is None:
            copy_streams = [[current_stream(d)] * len(batches) for d in devices]
        self.copy_streams = copy_streams
        if skip_layout is None:
            skip_layout = inspect_skip_layout(partitions)
        self.skip_layout = skip_layout
        self.checkpoint_stop = checkpoint_stop
        self.executors = executors
        self.timeline = timeline

    def run(self) -> None:
        m, n = len(self.batches), len(self.partitions)
        plan = StagePlan.build(m, n, self.checkpoint_stop)
        trackers = [SkipTrackerThroughPotals(self.skip_layout) for _ in self.batches]

        pool = self.executors
        transient = pool is None
        if transient:
            pool = ExecutorPool(self.devices)
        try:
            for cells in plan.ticks:
                self.fence(cells, trackers)
                self.compute(cells, trackers, pool)
        finally:
            if transient:
                pool.close()

    # -- phase 1 -----------------------------------------------------------
    def fence(self, schedule: Sequence[Tuple[int, int]],
              skip_trackers: List[SkipTrackerThroughPotals]) -> None:
        batches, streams, layout = self.batches, self.copy_streams, self.skip_layout
        timeline = self.timeline
        for i, j in schedule:
            if i != 0:
                depend(batches[i - 1], batches[i])

            next_stream = streams[j][i]
            for prev_j, ns, name in layout.copy_policy(j):
                skip_trackers[i].copy(batches[i], streams[prev_j][i], next_stream, ns, name)

            if j != 0:
                if timeline is not None:
                    timeline.begin('copy', i, j, next_stream)
                copy(batches[i], streams[j - 1][i], next_stream)
                if timeline is not None:
                    timeline.end('copy', i, j, next_stream)

    # -- phase 2 -----------------------------------------------------------
    def compute(self, schedule: Sequence[Tuple[int, int]],
                skip_trackers: List[SkipTrackerThroughPotals],
                pool: ExecutorPool) -> None:
        batches, partitions, devices = self.batches, self.partitions, self.devices
        copy_streams, stop, timeline = self.copy_streams, self.checkpoint_stop, self.timeline
        n = len(partitions)
        compute_streams = [current_stream(d) for d in devices]

        tickets: List[Ticket] = []
        for i, j in schedule:
            batch = batches[i]
            if j != 0:
                wait(batch, copy_streams[j][i], compute_streams[j])
            task = self._make_task(i, j, batch, partitions[j], skip_trackers[i],
                                   compute_streams[j], i < stop, timeline)
            tickets.append(pool[j].submit(task))

        failure: Optional[ExcInfo] = None
        for (i, j), ticket in zip(schedule, tickets):
            ok, payload = ticket.result()
            if failure is not None:
                continue            # keep draining so no worker is left mid-task
            if not ok:
                failure = payload   # type: ignore[assignment]
                continue

            task, batch = payload   # type: ignore[misc]
            try:
                check(batch.value)  # stage boundaries carry tensors only
            except TypeError:
                import sys
                failure = sys.exc_info()  # type: ignore[assignment]
                continue
            if j != n - 1:
                wait(batch, compute_streams[j], copy_streams[j][i])
            with use_device(devices[j]):
                task.finalize(batch)
            batches[i] = batch

        if failure is not None:
            raise failure[0].with_traceback(failure[1], failure[2])  # type: ignore[misc]

    @staticmethod
    def _make_task(i: int, j: int, batch: Batch, partition: nn.Sequential,
                   tracker: SkipTrackerThroughPotals, stream: AbstractStream,
                   checkpointed: bool, timeline: Optional[Timeline]) -> Task:
        if checkpointed:
            def function(input: TensorOrTensors) -> TensorOrTensors:
                with use_skip_tracker(tracker):
                    return partition(input)

            chk = Checkpointing(function, batch)
            run, finalize = chk.checkpoint, chk.recompute
        else:
            def run() -> Batch:
                with use_skip_tracker(tracker):
                    return batch.call(partition)

            finalize = None

        if nvtx.enabled():
            plain_run, plain_finalize = run, finalize

            def run() -> Batch:                                     # type: ignore[no-redef]
                with nvtx.annotate(f'fwd mb{i} stage{j}'):
                    return plain_run()

            if plain_finalize is not None:
                def finalize(b: Batch) -> None:                     # type: ignore[no-redef,misc]
                    with nvtx.annotate(f'arm recompute mb{i} stage{j}'):
                        plain_finalize(b)

        if timeline is None:
            return Task(stream, compute=run, finalize=finalize)

        def timed() -> Batch:
            timeline.begin('compute', i, j, stream)
            try:
                return run()
            finally:
                timeline.end('compute', i, j, stream)

        return Task(stream, compute=timed, finalize=finalize)
