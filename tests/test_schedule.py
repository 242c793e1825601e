"""The fill/drain schedule (``runtime/schedule.py``) against an independent simulation, and the
lock-step issue order of the eager engine. Behaviour pinned by the reference in
``tests/test_pipeline.py`` (tick tables for a few shapes, issue order of a slow second stage)."""
import threading
import time

import pytest
import torch
from torch import nn

from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.pipeline import Pipeline, clock_cycles
from torchgpipe_b200.runtime.schedule import StagePlan, bubble_fraction


def simulate(m, n):
    """Greedy simulation of an m-micro-batch, n-stage pipeline with unit-time cells: at every
    tick each stage takes its next micro-batch if the previous stage finished it earlier."""
    done_at = {}
    next_mb = [0] * n
    ticks = []
    tick = 0
    while any(k < m for k in next_mb):
        running = []
        for j in range(n):
            i = next_mb[j]
            if i < m and (j == 0 or done_at.get((i, j - 1), tick) < tick):
                running.append((i, j))
        for i, j in running:
            done_at[(i, j)] = tick
            next_mb[j] += 1
        ticks.append(running)
        tick += 1
    return ticks


@pytest.mark.parametrize('m, n', [(1, 1), (1, 3), (3, 1), (3, 3), (4, 2), (2, 5), (7, 4)])
def test_clock_cycles_equal_the_simulation(m, n):
    assert [list(t) for t in clock_cycles(m, n)] == simulate(m, n)


def test_small_tables_spelled_out():
    assert list(clock_cycles(2, 2)) == [[(0, 0)], [(1, 0), (0, 1)], [(1, 1)]]
    # stages ascend inside a tick, so the copy for the earliest stage is issued first
    for tick in clock_cycles(5, 4):
        assert [j for _, j in tick] == sorted(j for _, j in tick)
        assert len({i + j for i, j in tick}) == 1       # one anti-diagonal per tick


def test_every_cell_exactly_once():
    for m in range(1, 6):
        for n in range(1, 6):
            cells = [c for tick in clock_cycles(m, n) for c in tick]
            assert sorted(cells) == [(i, j) for i in range(m) for j in range(n)]
            assert len(list(clock_cycles(m, n))) == m + n - 1


def test_stage_plan_and_bubble():
    plan = StagePlan.build(4, 2, 3)
    assert len(plan.ticks) == 5
    assert plan.checkpointed(2) and not plan.checkpointed(3)
    assert bubble_fraction(32, 8) == 7 / 39
    assert bubble_fraction(1, 1) == 0.0
    # the bubble is the share of idle cells in the tick table
    for m, n in [(4, 2), (8, 4)]:
        busy = sum(len(t) for t in clock_cycles(m, n))
        assert abs(1 - busy / (n * (m + n - 1)) - bubble_fraction(m, n)) < 1e-12


def test_issue_order_is_tick_by_tick():
    """A fast first stage must not run ahead of the tick structure: cell (i, j) is issued only
    after every cell of the earlier ticks."""
    log = []
    guard = threading.Lock()

    class Stamp(nn.Module):
        def __init__(self, stage, seconds):
            super().__init__()
            self.stage, self.seconds = stage, seconds

        def forward(self, x):
            time.sleep(self.seconds)
            with guard:
                log.append((int(x.flatten()[0].item()), self.stage))
            return x

    m = 4
    batches = [Batch(torch.full((1, 1), float(i))) for i in range(m)]
    partitions = [nn.Sequential(Stamp(0, 0.0)), nn.Sequential(Stamp(1, 0.08)),
                  nn.Sequential(Stamp(2, 0.02))]
    Pipeline(batches, partitions).run()

    assert sorted(log) == [(i, j) for i in range(m) for j in range(3)]
    tick_of = {cell: k for k, tick in enumerate(clock_cycles(m, 3)) for cell in tick}
    finished_tick = [tick_of[cell] for cell in log]
    # the engine collects a whole tick before it issues the next one
    assert finished_tick == sorted(finished_tick)
