import threading
import time

import torch
from torch import nn

from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.pipeline import Pipeline, clock_cycles
from torchgpipe_b200.runtime.schedule import StagePlan, bubble_fraction


def test_clock_cycles_tables():
    assert list(clock_cycles(1, 1)) == [[(0, 0)]]
    assert list(clock_cycles(1, 3)) == [[(0, 0)], [(0, 1)], [(0, 2)]]
    assert list(clock_cycles(3, 1)) == [[(0, 0)], [(1, 0)], [(2, 0)]]
    assert list(clock_cycles(3, 3)) == [
        [(0, 0)],
        [(1, 0), (0, 1)],
        [(2, 0), (1, 1), (0, 2)],
        [(2, 1), (1, 2)],
        [(2, 2)],
    ]
    assert list(clock_cycles(4, 2)) == [
        [(0, 0)],
        [(1, 0), (0, 1)],
        [(2, 0), (1, 1)],
        [(3, 0), (2, 1)],
        [(3, 1)],
    ]


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


def test_lockstep_order():
    timeline = []
    lock = threading.Lock()

    class Delayed(nn.Module):
        def __init__(self, seconds, j):
            super().__init__()
            self.seconds, self.j = seconds, j

        def forward(self, x):
            time.sleep(self.seconds)
            with lock:
                timeline.append((int(x.item()), self.j))
            return x

    batches = [Batch(torch.tensor(float(i))) for i in range(3)]
    partitions = [nn.Sequential(Delayed(0.03, 0)), nn.Sequential(Delayed(0.15, 1))]
    Pipeline(batches, partitions).run()

    # Stage 0 is much faster, but a clock tick only ends when all of its
    # cells have been issued, so (2,0) cannot overtake (0,1).
    assert timeline == [(0, 0), (1, 0), (0, 1), (2, 0), (1, 1), (2, 1)]
