"""GPipe fill/drain schedule.

``clock_cycles(m, n)`` enumerates, clock tick by clock tick, the
``(micro-batch i, stage j)`` cells that may run concurrently: the anti-diagonal
wavefront ``i + j == k``. Same tables as reference
``torchgpipe/pipeline.py:49-65`` (pinned by ``tests/test_pipeline.py:10-29``).

``StagePlan`` precomputes everything about the schedule that is static for a
given ``(m, n, checkpoint_stop)`` so the per-forward Python work is a walk
over prebuilt lists.
"""
from typing import Iterator, List, NamedTuple, Tuple

__all__ = ['clock_cycles', 'StagePlan', 'bubble_fraction']

Cell = Tuple[int, int]


def clock_cycles(m: int, n: int) -> Iterator[List[Cell]]:
    """Yields the list of ``(i, j)`` cells of each of the ``m + n - 1`` ticks."""
    for k in range(m + n - 1):
        first_stage = max(k - m + 1, 0)
        last_stage = min(k, n - 1)
        yield [(k - j, j) for j in range(first_stage, last_stage + 1)]


def bubble_fraction(m: int, n: int) -> float:
    """Ideal idle fraction of a fill/drain pipeline with equal stage times: (n-1)/(m+n-1)."""
    if m <= 0 or n <= 0:
        return 0.0
    return (n - 1) / (m + n - 1)


class StagePlan(NamedTuple):
    """Static description of one forward pass of the pipeline."""

    m: int
    n: int
    checkpoint_stop: int
    ticks: Tuple[Tuple[Cell, ...], ...]

    @classmethod
    def build(cls, m: int, n: int, checkpoint_stop: int) -> 'StagePlan':
        ticks = tuple(tuple(cells) for cells in clock_cycles(m, n))
        return cls(m, n, checkpoint_stop, ticks)

    def checkpointed(self, i: int) -> bool:
        return i < self.checkpoint_stop
