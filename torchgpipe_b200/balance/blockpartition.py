"""Contiguous block partition of a cost sequence (Barany & Grinberg,
"Block Partitions of Sequences", arXiv:1308.2452).

Start from equal-length blocks and repeatedly slide one boundary element from
the neighbour towards the currently lightest block, walking from the heaviest
block's side, until the heaviest and lightest block differ by at most one
(costs normalised to [0, 1]). Not optimal, but O(k n^3) and good enough for
stage balancing. Contract: reference ``torchgpipe/balance/blockpartition.py:11-89``
(e.g. ``solve([1,2,3,4,5,6], 2) == [[1,2,3,4],[5,6]]``).
"""
from typing import List

__all__ = ['solve', 'solve_optimal']


def solve(sequence: List[int], partitions: int = 1) -> List[List[int]]:
    """Splits ``sequence`` into ``partitions`` contiguous blocks of similar sum."""
    if partitions < 1:
        raise ValueError(f'partitions must be a positive integer ({partitions} < 1)')

    n = len(sequence)
    if n < partitions:
        raise ValueError(f'sequence is shorter than intended partitions ({n} < {partitions})')

    lo = min(sequence)
    spread = max(sequence) - lo
    if spread == 0:
        weights = [0.0] * n
    else:
        weights = [(x - lo) / spread for x in sequence]

    # ends[b] is the exclusive end index of block b.
    ends = [n // partitions * (b + 1) for b in range(partitions - 1)] + [n]

    def weight_of(b: int) -> float:
        begin = ends[b - 1] if b > 0 else 0
        return sum(weights[begin:ends[b]])

    def blocks() -> List[List[int]]:
        begins = [0] + ends[:-1]
        return [sequence[a:z] for a, z in zip(begins, ends)]

    def spread_of() -> float:
        ws = [weight_of(b) for b in range(partitions)]
        return max(ws) - min(ws)

    def non_empty() -> bool:
        begins = [0] + ends[:-1]
        return all(a < z for a, z in zip(begins, ends))

    # The sliding procedure can cycle, or push a boundary past an emptied block, on some inputs
    # (the reference then never returns, e.g. solve([8,2,8,3,2,13,2,5,2,2,3,13], 10)): remember
    # the states seen and, on a repeat or an inconsistent state, answer with the most even state
    # without an empty block that was visited.
    seen = set()
    best_ends, best_spread = list(ends), spread_of()

    while True:
        # Heaviest block (ties -> highest index), as in the paper's step (1).
        heavy_weight, heavy = max((weight_of(b), b) for b in range(partitions))

        while True:
            light_weight, light = min((weight_of(b), b) for b in range(partitions))
            if heavy_weight <= light_weight + 1:
                return blocks()

            state = (tuple(ends), heavy, heavy_weight)
            consistent = all(0 <= z <= n for z in ends)
            if state in seen or not consistent:
                ends[:] = best_ends
                return blocks()
            seen.add(state)
            if non_empty():
                spread = spread_of()
                if spread < best_spread or not _all_non_empty(best_ends):
                    best_ends, best_spread = list(ends), spread

            if heavy < light:
                # Give the last element of the block left of `light` to `light`.
                neighbour = light - 1
                ends[neighbour] -= 1
            else:
                # Give the first element of the block right of `light` to `light`.
                neighbour = light + 1
                ends[light] += 1

            if neighbour == heavy:
                break   # the heaviest block changed: re-evaluate from step (1)


def _all_non_empty(ends: List[int]) -> bool:
    begins = [0] + list(ends[:-1])
    return all(a < z for a, z in zip(begins, ends))


def solve_optimal(sequence: List[float], partitions: int = 1) -> List[List[float]]:
    """Contiguous split of ``sequence`` into ``partitions`` non-empty blocks that minimises the
    heaviest block (the pipeline's bottleneck stage), ties broken by the smallest sum of squared
    block sums (evens out the other stages). Exact dynamic programme, O(partitions * n^2).

    Not in the reference: :func:`solve` (Barany & Grinberg's sliding heuristic) stays the default
    of the balancers for parity; pass ``solver='optimal'`` to ``balance_by_time`` /
    ``balance_by_size`` to use this one.
    """
    if partitions < 1:
        raise ValueError(f'partitions must be a positive integer ({partitions} < 1)')
    n = len(sequence)
    if n < partitions:
        raise ValueError(f'sequence is shorter than intended partitions ({n} < {partitions})')

    prefix = [0.0]
    for x in sequence:
        prefix.append(prefix[-1] + x)

    def block(a: int, z: int) -> float:
        return prefix[z] - prefix[a]

    inf = (float('inf'), float('inf'))
    # best[k][i]: (bottleneck, sum of squares) of the best split of the first i items into k blocks
    best = [[inf] * (n + 1) for _ in range(partitions + 1)]
    cut = [[0] * (n + 1) for _ in range(partitions + 1)]
    best[0][0] = (0.0, 0.0)
    for k in range(1, partitions + 1):
        for i in range(k, n - (partitions - k) + 1):
            for j in range(k - 1, i):
                head = best[k - 1][j]
                if head == inf:
                    continue
                last = block(j, i)
                cand = (max(head[0], last), head[1] + last * last)
                if cand < best[k][i]:
                    best[k][i] = cand
                    cut[k][i] = j
    ends = []
    i = n
    for k in range(partitions, 0, -1):
        ends.append(i)
        i = cut[k][i]
    ends.reverse()
    begins = [0] + ends[:-1]
    return [sequence[a:z] for a, z in zip(begins, ends)]
