"""Static map of which stage stashes and which stage pops each skip tensor.

Built once per :class:`~torchgpipe_b200.GPipe` from the partitions; the
scheduler asks it, for each destination stage, which skip tensors must be
moved there and from where. Skips whose stash and pop live on the same stage
never need a portal. Contract: reference ``torchgpipe/skip/layout.py:11-83``.
"""
from typing import Dict, Iterable, List, Sequence, Tuple

from torch import nn

from torchgpipe_b200.skip.namespace import Namespace

__all__ = ['SkipLayout', 'inspect_skip_layout']

SkipKey = Tuple[Namespace, str]


class SkipLayout:
    """Skip routes ``(ns, name) -> (source stage, destination stage)``."""

    def __init__(self, num_partitions: int, skip_routes: Dict[SkipKey, Tuple[int, int]]) -> None:
        self.by_ns_name = skip_routes
        self.by_partition: List[List[Tuple[int, Namespace, str]]] = [
            [] for _ in range(num_partitions)
        ]
        for (ns, name), (src, dst) in skip_routes.items():
            self.by_partition[dst].append((src, ns, name))
        for routes in self.by_partition:
            routes.sort()

    def copy_policy(self, next_j: int) -> Iterable[Tuple[int, Namespace, str]]:
        """Routes that end at stage ``next_j`` and start elsewhere, by ascending source."""
        for src, ns, name in self.by_partition[next_j]:
            if src != next_j:
                yield (src, ns, name)

    def requires_copy(self, ns: Namespace, name: str) -> bool:
        """Whether this skip crosses a stage boundary (and so needs a portal)."""
        route = self.by_ns_name.get((ns, name))
        if route is None:
            return False
        return route[0] != route[1]

    def cross_routes(self) -> List[Tuple[int, int, Namespace, str]]:
        """All cross-stage routes as ``(src, dst, ns, name)`` (used by the distributed mode)."""
        out = []
        for (ns, name), (src, dst) in self.by_ns_name.items():
            if src != dst:
                out.append((src, dst, ns, name))
        out.sort(key=lambda r: (r[0], r[1], r[3]))
        return out


def inspect_skip_layout(partitions: Sequence[nn.Sequential]) -> SkipLayout:
    """Walks the partitions in order and pairs each pop with its earlier stash."""
    from torchgpipe_b200.skip.skippable import Skippable

    pending: Dict[SkipKey, int] = {}
    routes: Dict[SkipKey, Tuple[int, int]] = {}
    for j, partition in enumerate(partitions):
        for layer in partition:
            if not isinstance(layer, Skippable):
                continue
            for key in layer.stashable():
                pending[key] = j
            for key in layer.poppable():
                routes[key] = (pending.pop(key), j)
    return SkipLayout(len(partitions), routes)
