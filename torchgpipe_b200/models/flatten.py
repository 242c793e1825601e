"""Flattening of nested ``nn.Sequential`` containers.

GPipe partitions only the top-level children of a sequential model, so models
are built as nested containers for readability and then flattened; nested
names are joined with ``_`` (``layer1`` > ``0`` > ``conv1`` becomes
``layer1_0_conv1``), the naming of reference
``benchmarks/models/*/flatten_sequential.py``.
"""
from collections import OrderedDict
from typing import Iterator, List, Tuple

from torch import nn

__all__ = ['flatten_sequential']


def _leaves(prefix: List[str], module: nn.Sequential) -> Iterator[Tuple[str, nn.Module]]:
    for name, child in module.named_children():
        path = prefix + [name]
        if isinstance(child, nn.Sequential):
            yield from _leaves(path, child)
        else:
            yield '_'.join(path), child


def flatten_sequential(module: nn.Sequential) -> nn.Sequential:
    """Returns a new ``nn.Sequential`` holding every non-sequential descendant in order."""
    if not isinstance(module, nn.Sequential):
        raise TypeError('not sequential')
    return nn.Sequential(OrderedDict(_leaves([], module)))
