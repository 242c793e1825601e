"""Optional NVTX ranges around pipeline cells (the reference has no tracer).

Off by default (zero cost: the context manager is a shared no-op object). ``TGB_NVTX=1`` in the
environment, or :func:`enable`, turns on ``torch.cuda.nvtx`` ranges named
``fwd mb<i> stage<j>`` / ``bwd mb<i> stage<j>`` / ``recompute ...`` that show up in Nsight Systems
timelines next to the kernels they launched.
"""
import os
from contextlib import contextmanager
from typing import Iterator

import torch

__all__ = ['enable', 'enabled', 'annotate']

_enabled = os.environ.get('TGB_NVTX', '0') not in ('', '0')


def enable(flag: bool = True) -> None:
    global _enabled
    _enabled = bool(flag)


def enabled() -> bool:
    return _enabled and torch.cuda.is_available()


class _Noop:
    def __enter__(self) -> None:
        return None

    def __exit__(self, *exc: object) -> bool:
        return False


_NOOP = _Noop()


@contextmanager
def _range(name: str) -> Iterator[None]:
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def annotate(name: str):
    """``with annotate('fwd mb3 stage1'): ...`` - an NVTX range when enabled, else nothing."""
    return _range(name) if enabled() else _NOOP
