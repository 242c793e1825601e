"""Alias of :mod:`torchgpipe_b200.runtime.pipeline` under the reference's module path."""
from torchgpipe_b200.runtime.pipeline import Pipeline, clock_cycles, copy, depend, wait  # noqa: F401

__all__ = ['Pipeline', 'clock_cycles', 'depend', 'copy', 'wait']
