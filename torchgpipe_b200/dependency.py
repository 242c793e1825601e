"""Alias: ``fork``/``join`` live in :mod:`torchgpipe_b200.autograd.edges`."""
from torchgpipe_b200.autograd.edges import Fork, Join, fork, join  # noqa: F401

__all__ = ['fork', 'join', 'Fork', 'Join']
