"""Alias: ``Copy``/``Wait`` live in :mod:`torchgpipe_b200.autograd.transfer`."""
from torchgpipe_b200.autograd.transfer import Copy, Wait  # noqa: F401

__all__ = ['Copy', 'Wait']
