"""Alias: ``get_phony`` lives in :mod:`torchgpipe_b200.autograd.edges`."""
from torchgpipe_b200.autograd.edges import get_phony  # noqa: F401

__all__ = ['get_phony']
