"""torchgpipe_b200: a B200-native (sm_100a) GPipe for PyTorch.

Same public API and ``state_dict`` format as torchgpipe v0.0.7::

    from torchgpipe_b200 import GPipe
    model = GPipe(nn.Sequential(...), balance=[...], chunks=8)

or, for code written against the reference, ``torchgpipe_b200.compat.install()``
makes ``import torchgpipe`` resolve to this package.
"""
from torchgpipe_b200.__version__ import __version__  # noqa: F401
from torchgpipe_b200.checkpoint import is_checkpointing, is_recomputing
from torchgpipe_b200.gpipe import GPipe

__all__ = ['GPipe', 'is_checkpointing', 'is_recomputing']
