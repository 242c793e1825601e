"""Alias of :mod:`torchgpipe_b200.runtime.streams` under the reference's module path."""
from torchgpipe_b200.runtime.streams import *  # noqa: F401,F403
from torchgpipe_b200.runtime.streams import __all__  # noqa: F401
