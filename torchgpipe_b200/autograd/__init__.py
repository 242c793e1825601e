"""Autograd-graph primitives: ordering edges, stream-aware transfers."""
from torchgpipe_b200.autograd.edges import fork, get_phony, join  # noqa: F401
from torchgpipe_b200.autograd.transfer import Copy, Wait  # noqa: F401
