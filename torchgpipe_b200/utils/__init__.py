"""Small helpers shared by the runtime (tracing hooks)."""
from torchgpipe_b200.utils import nvtx  # noqa: F401

__all__ = ['nvtx']
