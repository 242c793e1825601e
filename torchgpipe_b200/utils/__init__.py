"""Small helpers shared by the runtime (tracing hooks, measured machine constants)."""
from torchgpipe_b200.utils import nvtx  # noqa: F401
from torchgpipe_b200.utils.hw import link_bandwidth_gbs

__all__ = ['nvtx', 'link_bandwidth_gbs']
