"""Hand-written sm_100a operators and the module rewriter that installs them.

``convert_to_sm100(module)`` rewrites a (flat) ``nn.Sequential`` so its hot layers run on the
kernels in ``csrc/`` while keeping the number, order and names of the top-level layers (so
``balance`` lists and ``state_dict`` keys are unchanged).
"""
import torch
from torch import nn

from torchgpipe_b200.ops import _lib

__all__ = ['available', 'launch_count', 'reset_launch_count', 'convert_to_sm100']


def available() -> bool:
    """True when the native library is loaded and a CUDA device is present."""
    return torch.cuda.is_available() and _lib.load() is not None


def launch_count() -> int:
    """Kernels launched by the native library in this process so far."""
    lib = _lib.load()
    return int(lib.tgb_launch_count()) if lib is not None else 0


def reset_launch_count() -> None:
    lib = _lib.load()
    if lib is not None:
        lib.tgb_reset_launch_count()


def convert_to_sm100(module: nn.Module) -> nn.Module:
    """Installs the fused sm_100a operators into ``module`` (in place where possible)."""
    from torchgpipe_b200.ops.convert import convert_to_sm100 as _convert
    return _convert(module)
