"""Small fused reductions usable outside the converted models (e.g. DeferredBatchNorm)."""
import torch
from torch import Tensor

from torchgpipe_b200.ops import _lib

__all__ = ['available', 'channel_moments']


def available() -> bool:
    return torch.cuda.is_available() and _lib.load() is not None


def channel_moments(input: Tensor):
    """Per-channel (sum, sum of squares) of an ``[N,C,H,W]`` fp32 CUDA tensor in one pass.

    Uses the native kernel for channels-last fp32 inputs with ``C % 4 == 0``; anything else
    takes the two ATen reductions the reference uses (``torchgpipe/batchnorm.py:51-57``).
    """
    if (input.dtype == torch.float32 and input.dim() == 4 and input.size(1) % 4 == 0
            and input.is_contiguous(memory_format=torch.channels_last)):
        from torchgpipe_b200.ops.functional import channel_moments as native
        m = native(input)
        return m[:, 0], m[:, 1]
    dims = [0] + list(range(2, input.dim()))
    return input.sum(dims), (input * input).sum(dims)
