"""Symmetric-memory buffers and the fused scale + all-reduce over them.

``SymmetricBuffer`` allocates the same number of fp32 elements on every rank of a process group
through ``torch.distributed._symmetric_memory`` (cuMem allocations exported to the peers; a
multicast object on top when the NVSwitch supports it) plus a small zeroed signal pad.
``fused_allreduce_`` launches ``tgb_allreduce_scale_f32`` (``csrc/allreduce.cu``): one kernel per
rank that reduces over NVLink peer loads - or ``multimem.ld_reduce`` / ``multimem.st`` through
the switch - and applies the gradient scale (and an optional bf16 cast) on the way.

NCCL ``all_reduce`` on the same bucket is the baseline these numbers are compared with
(``benchmarks/allreduce_bench.py``); it is also the fallback when symmetric memory cannot be
set up (single rank, CPU, or a platform without peer access).
"""
import ctypes
import os
from typing import Any, Optional

import torch
import torch.distributed as dist
from torch import Tensor

from torchgpipe_b200.ops import _lib

__all__ = ['SymmetricBuffer', 'fused_allreduce_']

_MAX_BLOCKS = 148


class SymmetricBuffer:
    """``numel`` fp32 elements, addressable by every rank of ``group``."""

    def __init__(self, numel: int, group: Any = None, device: Optional[torch.device] = None,
                 ) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.device = device if device is not None else torch.device('cuda',
                                                                     torch.cuda.current_device())
        align = 4 * self.world
        self.numel = (numel + align - 1) // align * align
        enable = getattr(symm_mem, 'enable_symm_mem_for_group', None)
        if enable is not None:
            try:
                import warnings
                with warnings.catch_warnings():
                    warnings.simplefilter('ignore')
                    enable(self.group.group_name)
            except Exception:
                pass            # newer releases enable groups implicitly
        self.tensor = symm_mem.empty(self.numel, dtype=torch.float32, device=self.device)
        self.tensor.zero_()
        self.handle = symm_mem.rendezvous(self.tensor, self.group)
        lib = _lib.require()
        lib.tgb_allreduce_pad_bytes.restype = ctypes.c_ulonglong
        pad_words = int(lib.tgb_allreduce_pad_bytes(ctypes.c_int(_MAX_BLOCKS))) // 4
        self.pad = symm_mem.empty(pad_words, dtype=torch.int32, device=self.device)
        self.pad.zero_()
        self.pad_handle = symm_mem.rendezvous(self.pad, self.group)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.launch_index = 0
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)        # every pad is zeroed before anyone signals

        self._bufs = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        self._pads = (ctypes.c_void_p * self.world)(*[int(p) for p in
                                                      self.pad_handle.buffer_ptrs])
        # 0 when the allocation has no multicast object bound to it (no NVSwitch multicast on
        # this platform, or a backend that does not create one)
        try:
            mc = int(self.handle.multicast_ptr or 0)
        except Exception:       # attribute absent in other releases: treat as unsupported
            mc = 0
        self.multicast_ptr = mc

    @property
    def has_multicast(self) -> bool:
        return self.multicast_ptr != 0

    def check(self) -> None:
        """Raises if a fused all-reduce on this buffer gave up waiting for a peer.

        The kernel returns early (leaving the buffer partially reduced) when a peer misses a
        barrier for longer than ``timeout_ms`` and records that in ``error_flag``. Reading the
        flag synchronises with the stream the all-reduce ran on, so call it once per step,
        before ``optimizer.step()``. The flag is cleared when the error is raised."""
        if int(self.error_flag.item()) != 0:
            self.error_flag.zero_()
            raise RuntimeError(
                f'fused all-reduce timed out on rank {self.rank} of {self.world}: a peer did '
                'not reach the barrier; the gradients in this buffer are not fully reduced')


def fused_allreduce_(buf: SymmetricBuffer, scale: float = 1.0, cast_out: Optional[Tensor] = None,
                     multicast: Optional[bool] = None, timeout_ms: int = 20000,
                     max_blocks: int = _MAX_BLOCKS) -> Tensor:
    """In place: ``buf.tensor = scale * sum over ranks``; returns ``buf.tensor``.

    Enqueued on the current stream; every rank of ``buf.group`` must call it (same order).
    ``cast_out``: optional local bf16 tensor of ``buf.numel`` elements receiving a copy.
    ``multicast``: ``True`` selects the NVLS variant (``multimem.ld_reduce`` / ``multimem.st``
    through the switch), ``False`` the peer load / store variant. Default: NVLS when the buffer
    has a multicast mapping and there are more than two replicas (measured on 4 B200: 2.36 ms vs
    2.78 ms P2P vs 2.87 ms NCCL for 1 GiB; with two replicas P2P is the faster one, 1.69 vs 2.65
    ms); ``TGB_ALLREDUCE_NVLS=0`` turns the default off.
    Call ``buf.check()`` before using the result for an optimiser step.
    """
    lib = _lib.require()
    if multicast is None:
        multicast = buf.has_multicast and buf.world > 2 and \
            os.environ.get('TGB_ALLREDUCE_NVLS', '1') != '0'
    if multicast and not buf.has_multicast:
        raise RuntimeError('multicast requested but this buffer has no multicast mapping')
    if cast_out is not None:
        assert cast_out.dtype == torch.bfloat16 and cast_out.numel() >= buf.numel
    with torch.cuda.device(buf.device):
        rc = lib.tgb_allreduce_scale_f32(
            buf._bufs, buf._pads, ctypes.c_void_p(buf.multicast_ptr if multicast else 0),
            _lib.ptr(cast_out), ctypes.c_ulonglong(buf.numel), ctypes.c_float(scale),
            ctypes.c_int(buf.rank), ctypes.c_int(buf.world), ctypes.c_uint(buf.launch_index),
            ctypes.c_int(timeout_ms), _lib.ptr(buf.error_flag), ctypes.c_int(max_blocks),
            _lib.stream_ptr(buf.device))
    _lib.check(rc, 'tgb_allreduce_scale_f32')
    buf.launch_index += 1
    return buf.tensor
