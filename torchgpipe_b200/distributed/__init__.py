"""Multi-process pipeline mode: one OS process (and one GPU) per stage.

The reference's fork-added mode (``torchgpipe/distributed/``) stages every
activation through host memory and ``torch.distributed.rpc``. Here tensors stay
on the device and move with ``torch.distributed`` point-to-point operations
(NCCL over NVLink 5 / NVSwitch between GPUs; gloo for CPU tests); the
``context`` channel API of the reference is kept as the in-process backend.
Checkpointing, cross-stage ``@skippable`` connections and
``DeferredBatchNorm`` - all absent from the reference's distributed mode - work
here as they do in single-process mode.
"""
from torchgpipe_b200.distributed import context  # noqa: F401
from torchgpipe_b200.distributed.gpipe import (DistributedGPipe,  # noqa: F401
                                               DistributedGPipeDataLoader,
                                               get_module_partition)
from torchgpipe_b200.distributed.transport import (ChannelTransport,  # noqa: F401
                                                   ProcessGroupTransport, Transport)

__all__ = ['context', 'DistributedGPipe', 'DistributedGPipeDataLoader', 'get_module_partition',
           'Transport', 'ChannelTransport', 'ProcessGroupTransport']
