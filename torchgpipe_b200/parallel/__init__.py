"""Pipeline x data-parallel hybrid: replica groups, symmetric gradient buckets and the fused
NVLink / NVSwitch gradient all-reduce (absent from the reference, required by the north star)."""
from torchgpipe_b200.parallel.allreduce import SymmetricBuffer, fused_allreduce_
from torchgpipe_b200.parallel.buckets import GradientBucket
from torchgpipe_b200.parallel.hybrid import HybridLayout, HybridPipeline

__all__ = ['SymmetricBuffer', 'fused_allreduce_', 'GradientBucket', 'HybridLayout',
           'HybridPipeline']
