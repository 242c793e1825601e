"""Pipeline x data-parallel layout: ``stages`` pipeline stages, ``replicas`` copies of the pipeline.

Global rank ``g`` = ``replica * stages + stage``. The ranks of one replica form a pipeline (they
exchange activations through a :class:`ProcessGroupTransport` restricted to that replica); the
ranks holding the same stage in different replicas form a data-parallel group whose gradients are
all-reduced by the fused kernel as soon as that stage has finished its backward pass. Later stages
finish first, so their all-reduce overlaps the backward drain of the earlier stages.
"""
from typing import Any, Dict, Iterable, List, Optional

import torch
import torch.distributed as dist
from torch import nn

from torchgpipe_b200.parallel.buckets import GradientBucket

__all__ = ['HybridLayout', 'HybridPipeline']


class HybridLayout:
    """Rank arithmetic and process groups of an ``S x R`` hybrid."""

    def __init__(self, stages: int, replicas: int) -> None:
        world = dist.get_world_size()
        if stages * replicas != world:
            raise ValueError(f'stages x replicas = {stages * replicas} but world size is {world}')
        self.stages, self.replicas = stages, replicas
        self.global_rank = dist.get_rank()
        self.replica, self.stage = divmod(self.global_rank, stages)
        # every rank takes part in creating every group
        self.dp_groups = [dist.new_group([r * stages + s for r in range(replicas)])
                          for s in range(stages)]
        self.dp_group = self.dp_groups[self.stage]

    def stage_to_global(self) -> Dict[int, int]:
        """Stage rank -> global rank inside this rank's replica."""
        return {s: self.replica * self.stages + s for s in range(self.stages)}


class HybridPipeline:
    """A :class:`DistributedGPipe` per replica plus the fused gradient all-reduce per stage."""

    def __init__(self, module: nn.Sequential, balance: Iterable[int], chunks: int,
                 layout: HybridLayout, device: torch.device, *, checkpoint: str = 'except_last',
                 cuda_graphs: bool = False, use_symmetric: bool = True,
                 check_allreduce: bool = True) -> None:
        from torchgpipe_b200.distributed import DistributedGPipe, ProcessGroupTransport
        self.layout = layout
        workers = {s: f'replica{layout.replica}-stage{s}' for s in range(layout.stages)}
        transport = ProcessGroupTransport(device, ranks=layout.stage_to_global())
        self.pipe = DistributedGPipe(module, layout.stage, workers, list(balance), chunks,
                                     device=device, checkpoint=checkpoint, transport=transport,
                                     cuda_graphs=cuda_graphs)
        self.bucket = GradientBucket(self.pipe.model().parameters(), layout.dp_group,
                                     use_symmetric=use_symmetric)
        #: verify after every all-reduce that no replica timed out (one host sync per step, at
        #: the point where the optimiser needs the reduced gradients anyway)
        self.check_allreduce = check_allreduce

    def model(self) -> nn.Sequential:
        return self.pipe.model()

    def forward(self, batch: Optional[Any]) -> List[Any]:
        return self.pipe.forward(batch)

    def backward(self, losses: Optional[List[torch.Tensor]]) -> None:
        """Backward of this stage for all micro-batches, then the stage's gradient all-reduce."""
        self.pipe.backward(losses)
        self.bucket.allreduce()
        if self.check_allreduce:
            self.bucket.check()

    def zero_grad(self) -> None:
        self.bucket.zero()
