"""Rough automatic balancing of an ``nn.Sequential`` into pipeline stages.

::

    from torchgpipe_b200 import GPipe
    from torchgpipe_b200.balance import balance_by_time

    sample = torch.empty(128, 3, 224, 224)
    balance = balance_by_time(torch.cuda.device_count(), model, sample)
    gpipe = GPipe(model, balance, chunks=8)

Both balancers profile each layer in isolation on one device (the module
itself is left untouched) and then split the per-layer costs into contiguous
blocks of near-equal total with :func:`blockpartition.solve`.
API contract: reference ``torchgpipe/balance/__init__.py:38-156``.
"""
from typing import List, Tuple, Union

import torch
from torch import Tensor
import torch.nn as nn

from torchgpipe_b200.balance import blockpartition
from torchgpipe_b200.balance.profile import profile_sizes, profile_times

__all__ = ['balance_by_time', 'balance_by_size', 'balance_cost']

Device = Union[torch.device, int, str]
Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


def balance_cost(cost: List[int], partitions: int, solver: str = 'reference') -> List[int]:
    """Number of layers per stage for the given per-layer costs.

    ``solver='reference'``: the reference's sliding heuristic (:func:`blockpartition.solve`);
    ``solver='optimal'``: exact min-max split (:func:`blockpartition.solve_optimal`).
    """
    if solver not in ('reference', 'optimal'):
        raise ValueError("solver is not one of 'reference' or 'optimal'")
    solve = blockpartition.solve if solver == 'reference' else blockpartition.solve_optimal
    return [len(block) for block in solve(cost, partitions)]


def balance_by_time(partitions: int,
                    module: nn.Sequential,
                    sample: TensorOrTensors,
                    *,
                    timeout: float = 1.0,
                    device: Device = torch.device('cuda'),
                    solver: str = 'reference',
                    ) -> List[int]:
    """Balances by measured forward+backward time of each layer.

    Args:
        partitions: number of stages wanted.
        module: sequential module to balance.
        sample: example input (any batch size), on the same device as ``module``.

    Keyword Args:
        timeout: keep re-profiling until this many seconds have passed (default 1.0).
        device: CPU or CUDA device to profile on (default: current CUDA device).
            On CUDA, layers are timed with CUDA events on the device, not with
            the host clock.
        solver: ``'reference'`` (default, the reference's heuristic) or ``'optimal'``
            (exact min-max contiguous split; never yields an empty stage).

    Returns:
        list of layer counts, one per stage, for ``GPipe(balance=...)``.
    """
    times = profile_times(module, sample, timeout, torch.device(device))
    return balance_cost(times, partitions, solver)


def balance_by_size(partitions: int,
                    module: nn.Sequential,
                    input: TensorOrTensors,
                    *,
                    chunks: int = 1,
                    param_scale: float = 2.0,
                    device: Device = torch.device('cuda'),
                    solver: str = 'reference',
                    ) -> List[int]:
    """Balances by CUDA memory each layer needs for training.

    Per layer: ``activation bytes per sample x (batch / chunks) + parameter
    bytes x param_scale``. ``param_scale`` accounts for what the optimiser keeps
    per parameter besides the gradient:

    =========  =============  =========================================
    Optimizer  `param_scale`  Internal state
    =========  =============  =========================================
    SGD        2--3           (momentum_buffer)
    Adam       4--5           exp_avg, exp_avg_sq, (max_exp_avg_sq)
    Adadelta   4              square_avg, acc_delta
    Adagrad    3              sum
    RMSprop    3--5           square_avg, (momentum_buffer), (grad_avg)
    =========  =============  =========================================

    Args:
        partitions: number of stages wanted.
        module: sequential module to balance.
        input: example mini-batch of the size used in training.

    Keyword Args:
        chunks: number of micro-batches used in training (default 1).
        param_scale: copies of the parameters held during training (default 2.0).
        device: CUDA device to profile on (CUDA only).
        solver: as in :func:`balance_by_time`.

    Returns:
        list of layer counts, one per stage.
    """
    sizes = profile_sizes(module, input, chunks, param_scale, torch.device(device))
    return balance_cost(sizes, partitions, solver)
