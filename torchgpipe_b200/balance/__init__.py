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


def balance_by_time(partitions: int, module: nn.Sequential, sample: TensorOrTensors, *,
                    timeout: float = 1.0, device: Device = torch.device('cuda'),
                    solver: str = 'reference') -> List[int]:
    """Layer counts per stage such that the stages take about equally long.

    Every layer's forward+backward is measured in isolation on ``device`` with ``sample`` as the
    model input (any batch size; use one *micro*-batch to balance for the shapes the pipeline
    will actually run). On a CUDA device the measurement is device time taken on CUDA-graph
    replays of each layer - see :mod:`torchgpipe_b200.balance.profile` for why event pairs around
    eager calls are not good enough on a B200 - and ``timeout`` only bounds the eager fallback,
    which repeats its passes until that many seconds are over. On CPU the host clock is used.

    ``solver='reference'`` splits the measured costs with the reference's heuristic;
    ``solver='optimal'`` with an exact min-max contiguous split that never leaves a stage empty.
    The result goes straight into ``GPipe(module, balance=...)``. ``module`` itself is not
    moved, run or modified.
    """
    times = profile_times(module, sample, timeout, torch.device(device))
    return balance_cost(times, partitions, solver)


def balance_by_size(partitions: int, module: nn.Sequential, input: TensorOrTensors, *,
                    chunks: int = 1, param_scale: float = 2.0,
                    device: Device = torch.device('cuda'),
                    solver: str = 'reference') -> List[int]:
    """Layer counts per stage such that the stages need about equally much device memory.

    The estimate for a layer is::

        bytes its forward leaves allocated for ONE sample x (len(input) / chunks)
        + bytes of its parameters x param_scale

    ``input`` is a mini-batch of the training size and ``chunks`` the number of micro-batches, so
    the first term is what one micro-batch keeps alive. ``param_scale`` says how many
    parameter-sized tensors training holds per parameter: the parameter, its gradient, and
    whatever the optimiser keeps -

    * plain SGD: 2 (parameter + gradient), 3 with momentum;
    * Adagrad: 3 (its running sum); RMSprop: 3, up to 5 with momentum and centred variance;
    * Adadelta: 4 (two running averages);
    * Adam / AdamW: 4 (two moments), 5 with ``amsgrad``.

    CUDA devices only (the numbers come from the caching allocator's counters). ``solver`` as in
    :func:`balance_by_time`.
    """
    sizes = profile_sizes(module, input, chunks, param_scale, torch.device(device))
    return balance_cost(sizes, partitions, solver)
