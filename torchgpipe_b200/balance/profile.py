"""Per-layer profilers behind the automatic balancers.

Each layer is deep-copied to the profiling device in training mode, so the
user's module keeps its device, mode, buffers and (absent) gradients.
Contract: reference ``torchgpipe/balance/profile.py:21-118``. Timing on CUDA
uses CUDA events recorded on the device around forward+backward (the reference
uses the host clock between two device synchronisations).
"""
import copy
import time
from typing import Generator, List, Tuple, Union

import torch
from torch import Tensor
import torch.nn as nn

from torchgpipe_b200.microbatch import Batch

__all__ = ['profile_times', 'profile_sizes', 'layerwise_sandbox']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


def layerwise_sandbox(module: nn.Sequential, device: torch.device,
                      ) -> Generator[nn.Module, None, None]:
    """Yields a throw-away training-mode copy of each layer on ``device``."""
    for layer in module:
        clone = copy.deepcopy(layer)
        clone.to(device)
        clone.train()
        yield clone


def _cut_graph(batch: Batch) -> None:
    for k, x in enumerate(batch):
        batch[k] = x.detach().requires_grad_(x.requires_grad)


def profile_times(module: nn.Sequential, sample: TensorOrTensors, timeout: float,
                  device: torch.device) -> List[int]:
    """Total forward+backward microseconds per layer, accumulated until ``timeout``."""
    if any(p.grad is not None for p in module.parameters()):
        raise ValueError('some parameter already has gradient')

    seed = Batch(sample)
    for k, x in enumerate(seed):
        seed[k] = x.detach().to(device).requires_grad_(x.requires_grad)

    on_cuda = device.type == 'cuda'
    totals_us = [0 for _ in module]
    started = time.time()

    while time.time() - started < timeout:
        batch = seed
        for index, layer in enumerate(layerwise_sandbox(module, device)):
            _cut_graph(batch)

            if on_cuda:
                torch.cuda.synchronize(device)
                begin = torch.cuda.Event(enable_timing=True)
                end = torch.cuda.Event(enable_timing=True)
                begin.record(torch.cuda.current_stream(device))
            tick = time.time()

            batch = batch.call(layer)
            heads = tuple(y for y in batch if y.requires_grad)
            if heads:
                torch.autograd.backward(heads, heads)

            if on_cuda:
                end.record(torch.cuda.current_stream(device))
                torch.cuda.synchronize(device)
                elapsed = begin.elapsed_time(end) * 1e-3
            else:
                elapsed = time.time() - tick

            totals_us[index] += int(elapsed * 1_000_000)

    return totals_us


def profile_sizes(module: nn.Sequential, input: TensorOrTensors, chunks: int,
                  param_scale: float, device: torch.device) -> List[int]:
    """Estimated training bytes per layer (activations of one micro-batch + scaled params)."""
    if device.type != 'cuda':
        raise ValueError('size profiler supports only CUDA device')

    batch = Batch(input)
    samples_per_microbatch = batch[0].size(0) / chunks

    # Profile with a single sample, then scale.
    for k, x in enumerate(batch):
        batch[k] = x[:1].detach().to(device).requires_grad_(x.requires_grad)

    sizes: List[int] = []
    for layer in layerwise_sandbox(module, device):
        _cut_graph(batch)

        before = torch.cuda.memory_allocated(device)
        batch = batch.call(layer)
        after = torch.cuda.memory_allocated(device)
        activation_bytes = after - before

        param_bytes = sum(p.numel() * p.element_size() for p in layer.parameters())

        sizes.append(int(activation_bytes * samples_per_microbatch + param_bytes * param_scale))

    return sizes
