"""Per-layer cost measurements behind the automatic balancers.

What is measured follows reference ``torchgpipe/balance/profile.py:21-118`` (forward+backward
time per layer; activation bytes of one micro-batch plus scaled parameter bytes per layer, on
throw-away training-mode copies so the user's module keeps its device, mode, buffers and absent
gradients). How it is measured is different:

* the copies live in a :class:`LayerLab` that is built ONCE per profiling run. The reference
  deep-copies and uploads every layer again on every pass of its timeout loop; for a
  multi-billion-parameter model that upload dominates what is being timed. Layers stay resident
  on the device while they fit a share of its free memory and are streamed one at a time beyond
  that;
* on CUDA every layer's forward+backward is captured into a CUDA graph and timed on replays, so
  the numbers are device time even where Python cannot launch kernels as fast as a B200 runs
  them (``TGB_PROFILE_MODE=eager``, or a layer that cannot be captured, falls back to one event
  pair per layer around eager calls with a single synchronisation per pass - the reference uses
  a host clock between two device-wide synchronisations per layer);
* the per-layer result is the median over the passes (times the number of passes, to keep the
  reference's "accumulated microseconds" scale), so a lazy-initialisation outlier in the first
  pass does not tilt the balance.
"""
import copy
import gc
import os
import statistics
import time
from typing import Iterator, List, Optional, Tuple, Union

import torch
from torch import Tensor
import torch.nn as nn

from torchgpipe_b200.microbatch import Batch

__all__ = ['profile_times', 'profile_sizes', 'layerwise_sandbox', 'LayerLab']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


def _nbytes(layer: nn.Module) -> int:
    own = list(layer.parameters()) + list(layer.buffers())
    return sum(t.numel() * t.element_size() for t in own)


class LayerLab:
    """Training-mode copies of a sequential model's layers on the profiling device.

    ``resident_share`` of the device's free memory (CUDA) may be spent on keeping copies alive
    between passes; layers beyond that budget are copied each time they are visited and dropped
    right after (``streamed``). On CPU everything is resident.
    """

    def __init__(self, module: nn.Sequential, device: torch.device,
                 resident_share: float = 0.25) -> None:
        self.device = device
        self.sources: List[nn.Module] = list(module)
        budget: Optional[int] = None
        if device.type == 'cuda':
            free, _ = torch.cuda.mem_get_info(device)
            budget = int(free * resident_share)
        self.copies: List[Optional[nn.Module]] = []
        self.streamed = 0
        spent = 0
        for layer in self.sources:
            need = _nbytes(layer)
            if budget is not None and spent + need > budget:
                self.copies.append(None)
                self.streamed += 1
                continue
            spent += need
            self.copies.append(self._clone(layer))

    def _clone(self, layer: nn.Module) -> nn.Module:
        return copy.deepcopy(layer).to(self.device).train()

    def __len__(self) -> int:
        return len(self.sources)

    def __iter__(self) -> Iterator[nn.Module]:
        for source, kept in zip(self.sources, self.copies):
            yield kept if kept is not None else self._clone(source)

    def clear_gradients(self) -> None:
        for kept in self.copies:
            if kept is not None:
                for p in kept.parameters():
                    p.grad = None


def layerwise_sandbox(module: nn.Sequential, device: torch.device) -> Iterator[nn.Module]:
    """One training-mode copy of each layer on ``device``, in order (a single pass over a
    :class:`LayerLab`; kept for callers of the reference's helper of this name)."""
    return iter(LayerLab(module, device))


def _fresh_leaves(values: TensorOrTensors, device: Optional[torch.device] = None,
                  first_n: Optional[int] = None) -> Batch:
    """A batch of leaf tensors holding ``values`` (graph cut, ``requires_grad`` preserved)."""
    batch = Batch(values)
    for k, x in enumerate(batch):
        y = x.detach()
        if first_n is not None:
            y = y[:first_n]
        if device is not None:
            y = y.to(device)
        batch[k] = y.requires_grad_(x.requires_grad)
    return batch


def _forward_backward(layer: nn.Module, batch: Batch) -> Batch:
    out = batch.call(layer)
    heads = tuple(y for y in out if y.requires_grad)
    if heads:
        torch.autograd.backward(heads, heads)       # any cotangent of the right shape will do
    return out


def _one_pass_cuda(lab: LayerLab, seed: Batch) -> List[float]:
    device = lab.device
    stream = torch.cuda.current_stream(device)
    marks = []
    batch = seed
    for layer in lab:
        batch = _fresh_leaves(batch.tensor_or_tensors)
        begin = torch.cuda.Event(enable_timing=True)
        end = torch.cuda.Event(enable_timing=True)
        begin.record(stream)
        batch = _forward_backward(layer, batch)
        end.record(stream)
        marks.append((begin, end))
    torch.cuda.synchronize(device)
    return [b.elapsed_time(e) * 1e-3 for b, e in marks]


def _graphed_times(lab: LayerLab, seed: Batch, replays: int) -> Optional[List[List[float]]]:
    """Per-layer GPU time of forward+backward measured on CUDA-graph replays.

    On a B200 most layers of the benchmark models finish faster than Python can launch their
    kernels, so events around eager calls measure the host, not the device, and a balance built
    from them is off. Each layer's forward+backward is therefore captured once into a CUDA graph
    (static input = the static output of the previous layer's graph) and replayed ``replays``
    times between two events: launch gaps disappear and what remains is device time.
    Returns ``replays`` samples per layer, or ``None`` when some layer cannot be captured (it
    synchronises with the host, for instance) - the caller then falls back to eager timing.
    """
    from torchgpipe_b200.skip.tracker import SkipTracker, use_skip_tracker

    class _KeepingTracker(SkipTracker):
        """Every layer runs twice here (eagerly, then under capture): a popped skip tensor
        stays available for the second run."""

        def load(self, batch, ns, name):        # type: ignore[override]
            return self.tensors[(ns, name)]

    device = lab.device
    stream = torch.cuda.Stream(device)
    graphs, marks = [], []
    batch = seed
    try:
        with torch.cuda.device(device), use_skip_tracker(_KeepingTracker()):
            stream.wait_stream(torch.cuda.current_stream(device))
            for layer in lab:
                batch = _fresh_leaves(batch.tensor_or_tensors)
                with torch.cuda.stream(stream):            # lazy initialisation outside capture
                    _forward_backward(layer, batch)
                stream.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=stream, capture_error_mode='relaxed'):
                    out = _forward_backward(layer, batch)
                samples = []
                with torch.cuda.stream(stream):
                    graph.replay()                          # outputs hold real values
                    for _ in range(replays):
                        begin = torch.cuda.Event(enable_timing=True)
                        end = torch.cuda.Event(enable_timing=True)
                        begin.record(stream)
                        graph.replay()
                        end.record(stream)
                        samples.append((begin, end))
                graphs.append(graph)                        # keeps the static tensors alive
                marks.append(samples)
                batch = out
            stream.synchronize()
            times = [[b.elapsed_time(e) * 1e-3 for b, e in samples] for samples in marks]
    except Exception:
        torch.cuda.synchronize(device)
        return None
    finally:
        del graphs
    return times


def _one_pass_host(lab: LayerLab, seed: Batch) -> List[float]:
    seconds = []
    batch = seed
    for layer in lab:
        batch = _fresh_leaves(batch.tensor_or_tensors)
        tick = time.perf_counter()
        batch = _forward_backward(layer, batch)
        seconds.append(time.perf_counter() - tick)
    return seconds


def profile_times(module: nn.Sequential, sample: TensorOrTensors, timeout: float,
                  device: torch.device) -> List[int]:
    """Forward+backward cost of every layer in microseconds: ``median over passes x passes``,
    with passes repeated until ``timeout`` seconds have gone by (at least one)."""
    if any(p.grad is not None for p in module.parameters()):
        raise ValueError('some parameter already has gradient')

    lab = LayerLab(module, device)
    seed = _fresh_leaves(sample, device)
    one_pass = _one_pass_cuda if device.type == 'cuda' else _one_pass_host

    if device.type == 'cuda' and os.environ.get('TGB_PROFILE_MODE', 'graph') == 'graph' \
            and lab.streamed == 0:
        graphed = _graphed_times(lab, seed, replays=5)
        lab.clear_gradients()
        if graphed is not None:
            return [max(1, int(statistics.median(bucket) * len(bucket) * 1_000_000))
                    for bucket in graphed]

    samples: List[List[float]] = [[] for _ in range(len(lab))]
    deadline = time.monotonic() + timeout
    passes = 0
    while passes == 0 or time.monotonic() < deadline:
        for bucket, seconds in zip(samples, one_pass(lab, seed)):
            bucket.append(seconds)
        lab.clear_gradients()
        passes += 1
    return [int(statistics.median(bucket) * passes * 1_000_000) for bucket in samples]


def profile_sizes(module: nn.Sequential, input: TensorOrTensors, chunks: int,
                  param_scale: float, device: torch.device) -> List[int]:
    """Estimated training bytes per layer: what one sample's forward leaves allocated, scaled
    to a micro-batch, plus ``param_scale`` times the layer's parameter bytes."""
    if device.type != 'cuda':
        raise ValueError('size profiler supports only CUDA device')

    # garbage from earlier work must not be freed in the middle of a measurement: the allocator
    # counter would drop and a layer would seem to release memory
    gc.collect()
    # cuBLAS takes its per-stream workspace from the caching allocator on the first matmul of a
    # stream; get that out of the way so it is not billed to the first layer that multiplies
    warm = torch.zeros(8, 8, device=device)
    torch.mm(warm, warm)
    torch.nn.functional.linear(warm, warm, warm[0])     # addmm: the cuBLASLt workspace
    del warm
    whole = Batch(input)
    per_microbatch = whole[0].size(0) / chunks
    batch = _fresh_leaves(input, device, first_n=1)     # measure with one sample, then scale

    # stream the layers: only the layer under measurement may add to the allocator's count
    lab = LayerLab(module, device, resident_share=0.0)
    sizes: List[int] = []
    for layer in lab:
        batch = _fresh_leaves(batch.tensor_or_tensors)
        held_before = torch.cuda.memory_allocated(device)
        batch = batch.call(layer)
        grown = max(0, torch.cuda.memory_allocated(device) - held_before)
        weights = sum(p.numel() * p.element_size() for p in layer.parameters())
        sizes.append(int(grown * per_microbatch + weights * param_scale))
    return sizes
