"""Activation checkpointing whose recomputation can start *before* the
gradient arrives.

``torch.utils.checkpoint`` re-runs the forward inside the same autograd node
that then back-propagates, so recomputation cannot begin until the upstream
gradient has been produced (and, in a pipeline, copied over NVLink). Here the
two halves are separate autograd nodes that share a one-slot mailbox:

* ``Recompute`` hangs off a phony branch of the stage output. It has no data
  dependency on the incoming gradient, so the autograd engine may run it as
  soon as the stage's device thread is free: the re-forward of micro-batch
  ``i`` on GPU ``j`` overlaps with the gradient transfer from GPU ``j+1``.
* ``Checkpoint`` sits on the main lane behind the gradient ``Wait``. It takes
  the recomputed graph out of the mailbox and back-propagates through it.

RNG state (CPU + the stage's CUDA device) is captured at the first forward and
restored around recomputation so dropout masks are identical.

Contract: reference ``torchgpipe/checkpoint.py:1-308``.
"""
from collections import deque
from contextlib import contextmanager
import threading
from typing import Callable, Deque, Generator, List, Optional, Tuple, Union

import torch
from torch import Tensor

from torchgpipe_b200.autograd.edges import fork, get_phony, join
from torchgpipe_b200.microbatch import Batch

__all__ = ['is_checkpointing', 'is_recomputing', 'checkpoint', 'Checkpointing']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]
Function = Callable[[TensorOrTensors], TensorOrTensors]


class _Phase(threading.local):
    def __init__(self) -> None:
        self.checkpointing = False
        self.recomputing = False


_phase = _Phase()


@contextmanager
def enable_checkpointing() -> Generator[None, None, None]:
    previous, _phase.checkpointing = _phase.checkpointing, True
    try:
        yield
    finally:
        _phase.checkpointing = previous


@contextmanager
def enable_recomputing() -> Generator[None, None, None]:
    previous, _phase.recomputing = _phase.recomputing, True
    try:
        yield
    finally:
        _phase.recomputing = previous


def is_checkpointing() -> bool:
    """True while a partition runs its first, graph-less forward under checkpointing."""
    return _phase.checkpointing


def is_recomputing() -> bool:
    """True while a partition is being re-run during backward.

    Use it to suppress side effects that must happen once per micro-batch
    (e.g. statistics tracking in :class:`~torchgpipe_b200.batchnorm.DeferredBatchNorm`).
    """
    return _phase.recomputing


class _RNGSnapshot:
    """CPU and (optionally) one CUDA device's generator state."""

    __slots__ = ('cpu', 'cuda', 'device')

    def __init__(self, device: torch.device) -> None:
        self.device = device
        self.cpu = torch.get_rng_state()
        self.cuda = torch.cuda.get_rng_state(device) if device.type == 'cuda' else None

    @contextmanager
    def restored(self) -> Generator[None, None, None]:
        cuda_devices = [self.device] if self.cuda is not None else []
        with torch.random.fork_rng(cuda_devices):
            torch.set_rng_state(self.cpu)
            if self.cuda is not None:
                torch.cuda.set_rng_state(self.cuda, self.device)
            yield


class _Shared:
    """State shared by the Checkpoint/Recompute pair of one (micro-batch, stage)."""

    __slots__ = ('function', 'atomic', 'rng', 'recomputed')

    def __init__(self, function: Function, atomic: bool) -> None:
        self.function = function
        self.atomic = atomic
        self.rng: Deque[_RNGSnapshot] = deque(maxlen=1)
        self.recomputed: Deque[Tuple[TensorOrTensors, Tensors]] = deque(maxlen=1)

    def run(self, tensors: Tensors) -> TensorOrTensors:
        return self.function(tensors[0] if self.atomic else tensors)


class Checkpointing:
    """Builds the ``Checkpoint`` / ``Recompute`` pair for one micro-batch on one stage."""

    def __init__(self, function: Function, batch: Batch) -> None:
        self.function = function
        self.batch = batch
        self._shared = _Shared(function, batch.atomic)

    # Attributes exposed for tests / introspection.
    @property
    def recomputed(self) -> Deque[Tuple[TensorOrTensors, Tensors]]:
        return self._shared.recomputed

    @property
    def rng_states(self) -> Deque[_RNGSnapshot]:
        return self._shared.rng

    def checkpoint(self) -> Batch:
        """Runs the forward without building a graph; returns the output batch."""
        inputs = tuple(self.batch)
        # A phony that requires grad guarantees the output requires grad, so
        # Checkpoint.backward is reached even when no input requires grad.
        phony = get_phony(inputs[0].device, requires_grad=True)
        output = Checkpoint.apply(phony, self._shared, *inputs)
        return Batch(output)

    def recompute(self, batch: Batch) -> None:
        """Attaches the recomputation to ``batch`` (the stage output), in place."""
        inputs = tuple(self.batch)
        batch[0], phony = fork(batch[0])
        phony = Recompute.apply(phony, self._shared, *inputs)
        batch[0] = join(batch[0], phony)


class Checkpoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, phony: Tensor, shared: _Shared, *inputs: Tensor):  # type: ignore[override]
        ctx.shared = shared
        shared.rng.append(_RNGSnapshot(inputs[0].device))
        ctx.save_for_backward(*inputs)
        with torch.no_grad(), enable_checkpointing():
            return shared.run(inputs)

    @staticmethod
    def backward(ctx, *grad_output: Tensor):  # type: ignore[override]
        output, leaves = ctx.shared.recomputed.pop()
        outputs = output if isinstance(output, tuple) else (output,)
        if any(y.requires_grad for y in outputs):
            pairs = [(y, g) for y, g in zip(outputs, grad_output) if y.requires_grad]
            torch.autograd.backward([y for y, _ in pairs], [g for _, g in pairs])
        grads: List[Optional[Tensor]] = [None, None]
        grads.extend(leaf.grad for leaf in leaves)
        return tuple(grads)


class Recompute(torch.autograd.Function):
    @staticmethod
    def forward(ctx, phony: Tensor, shared: _Shared, *inputs: Tensor):  # type: ignore[override]
        ctx.shared = shared
        ctx.save_for_backward(*inputs)
        return phony

    @staticmethod
    def backward(ctx, *grad_output: Tensor):  # type: ignore[override]
        shared: _Shared = ctx.shared
        inputs = ctx.saved_tensors
        leaves = tuple(x.detach().requires_grad_(x.requires_grad) for x in inputs)
        with shared.rng.pop().restored():
            with torch.enable_grad(), enable_recomputing():
                output = shared.run(leaves)
        shared.recomputed.append((output, leaves))
        return (None, None) + tuple(None for _ in inputs)


def checkpoint(function: Function, input: TensorOrTensors) -> TensorOrTensors:
    """Convenience wrapper with the call shape of ``torch.utils.checkpoint.checkpoint``."""
    batch = Batch(input)
    chk = Checkpointing(function, batch)
    out = chk.checkpoint()
    chk.recompute(out)
    return out.tensor_or_tensors
