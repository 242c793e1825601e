"""Micro-batch container and mini-batch <-> micro-batch conversion.

``Batch`` hides whether a stage boundary carries one tensor or a tuple of
tensors. ``scatter`` follows ``torch.chunk`` semantics, so a mini-batch whose
size is not divisible by ``chunks`` (or is smaller than ``chunks``) is legal
and simply yields fewer / uneven micro-batches. Contract: reference
``torchgpipe/microbatch.py:17-177``.
"""
from typing import Callable, Iterator, List, Tuple, Union

import torch
from torch import Tensor

__all__ = ['Batch', 'check', 'scatter', 'gather']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


class Batch:
    """One micro-batch: an *atomic* tensor or a tuple of tensors."""

    __slots__ = ('value', 'atomic')

    def __init__(self, value: TensorOrTensors) -> None:
        self.value = value
        self.atomic = torch.is_tensor(value)

    # -- typed accessors -------------------------------------------------
    @property
    def tensor(self) -> Tensor:
        if not self.atomic:
            raise AttributeError('not atomic batch')
        return self.value  # type: ignore[return-value]

    @property
    def tensors(self) -> Tensors:
        if self.atomic:
            raise AttributeError('batch is atomic')
        return self.value  # type: ignore[return-value]

    @property
    def tensor_or_tensors(self) -> TensorOrTensors:
        return self.value

    def call(self, function: Callable[[TensorOrTensors], TensorOrTensors]) -> 'Batch':
        """Applies ``function`` to the payload and wraps the result."""
        return Batch(function(self.value))

    # -- sequence protocol -----------------------------------------------
    def __repr__(self) -> str:
        return f'Batch[atomic={self.atomic!r}]({self.value!r})'

    def __iter__(self) -> Iterator[Tensor]:
        if self.atomic:
            yield self.value  # type: ignore[misc]
        else:
            yield from self.value  # type: ignore[misc]

    def __len__(self) -> int:
        return 1 if self.atomic else len(self.value)  # type: ignore[arg-type]

    def __getitem__(self, index: int) -> Tensor:
        if self.atomic:
            if index != 0:
                raise IndexError('atomic batch allows index 0 only')
            return self.value  # type: ignore[return-value]
        return self.value[index]  # type: ignore[index]

    def __setitem__(self, index: Union[int, slice], value: TensorOrTensors) -> None:
        if isinstance(index, slice):
            if not (index.start is None and index.stop is None and index.step is None):
                raise NotImplementedError('only slice [:] supported')
            if self.atomic:
                if len(value) != 1:  # type: ignore[arg-type]
                    raise IndexError('atomic batch cannot be replaced with multiple tensors')
                self.value = value[0]  # type: ignore[index]
            else:
                self.value = tuple(value)  # type: ignore[arg-type]
            return

        if self.atomic:
            if index != 0:
                raise IndexError('atomic batch allows index 0 only')
            self.value = value
            return
        items = list(self.value)  # type: ignore[arg-type]
        items[index] = value  # type: ignore[call-overload]
        self.value = tuple(items)


def check(input: TensorOrTensors) -> None:
    """Raises ``TypeError`` unless ``input`` is a tensor or a (nested) tuple of tensors."""
    if isinstance(input, tuple):
        for x in input:
            check(x)
        return
    if not isinstance(input, Tensor):
        raise TypeError(f'expected Tensor, but got {input.__class__.__name__}')


def scatter(input: TensorOrTensors, chunks: int) -> List[Batch]:
    """Splits a mini-batch along dim 0 into at most ``chunks`` micro-batches (views)."""
    if isinstance(input, Tensor):
        return [Batch(piece) for piece in input.chunk(chunks)]
    per_tensor = [t.chunk(chunks) for t in input]
    return [Batch(tuple(pieces)) for pieces in zip(*per_tensor)]


def gather(outputs: List[Batch]) -> TensorOrTensors:
    """Concatenates micro-batch outputs back into one mini-batch."""
    if outputs[0].atomic:
        return torch.cat([b.tensor for b in outputs])
    columns = zip(*(b.tensors for b in outputs))
    return tuple(torch.cat(col) for col in columns)
