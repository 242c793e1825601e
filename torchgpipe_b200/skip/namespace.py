"""Namespaces that let several skip tensors share one name in a model.

Contract: reference ``torchgpipe/skip/namespace.py:10-43`` - unique, hashable,
totally ordered, equal to its copies, and ``None`` counts as "the default
namespace" for ``isinstance`` checks.
"""
import abc
from functools import total_ordering
from typing import Any
import uuid

__all__ = ['Namespace']


@total_ordering
class Namespace(metaclass=abc.ABCMeta):
    """An opaque scope for skip-tensor names; see ``Skippable.isolate``."""

    __slots__ = ('id',)

    def __init__(self) -> None:
        self.id = uuid.uuid4()

    def __repr__(self) -> str:
        return f"<Namespace '{self.id}'>"

    def __hash__(self) -> int:
        return hash(self.id)

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, Namespace) and other is not None and self.id == other.id

    def __lt__(self, other: Any) -> bool:
        # ``None`` (the default namespace) never sorts after a real namespace.
        if isinstance(other, Namespace) and other is not None:
            return self.id < other.id
        return False


# ``isinstance(None, Namespace)`` is True: None is the default namespace.
Namespace.register(type(None))
