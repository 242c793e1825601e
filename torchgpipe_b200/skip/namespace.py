"""Namespaces that let several skip tensors share one name in a model.

Contract (reference ``torchgpipe/skip/namespace.py:10-43``): a namespace is unique, hashable,
totally ordered, equal to its copies (``copy.deepcopy`` / pickling a model keeps the skip routes
intact), and ``None`` counts as "the default namespace" for ``isinstance`` checks.

Built differently here: the identity is a ``(creation index, random salt)`` pair instead of a
UUID. Within one process namespaces therefore sort in creation order - skip layouts, which sort
``(namespace, name)`` pairs, come out the same on every run - and the salt keeps namespaces made
in different processes (every rank of the multi-process mode builds its own model) from ever
comparing equal by accident. ``None`` is recognised by the metaclass, not by ABC registration.
"""
import itertools
import secrets
from typing import Any, Tuple

__all__ = ['Namespace']

_creation_index = itertools.count()


class _NamespaceType(type):
    """``isinstance(None, Namespace)`` is True: ``None`` is the default namespace."""

    def __instancecheck__(cls, obj: Any) -> bool:
        return obj is None or type.__instancecheck__(cls, obj)


class Namespace(metaclass=_NamespaceType):
    """An opaque scope for skip-tensor names; see ``Skippable.isolate``."""

    __slots__ = ('key',)

    def __init__(self) -> None:
        self.key: Tuple[int, int] = (next(_creation_index), secrets.randbits(64))

    @property
    def id(self) -> str:
        """Printable identity (the reference exposes a UUID under this name)."""
        return '%d-%016x' % self.key

    def __repr__(self) -> str:
        return f"<Namespace '{self.id}'>"

    def __hash__(self) -> int:
        return hash(self.key)

    @staticmethod
    def _key_of(other: Any) -> Any:
        """The key of a real namespace; ``None`` for the default namespace and foreign objects,
        which compare unequal to every namespace and sort before all of them."""
        if other is not None and isinstance(other, Namespace):
            return other.key
        return None

    def __eq__(self, other: Any) -> bool:
        key = self._key_of(other)
        return key is not None and key == self.key

    def __ne__(self, other: Any) -> bool:
        return not self.__eq__(other)

    def __lt__(self, other: Any) -> bool:
        key = self._key_of(other)
        return key is not None and self.key < key

    def __le__(self, other: Any) -> bool:
        key = self._key_of(other)
        return key is not None and self.key <= key

    def __gt__(self, other: Any) -> bool:
        key = self._key_of(other)
        return key is None or self.key > key

    def __ge__(self, other: Any) -> bool:
        key = self._key_of(other)
        return key is None or self.key >= key
