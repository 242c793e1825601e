"""User-facing API for skip connections: ``@skippable``, ``stash``, ``pop``.

A skippable module's ``forward`` is a generator. It hands tensors to the
framework with ``yield stash(name, tensor)`` and asks for them back with
``tensor = yield pop(name)``; the framework decides whether the tensor stays in
a dictionary (no pipeline / same stage) or travels through a portal straight
to the stage that pops it. Contract: reference
``torchgpipe/skip/skippable.py:31-416``; message texts pinned by reference
``tests/skip/test_verify_skippables.py`` and ``tests/skip/test_stash_pop.py``.
"""
from typing import (Any, Callable, ClassVar, Dict, FrozenSet, Generator, Iterable, Iterator, List,
                    Optional, Set, Tuple, Type, TypeVar, Union)

from torch import Tensor, nn

from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.skip.namespace import Namespace
from torchgpipe_b200.skip.tracker import current_skip_tracker

__all__ = ['skippable', 'stash', 'pop', 'verify_skippables', 'Skippable']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]
T = TypeVar('T', bound='Skippable')
SkipKey = Tuple[Namespace, str]


class stash:  # noqa: N801 (command objects are lower-case by API)
    """``yield stash(name, tensor)`` - hand ``tensor`` (or ``None``) to the skip named ``name``."""

    __slots__ = ('name', 'tensor')

    def __init__(self, name: str, tensor: Optional[Tensor]) -> None:
        self.name = name
        self.tensor = tensor


class pop:  # noqa: N801
    """``tensor = yield pop(name)`` - take back the tensor stashed under ``name``."""

    __slots__ = ('name',)

    def __init__(self, name: str) -> None:
        self.name = name


class Skippable(nn.Module):
    """Base of every class produced by :func:`skippable`; do not subclass by hand."""

    module_cls: ClassVar[Type[nn.Module]]
    stashable_names: ClassVar[FrozenSet[str]]
    poppable_names: ClassVar[FrozenSet[str]]

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__()
        self.module = self.module_cls(*args, **kwargs)
        self.namespaces: Dict[str, Namespace] = {}

    def __repr__(self) -> str:
        return f'@skippable({self.module})'

    # -- naming ------------------------------------------------------------
    def namespaced(self, name: str) -> SkipKey:
        return (self.namespaces.get(name), name)  # type: ignore[return-value]

    def stashable(self) -> Iterator[SkipKey]:
        for name in self.stashable_names:
            yield self.namespaced(name)

    def poppable(self) -> Iterator[SkipKey]:
        for name in self.poppable_names:
            yield self.namespaced(name)

    def isolate(self: T, ns: Namespace, *, only: Optional[Iterable[str]] = None) -> T:
        """Puts all (or ``only`` the given) skip names of this layer into namespace ``ns``.

        Two stash/pop pairs may reuse a name inside one ``nn.Sequential`` only
        if they are isolated in different namespaces. Returns ``self`` so calls
        can be chained.
        """
        names = (self.stashable_names | self.poppable_names) if only is None else set(only)
        for name in names:
            self.namespaces[name] = ns
        return self

    # -- execution ---------------------------------------------------------
    def dispatch(self,
                 input: TensorOrTensors,
                 handle_stash: Callable[[str, Optional[Tensor]], None],
                 handle_pop: Callable[[str], Optional[Tensor]],
                 ) -> TensorOrTensors:
        """Drives the wrapped ``forward``; a non-generator ``forward`` passes through."""
        result = self.module(input)
        if not isinstance(result, Generator):
            return result

        gen = result
        try:
            command = next(gen)
            while True:
                if isinstance(command, stash):
                    handle_stash(command.name, command.tensor)
                    command = next(gen)
                elif isinstance(command, pop):
                    command = gen.send(handle_pop(command.name))
                else:
                    raise TypeError('%r is not a command from @skippable' % command)
        except StopIteration as stop:
            return stop.args[0]

    def forward(self, input: TensorOrTensors) -> TensorOrTensors:  # type: ignore[override]
        tracker = current_skip_tracker()
        batch = Batch(input)

        # Load everything this layer may pop *before* running it, so the
        # portal/fork nodes precede the layer's own nodes in the graph.
        available: Dict[str, Optional[Tensor]] = {}
        for ns, name in self.poppable():
            try:
                available[name] = tracker.load(batch, ns, name)
            except KeyError:
                raise RuntimeError(f"'{name}' has not been stashed")
        input = batch.tensor_or_tensors

        stashed: Dict[str, Optional[Tensor]] = {}

        def handle_stash(name: str, tensor: Optional[Tensor]) -> None:
            if name not in self.stashable_names:
                raise RuntimeError(f"'{name}' has not been declared as stashable")
            stashed[name] = tensor

        def handle_pop(name: str) -> Optional[Tensor]:
            if name not in self.poppable_names:
                raise RuntimeError(f"'{name}' has not been declared as poppable")
            return available.pop(name)

        output = self.dispatch(input, handle_stash, handle_pop)

        missing = self.stashable_names - stashed.keys()
        if missing:
            listed = ', '.join("'%s'" % n for n in missing)
            raise RuntimeError(f'{listed} must be stashed but have not')
        if available:
            listed = ', '.join("'%s'" % n for n in available)
            raise RuntimeError(f'{listed} must be popped but have not')

        batch = Batch(output)
        for ns, name in self.stashable():
            tracker.save(batch, ns, name, stashed[name])
        return batch.tensor_or_tensors


def skippable(stash: Iterable[str] = (),  # noqa: A002 (API names)
              pop: Iterable[str] = (),
              ) -> Callable[[Type[nn.Module]], Type[Skippable]]:
    """Class decorator declaring which skip names a module stashes and pops.

    ::

        @skippable(stash=['1to3'])
        class Layer1(nn.Module):
            def forward(self, input):
                yield stash('1to3', input)
                return f1(input)

        @skippable(pop=['1to3'])
        class Layer3(nn.Module):
            def forward(self, input):
                skip = yield pop('1to3')
                return f3(input) + skip

    The decorated class keeps its name, is a :class:`Skippable`, and holds the
    original module as ``.module``. It works with or without
    :class:`~torchgpipe_b200.GPipe`.
    """
    stashable_names = frozenset(stash)
    poppable_names = frozenset(pop)

    def wrap(module_cls: Type[nn.Module]) -> Type[Skippable]:
        return type(module_cls.__name__, (Skippable,), {
            'module_cls': module_cls,
            'stashable_names': stashable_names,
            'poppable_names': poppable_names,
            '__doc__': module_cls.__doc__,
        })

    return wrap


def verify_skippables(module: nn.Sequential) -> None:
    """Checks statically that every skip has exactly one stash followed by one pop.

    Raises one ``TypeError`` listing every violation.
    """
    stashed: Set[SkipKey] = set()
    popped: Set[SkipKey] = set()
    problems: List[str] = []

    for layer_name, layer in module.named_children():
        if not isinstance(layer, Skippable):
            continue

        for name in layer.stashable_names & layer.poppable_names:
            problems.append(f"'{layer_name}' declared '{name}' both as stashable and as poppable")

        for ns, name in layer.stashable():
            if name in layer.poppable_names:
                continue
            if (ns, name) in stashed:
                problems.append(f"'{layer_name}' redeclared '{name}' as stashable "
                                'but not isolated by namespace')
                continue
            stashed.add((ns, name))

        for ns, name in layer.poppable():
            if name in layer.stashable_names:
                continue
            if (ns, name) in popped:
                problems.append(f"'{layer_name}' redeclared '{name}' as poppable "
                                'but not isolated by namespace')
                continue
            if (ns, name) not in stashed:
                problems.append(f"'{layer_name}' declared '{name}' as poppable "
                                'but it was not stashed')
                continue
            popped.add((ns, name))

    for _, name in stashed - popped:
        problems.append(f"no module declared '{name}' as poppable but stashed")

    if problems:
        raise TypeError('one or more pairs of stash and pop do not match:\n\n%s'
                        % '\n'.join('* %s' % p for p in problems))
