"""``GPipe``: wrap an ``nn.Sequential`` to train it as a micro-batched pipeline.

Public surface, argument coercions, error types/messages and ``state_dict``
keys match reference ``torchgpipe/gpipe.py:53-380`` (checked by
``tests/test_gpipe_api.py`` here). What is different underneath is the
runtime: worker threads and side streams are created once per module, not per
call, and an optional :class:`~torchgpipe_b200.runtime.timeline.Timeline`
reports pipeline bubble and exposed communication per step.
"""
from collections import OrderedDict
from typing import Any, Dict, Iterable, Iterator, List, Optional, Tuple, Union

import torch
from torch import Tensor, nn

from torchgpipe_b200 import microbatch
from torchgpipe_b200.batchnorm import DeferredBatchNorm
from torchgpipe_b200.runtime.executor import ExecutorPool
from torchgpipe_b200.runtime.pipeline import Pipeline
from torchgpipe_b200.runtime.streams import StreamPool
from torchgpipe_b200.runtime.timeline import Timeline
from torchgpipe_b200.skip.layout import inspect_skip_layout
from torchgpipe_b200.skip.skippable import verify_skippables

__all__ = ['GPipe', 'BalanceError', 'verify_module', 'split_module', 'recommend_auto_balance']

Device = Union[torch.device, int, str]
Devices = Iterable[Device]
Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]

_CHECKPOINT_MODES = ('always', 'except_last', 'never')


def recommend_auto_balance(message: str) -> str:
    """Appends a pointer to the automatic balancers to an error message."""
    return f'''{message}

If your model is still under development, its optimal balance would change
frequently. In this case, we highly recommend 'torchgpipe.balance' for naive
automatic balancing:

  from torchgpipe_b200 import GPipe
  from torchgpipe_b200.balance import balance_by_time

  partitions = torch.cuda.device_count()
  sample = torch.empty(...)
  balance = balance_by_time(partitions, model, sample)

  model = GPipe(model, balance, ...)
'''


class BalanceError(ValueError):
    """``balance`` does not describe the module (wrong sum or non-positive entry)."""


def verify_module(module: nn.Sequential) -> None:
    """Rejects anything that cannot be cut into disjoint consecutive stages."""
    if not isinstance(module, nn.Sequential):
        raise TypeError('module must be nn.Sequential to be partitioned')

    if len(list(module.named_children())) != len(module):
        raise ValueError('module with duplicate children is not supported')

    total = len(list(module.parameters()))
    per_child = sum(len(list(child.parameters())) for child in module.children())
    if total != per_child:
        raise ValueError('module with duplicate parameters in distinct children is not supported')


def check_balance(module: nn.Sequential, balance: List[int]) -> None:
    if len(module) != sum(balance):
        raise BalanceError('module and sum of balance have different length '
                           f'(module: {len(module)}, sum of balance: {sum(balance)})')
    if any(x <= 0 for x in balance):
        raise BalanceError(f'all balance numbers must be positive integer (balance: {balance})')


def split_module(module: nn.Sequential,
                 balance: Iterable[int],
                 devices: List[torch.device],
                 ) -> Tuple[nn.ModuleList, List[int], List[torch.device]]:
    """Cuts ``module`` into ``len(balance)`` stages and places stage ``j`` on ``devices[j]``.

    Children keep their original names inside each stage, so the wrapper's
    ``state_dict`` keys are ``partitions.<j>.<child name>.<...>``. Surplus
    devices are dropped from the returned list.
    """
    balance = list(balance)
    check_balance(module, balance)
    if len(balance) > len(devices):
        raise IndexError('too few devices to hold given partitions '
                         f'(devices: {len(devices)}, partitions: {len(balance)})')

    named = iter(module.named_children())
    stages = []
    for j, count in enumerate(balance):
        stage = nn.Sequential(OrderedDict(next(named) for _ in range(count)))
        verify_fusion_partners(stage, j)
        stage.to(devices[j])
        stages.append(stage)

    return nn.ModuleList(stages), balance, devices[:len(balance)]


def verify_fusion_partners(stage: nn.Module, index: int) -> None:
    """A fused layer made by ``ops.convert_to_sm100`` may use the parameters of a neighbouring
    placeholder layer (``fusion_partners()``); both must land in the same stage."""
    members = None
    for name, layer in stage.named_modules():
        partners = getattr(layer, 'fusion_partners', None)
        if partners is None:
            continue
        if members is None:
            members = {id(m) for m in stage.modules()}
        for partner in partners():
            if id(partner) not in members:
                raise ValueError(
                    f'balance cuts a fused run: layer {name!r} of stage {index} uses the '
                    f'parameters of a {type(partner).__name__} layer placed in another stage; '
                    'move the cut by one layer or convert the model after choosing the balance '
                    'with that run left unfused')


_MOVING_DENIED = ('denied to move parameters and buffers, '
                  'because GPipe should manage device placement')


class GPipe(nn.Module):
    """Pipeline-parallel wrapper for an :class:`torch.nn.Sequential`.

    ::

        model = nn.Sequential(a, b, c, d)
        model = GPipe(model, balance=[1, 1, 1, 1], chunks=8)
        output = model(input)          # input on devices[0], output on devices[-1]

    Args:
        module: sequential module to split into stages.
        balance: number of layers in each stage (required; see
            :mod:`torchgpipe_b200.balance` for automatic balancing).

    Keyword Args:
        devices: devices for the stages (default: every CUDA device, in order).
            CPU devices and repeated devices are allowed.
        chunks: number of micro-batches (default ``1``).
        checkpoint: ``'always'``, ``'except_last'`` (default) or ``'never'``.
        deferred_batch_norm: track BatchNorm running statistics per mini-batch
            instead of per micro-batch.
        cuda_graphs: replay every stage pass from CUDA graphs
            (:mod:`torchgpipe_b200.runtime.graph_pipeline`): one graph launch per
            micro-batch and stage instead of hundreds of kernel launches, which is
            what keeps B200-class devices busy at the published micro-batch
            counts. Same call pattern (``out = model(x); loss(out).backward()``);
            parameter gradients arrive in ``param.grad``. Needs CUDA devices,
            static micro-batch shapes and host-sync-free layers; every
            micro-batch is recomputed from its stage input (``checkpoint`` then
            only names the memory behaviour you get anyway), the pipeline input
            receives no gradient, and ``deferred_batch_norm`` is refused.

    Raises:
        TypeError: ``module`` is not ``nn.Sequential`` / skips do not match.
        ValueError: invalid arguments or wrong balance.
        IndexError: fewer devices than stages.
    """

    #: Number of layers per stage.
    balance: List[int] = []
    #: Device of each stage; ``devices[-1]`` is where the output lives.
    devices: List[torch.device] = []
    #: Number of micro-batches.
    chunks: int = 1
    #: One of ``'always'``, ``'except_last'``, ``'never'``.
    checkpoint: str = 'except_last'

    def __init__(self,
                 module: nn.Sequential,
                 balance: Optional[Iterable[int]] = None,
                 *,
                 devices: Optional[Devices] = None,
                 chunks: int = chunks,
                 checkpoint: str = checkpoint,
                 deferred_batch_norm: bool = False,
                 cuda_graphs: bool = False,
                 ) -> None:
        super().__init__()

        chunks = int(chunks)
        checkpoint = str(checkpoint)

        if balance is None:
            raise ValueError(recommend_auto_balance('balance is required'))
        if chunks <= 0:
            raise ValueError('number of chunks must be positive integer')
        if checkpoint not in _CHECKPOINT_MODES:
            raise ValueError("checkpoint is not one of 'always', 'except_last', or 'never'")

        verify_module(module)
        verify_skippables(module)   # static: can be checked before any forward

        self.chunks = chunks
        self.checkpoint = checkpoint

        if deferred_batch_norm and cuda_graphs:
            raise ValueError(
                'cuda_graphs=True cannot be combined with deferred_batch_norm=True: the '
                'micro-batch counter and the commit of the accumulated statistics are host-side '
                'state that a graph replay does not execute')
        if deferred_batch_norm:
            module = DeferredBatchNorm.convert_deferred_batch_norm(module, chunks)

        if devices is None:
            devices = range(torch.cuda.device_count())
        device_list = [torch.device(d) for d in devices]

        try:
            self.partitions, self.balance, self.devices = \
                split_module(module, balance, device_list)
        except BalanceError as exc:
            raise ValueError(recommend_auto_balance(str(exc)))

        self._skip_layout = inspect_skip_layout(self.partitions)
        # Runtime resources are created on first use and then kept.
        self._copy_streams: Optional[StreamPool] = None
        self._executors: Optional[ExecutorPool] = None
        #: Set to a :class:`Timeline` to record per-cell CUDA events.
        self.timeline: Optional[Timeline] = None
        #: Whether stage passes are replayed from CUDA graphs (see the class docstring).
        self.cuda_graphs = bool(cuda_graphs)
        if self.cuda_graphs and any(d.type != 'cuda' for d in self.devices):
            raise ValueError('cuda_graphs=True needs every partition on a CUDA device')
        self._graph_engine: Any = None

    # -- sequence protocol over the ORIGINAL layers ---------------------------
    def __len__(self) -> int:
        return sum(len(p) for p in self.partitions)

    def __getitem__(self, index: int) -> nn.Module:
        total = len(self)
        if index < 0:
            index += total
        if not 0 <= index < total:
            raise IndexError
        for partition in self.partitions:
            if index < len(partition):
                return partition[index]
            index -= len(partition)
        raise IndexError

    def __iter__(self) -> Iterator[nn.Module]:
        for partition in self.partitions:
            yield from partition

    # -- placement is owned by GPipe ------------------------------------------
    def cuda(self, device: Optional[Device] = None) -> 'GPipe':
        raise TypeError(_MOVING_DENIED)

    def cpu(self) -> 'GPipe':
        raise TypeError(_MOVING_DENIED)

    def to(self, *args: Any, **kwargs: Any) -> 'GPipe':
        # to(dtype[, non_blocking]) is fine; anything that names a device is not.
        if 'device' in kwargs or 'tensor' in kwargs:
            raise TypeError(_MOVING_DENIED)
        if args and (isinstance(args[0], (torch.device, int, str)) or torch.is_tensor(args[0])):
            raise TypeError(_MOVING_DENIED)
        return super().to(*args, **kwargs)

    # -- state_dict helpers (balance-independent checkpoints) -----------------
    def sequential_state_dict(self) -> 'OrderedDict[str, Tensor]':
        """``state_dict`` with the ``partitions.<j>.`` prefix stripped.

        The result loads into the plain ``nn.Sequential`` the pipeline was
        built from, and into a :class:`GPipe` with any other ``balance`` via
        :meth:`load_sequential_state_dict`.
        """
        out: 'OrderedDict[str, Tensor]' = OrderedDict()
        for key, value in self.state_dict().items():
            out[_strip_partition_prefix(key)] = value
        return out

    def load_sequential_state_dict(self, state_dict: Dict[str, Tensor],
                                   strict: bool = True) -> Any:
        """Loads a plain-``nn.Sequential`` (or any-balance GPipe) ``state_dict``."""
        where: Dict[str, int] = {}
        for j, partition in enumerate(self.partitions):
            for name, _ in partition.named_children():
                where[name] = j
        remapped: 'OrderedDict[str, Tensor]' = OrderedDict()
        for key, value in state_dict.items():
            key = _strip_partition_prefix(key)
            child = key.split('.', 1)[0]
            if child in where:
                key = f'partitions.{where[child]}.{key}'
            remapped[key] = value
        return self.load_state_dict(remapped, strict=strict)

    # -- runtime ----------------------------------------------------------------
    def _ensure_copy_streams(self) -> StreamPool:
        if self._copy_streams is None:
            self._copy_streams = StreamPool(self.devices, self.chunks)
        return self._copy_streams

    def _ensure_executors(self) -> ExecutorPool:
        if self._executors is None:
            self._executors = ExecutorPool(self.devices)
        return self._executors

    def native_launches_replayed(self) -> int:
        """Native-kernel launches issued through CUDA-graph replays so far (graph mode)."""
        return 0 if self._graph_engine is None else self._graph_engine.native_launches_replayed()

    def _checkpoint_stop(self) -> int:
        if not self.training:
            return 0
        return {'always': self.chunks, 'except_last': self.chunks - 1, 'never': 0}[self.checkpoint]

    def forward(self, input: TensorOrTensors) -> TensorOrTensors:  # type: ignore[override]
        """Runs the mini-batch through the pipeline.

        ``input`` is a tensor or a tuple of tensors on ``devices[0]``; the same
        restriction applies at every stage boundary. Returns the concatenated
        micro-batch outputs on ``devices[-1]``.
        """
        microbatch.check(input)

        if not self.devices:
            return input        # an empty nn.Sequential is the identity

        if self.cuda_graphs:
            if self._graph_engine is None:
                from torchgpipe_b200.runtime.graph_pipeline import GraphedPipeline
                self._graph_engine = GraphedPipeline(self.partitions, self.devices, self.chunks,
                                                     self.checkpoint)
            return self._graph_engine(input)

        batches = microbatch.scatter(input, self.chunks)
        pipeline = Pipeline(batches,
                            self.partitions,
                            self.devices,
                            self._ensure_copy_streams(),
                            self._skip_layout,
                            self._checkpoint_stop(),
                            executors=self._ensure_executors(),
                            timeline=self.timeline)
        pipeline.run()
        return microbatch.gather(batches)


def _strip_partition_prefix(key: str) -> str:
    parts = key.split('.', 2)
    if len(parts) == 3 and parts[0] == 'partitions' and parts[1].isdigit():
        return parts[2]
    return key


def _gpipe_getstate(self: GPipe) -> Dict[str, Any]:
    # Threads, streams and event recorders are process-local: rebuild lazily.
    state = dict(self.__dict__)
    state['_copy_streams'] = None
    state['_executors'] = None
    state['_graph_engine'] = None
    state['timeline'] = None
    return state


GPipe.__getstate__ = _gpipe_getstate  # type: ignore[assignment]
