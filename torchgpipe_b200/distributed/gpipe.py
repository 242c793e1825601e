"""``DistributedGPipe``: one process per pipeline stage.

Each rank builds the full ``nn.Sequential``, keeps only its own stage
(:func:`get_module_partition`) and exchanges micro-batches with its
neighbours through a :class:`~torchgpipe_b200.distributed.transport.Transport`.

Public surface follows reference ``torchgpipe/distributed/gpipe.py:26-275``:
``DistributedGPipe(module, rank, workers, balance, microbatch_chunks, *, device,
deferred_batch_norm)`` with ``forward(batch_or_None) -> list of per-micro-batch
outputs``, ``backward(losses_or_None)``, ``model()``; static hooks
``_get/_put`` remain the patch points of the queue backend. Added on top
(absent in the reference): ``checkpoint=`` modes, cross-stage
``@skippable`` routes (sent directly from the stashing rank to the popping
rank), a working ``deferred_batch_norm``, and device-resident NCCL transfers.

Per micro-batch the stage is one function ``(boundary inputs + popped skips)
-> (outputs + stashed skips)``; with checkpointing that function is wrapped by
the same ``Checkpoint``/``Recompute`` pair as in single-process mode, so
``backward`` is a single ``torch.autograd.backward`` whose roots are the
stage outputs and the stashed skips.
"""
import contextlib
from collections import OrderedDict
from typing import Any, Dict, Iterable, Iterator, List, Optional, Tuple, Union

import torch
from torch import Tensor, nn

from torchgpipe_b200 import microbatch
from torchgpipe_b200.batchnorm import DeferredBatchNorm
from torchgpipe_b200.checkpoint import Checkpointing
from torchgpipe_b200.distributed import context
from torchgpipe_b200.distributed.transport import ChannelTransport, Transport
from torchgpipe_b200.distributed.utils import to
from torchgpipe_b200.gpipe import (BalanceError, check_balance, recommend_auto_balance,
                                   verify_fusion_partners, verify_module)
from torchgpipe_b200.microbatch import Batch
from torchgpipe_b200.skip.layout import inspect_skip_layout
from torchgpipe_b200.skip.namespace import Namespace
from torchgpipe_b200.skip.skippable import Skippable, verify_skippables
from torchgpipe_b200.skip.tracker import SkipTracker, use_skip_tracker
from torchgpipe_b200.utils import nvtx

__all__ = ['DistributedGPipe', 'DistributedGPipeDataLoader', 'get_module_partition']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]
SkipKey = Tuple[Namespace, str]


def _stage_slices(module: nn.Sequential, balance: List[int]) -> List['OrderedDict[str, nn.Module]']:
    named = iter(module.named_children())
    return [OrderedDict(next(named) for _ in range(count)) for count in balance]


def get_module_partition(module: nn.Sequential,
                         rank: int,
                         balance: Iterable[int],
                         device: Optional[torch.device],
                         ) -> nn.Sequential:
    """Returns stage ``rank`` of ``module`` under ``balance`` (children keep their names),
    moved to ``device`` unless that is ``None``.

    A model built on the ``meta`` device (``with torch.device('meta'): model = ...``) is
    materialised for this stage only: its parameters are allocated on ``device`` uninitialised
    (``to_empty``) and the caller initialises or loads them. That is how a 7.8 B-parameter model
    is set up with one process per GPU without every process building all of it in host memory.
    """
    balance = list(balance)
    check_balance(module, balance)
    if not 0 <= rank < len(balance):
        raise RuntimeError('module and balance mismatch')
    partition = nn.Sequential(_stage_slices(module, balance)[rank])
    verify_fusion_partners(partition, rank)
    if device is not None:
        if any(t.is_meta for t in list(partition.parameters()) + list(partition.buffers())):
            partition.to_empty(device=device)
        else:
            partition.to(device)
    return partition


class _StageSkipTracker(SkipTracker):
    """Tracker of one micro-batch on one rank: cross-stage skips use in/out boxes."""

    def __init__(self, incoming: Dict[SkipKey, Optional[Tensor]], outgoing_keys: List[SkipKey]):
        super().__init__()
        self.incoming = incoming
        self.outgoing_keys = set(outgoing_keys)
        self.outgoing: Dict[SkipKey, Optional[Tensor]] = {}

    def save(self, batch: Batch, ns: Namespace, name: str, tensor: Optional[Tensor]) -> None:
        if (ns, name) in self.outgoing_keys:
            self.outgoing[(ns, name)] = tensor
        else:
            super().save(batch, ns, name, tensor)

    def load(self, batch: Batch, ns: Namespace, name: str) -> Optional[Tensor]:
        if (ns, name) in self.incoming:
            return self.incoming.pop((ns, name))
        return super().load(batch, ns, name)


class _Flight:
    """Everything a rank keeps about one micro-batch between forward and backward."""

    __slots__ = ('leaves', 'n_boundary', 'outputs', 'n_outputs', 'atomic', 'atomic_in', 'rng',
                 'staged')

    def __init__(self) -> None:
        self.leaves: Tensors = ()
        self.n_boundary = 0
        self.outputs: Tensors = ()
        self.n_outputs = 0
        self.atomic = True
        self.atomic_in = True
        self.rng = None
        self.staged = None


class DistributedGPipe:
    """Stage-per-process GPipe. Not an ``nn.Module``; see :meth:`model`.

    Args:
        module: the *whole* sequential model (every rank builds the same one).
        rank: this process's stage index.
        workers: ``{stage rank: unique worker name}``.
        balance: layers per stage.
        microbatch_chunks: number of micro-batches.

    Keyword Args:
        device: device of this stage (``None`` = leave the module where it is).
        deferred_batch_norm: convert BatchNorm layers of this stage.
        checkpoint: ``'never'`` (default, reference behaviour), ``'except_last'``
            or ``'always'``.
        transport: a :class:`Transport`; default is the in-process channel
            backend driven through ``_put``/``_get``.
        cuda_graphs: replay each stage pass from CUDA graphs
            (:class:`~torchgpipe_b200.runtime.graphs.GraphedStage`): one graph
            launch per micro-batch instead of hundreds of kernel launches.
            Implies checkpointing of every micro-batch while training; needs a
            CUDA ``device``, static shapes per micro-batch and host-sync-free layers.
    """

    #: The number of micro-batches.
    chunks: int = 1

    # Patch points of the channel backend (reference gpipe.py:83-96).
    @staticmethod
    def _get(name: str, id: int, backward: bool = False) -> TensorOrTensors:
        func = context.get_backward if backward else context.get_forward
        return func(name, id)

    @staticmethod
    def _put(name: str, id: int, values: TensorOrTensors, backward: bool = False) -> None:
        func = context.put_backward if backward else context.put_forward
        func(name, id, values)

    def __init__(self,
                 module: nn.Sequential,
                 rank: int,
                 workers: Dict[int, str],
                 balance: Optional[Iterable[int]] = None,
                 microbatch_chunks: int = chunks,
                 *,
                 device: Optional[torch.device] = None,
                 deferred_batch_norm: bool = False,
                 checkpoint: str = 'never',
                 transport: Optional[Transport] = None,
                 cuda_graphs: bool = False,
                 partitions: Optional[List[nn.Sequential]] = None,
                 ) -> None:
        microbatch_chunks = int(microbatch_chunks)
        checkpoint = str(checkpoint)

        if balance is None and partitions is None:
            raise ValueError(recommend_auto_balance('balance is required'))
        if microbatch_chunks <= 0:
            raise ValueError('number of chunks must be positive integer')
        if checkpoint not in ('always', 'except_last', 'never'):
            raise ValueError("checkpoint is not one of 'always', 'except_last', or 'never'")

        if partitions is not None:
            # stages that already exist, each on its device (the single-process GPipe hands its
            # own partitions over when it runs them through this engine): nothing to split
            stages = list(partitions)
            balance = [len(s) for s in stages]
        else:
            verify_module(module)
            verify_skippables(module)
            balance = list(balance)
            try:
                check_balance(module, balance)
            except BalanceError as exc:
                raise ValueError(recommend_auto_balance(str(exc)))
            stages = [nn.Sequential(s) for s in _stage_slices(module, balance)]

        # Skip routes are numbered in model order, identically on every rank.
        layout = inspect_skip_layout(stages)
        order: Dict[SkipKey, int] = {}
        for stage in stages:
            for layer in stage:
                if isinstance(layer, Skippable):
                    for key in layer.poppable():
                        order.setdefault(key, len(order))
        routes = [(order[(ns, name)], src, dst, (ns, name))
                  for src, dst, ns, name in layout.cross_routes()]
        routes.sort()
        self._routes_out = [(slot, dst, key) for slot, src, dst, key in routes if src == rank]
        self._routes_in = [(slot, src, key) for slot, src, dst, key in routes if dst == rank]

        if partitions is not None:
            partition = stages[rank]
        else:
            partition = get_module_partition(module, rank, balance, device)
        if deferred_batch_norm:
            partition = DeferredBatchNorm.convert_deferred_batch_norm(partition, microbatch_chunks)
        self.module = partition

        self.rank = rank
        self.world_size = len(workers)
        self.workers = workers
        self.chunks = microbatch_chunks
        self.checkpoint = checkpoint
        self.device = device
        self.name = workers[rank]
        self.balance = balance

        if transport is None:
            transport = ChannelTransport(workers, rank, type(self))
        self.transport = transport
        # every pair of stages that will talk: neighbours, and the two ends of each skip route
        pairs = [(r, r + 1) for r in range(len(workers) - 1)]
        pairs += sorted({(src, dst) for _, src, dst, _ in routes})
        transport.prepare_pairs(pairs)

        self._outputs: List[Optional[TensorOrTensors]] = [None] * microbatch_chunks
        self._flights: List[Optional[_Flight]] = [None] * microbatch_chunks
        self._live = 0      # micro-batches produced by the latest forward()
        self.cuda_graphs = bool(cuda_graphs) and device is not None and \
            torch.device(device).type == 'cuda'
        if self.cuda_graphs:
            self._refuse_host_state_under_graphs(partition, deferred_batch_norm)
        self._graphed: Dict[Tuple[int, bool], Any] = {}
        #: set to True to record CUDA events around every stage pass (see timeline_report)
        self.timeline = False
        self._marks: List[Tuple[str, Any, Any, Any]] = []
        self._bytes_received = 0
        #: Held around everything this stage enqueues on its device for one micro-batch pass
        #: (never while waiting for a message). A no-op unless several stages of one process
        #: share a device: then it is one lock per device, because a CUDA-graph capture must not
        #: overlap with work another thread puts on the same device's default stream.
        self.device_lock: Any = contextlib.nullcontext()

    @staticmethod
    def _refuse_host_state_under_graphs(partition: nn.Module, deferred_batch_norm: bool) -> None:
        """A replayed graph runs kernels only: Python-side counters and branches are frozen at
        capture time. Refuse the layer configurations whose correctness depends on them."""
        from torch.nn.modules.batchnorm import _BatchNorm
        if deferred_batch_norm or any(isinstance(m, DeferredBatchNorm) or
                                      getattr(m, 'deferred_chunks', 0) > 0
                                      for m in partition.modules()):
            raise ValueError(
                'cuda_graphs=True cannot be combined with deferred_batch_norm=True: the '
                'micro-batch counter and the commit of the accumulated statistics are host-side '
                'state that a graph replay does not execute')
        for m in partition.modules():
            tracks = getattr(m, 'track_running_stats', False)
            native = type(m).__name__ == 'AbsorbedBatchNorm2d'    # counter kept on the device
            if isinstance(m, _BatchNorm) and tracks and not native \
                    and getattr(m, 'momentum', 0.1) is None:
                raise ValueError(
                    'cuda_graphs=True needs BatchNorm layers with a numeric momentum: '
                    'momentum=None (cumulative average) reads num_batches_tracked on the host')

    # -- topology ----------------------------------------------------------------
    @property
    def is_first(self) -> bool:
        return self.rank == 0

    @property
    def is_last(self) -> bool:
        return self.rank == self.world_size - 1

    def _previous_worker(self) -> Optional[str]:
        return None if self.is_first else self.workers[self.rank - 1]

    def _next_worker(self) -> Optional[str]:
        return None if self.is_last else self.workers[self.rank + 1]

    # -- per-pass timeline -------------------------------------------------------
    def _mark(self) -> Any:
        if not (self.timeline and self.device is not None
                and torch.device(self.device).type == 'cuda'):
            return None
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(self.device))
        return ev

    def timeline_report(self, reset: bool = True) -> Dict[str, float]:
        """Summary of the passes recorded since the last reset (synchronises the device).

        ``busy_ms``: time the compute stream spent inside stage passes; ``recv_wait_ms``: time
        it sat waiting for activations / gradients from the neighbours (pipeline bubble plus
        any communication that was not overlapped); ``span_ms``: first pass start to last pass
        end; ``bubble_frac`` = 1 - busy/span; ``comm_ms_bound``: bytes received over the
        measured peer bandwidth (``utils.link_bandwidth_gbs``: 775 GB/s) - an upper bound on the exposed communication, which
        is also bounded by ``recv_wait_ms``.
        """
        if not self._marks:
            return {}
        torch.cuda.synchronize(self.device)
        busy = sum(t1.elapsed_time(t2) for _, _, t1, t2 in self._marks)
        wait = sum(t0.elapsed_time(t1) for _, t0, t1, _ in self._marks)
        span = self._marks[0][1].elapsed_time(self._marks[-1][3])
        from torchgpipe_b200.utils import link_bandwidth_gbs
        comm_bound = self._bytes_received / (link_bandwidth_gbs() * 1e9) * 1e3
        report = {'busy_ms': busy, 'recv_wait_ms': wait, 'span_ms': span,
                  'bubble_frac': max(0.0, 1.0 - busy / span) if span > 0 else 0.0,
                  'bytes_received': float(self._bytes_received), 'comm_ms_bound': comm_bound,
                  'exposed_comm_ms_bound': min(wait, comm_bound)}
        if reset:
            self._marks = []
            self._bytes_received = 0
        return report

    def _count_bytes(self, value: Any) -> None:
        if self.timeline:
            for t in ((value,) if torch.is_tensor(value) else value):
                self._bytes_received += t.numel() * t.element_size()

    def native_launches_replayed(self) -> int:
        """Native-kernel launches issued through CUDA-graph replays so far (graph mode)."""
        return sum(st.native_launches_replayed for st in self._graphed.values())

    def model(self) -> nn.Sequential:
        """This rank's stage (give its parameters to the local optimiser)."""
        return self.module

    # -- model persistence -------------------------------------------------------
    def sequential_state_dict(self) -> 'OrderedDict[str, Tensor]':
        """This stage's parameters and buffers under the keys of the un-partitioned
        ``nn.Sequential`` (stages keep the original child names, so the union over all ranks IS
        the plain model's ``state_dict``). The reference keeps per-process partitions and has no
        way to put them back together."""
        return self.module.state_dict()

    def gather_sequential_state_dict(self, dst: int = 0, group: Any = None,
                                     ) -> Optional['OrderedDict[str, Tensor]']:
        """Collects every stage's :meth:`sequential_state_dict` on rank ``dst`` (CPU tensors, in
        stage order); returns ``None`` on the other ranks. Needs an initialised
        ``torch.distributed`` process group; every rank of the pipeline must call it."""
        import torch.distributed as dist
        mine = OrderedDict((k, v.detach().cpu()) for k, v in self.module.state_dict().items())
        me = dist.get_rank(group)
        parts: Optional[List[Any]] = [None] * dist.get_world_size(group) if me == dst else None
        dist.gather_object((self.rank, mine), parts, dst=dst, group=group)
        if parts is None:
            return None
        merged: 'OrderedDict[str, Tensor]' = OrderedDict()
        for _, part in sorted(parts, key=lambda item: item[0]):
            merged.update(part)
        return merged

    def load_sequential_state_dict(self, state: Dict[str, Tensor], strict: bool = True) -> Any:
        """Loads this stage's entries out of a state dict of the whole ``nn.Sequential`` (or of
        any superset: entries of layers owned by other stages are ignored)."""
        names = {name for name, _ in self.module.named_children()}
        mine = OrderedDict((k, v) for k, v in state.items() if k.split('.', 1)[0] in names)
        return self.module.load_state_dict(mine, strict=strict)

    def _checkpoint_stop(self, m: int) -> int:
        if not (self.module.training and torch.is_grad_enabled()):
            return 0
        return {'always': m, 'except_last': m - 1, 'never': 0}[self.checkpoint]

    # -- forward -------------------------------------------------------------------
    def _stage_function(self, n_boundary: int, atomic_in: bool, flight: _Flight):
        routes_in, routes_out = self._routes_in, self._routes_out
        module = self.module

        def function(inputs: Tensors) -> Tensors:
            if torch.is_tensor(inputs):
                inputs = (inputs,)
            boundary = inputs[:n_boundary]
            incoming = {key: t for (_, _, key), t in zip(routes_in, inputs[n_boundary:])}
            tracker = _StageSkipTracker(incoming, [key for _, _, key in routes_out])
            with use_skip_tracker(tracker):
                out = module(boundary[0] if atomic_in else tuple(boundary))
            flight.atomic = torch.is_tensor(out)
            outs = (out,) if flight.atomic else tuple(out)
            microbatch.check(outs)
            flight.n_outputs = len(outs)
            stashes = []
            for _, _, key in routes_out:
                t = tracker.outgoing[key]
                if t is None:
                    raise TypeError('None skip tensors cannot cross processes')
                stashes.append(t)
            return outs + tuple(stashes)

        return function

    def forward(self, batch: Optional[TensorOrTensors]) -> List[TensorOrTensors]:
        """Runs every micro-batch through this stage.

        Rank 0 passes the mini-batch; other ranks pass ``None`` and receive
        their inputs from the previous stage. Returns the list of per
        micro-batch outputs of *this* stage (the last rank computes one loss
        per entry and hands the list of losses to :meth:`backward`).
        """
        grad_mode = torch.is_grad_enabled()
        if batch is not None:
            assert self.is_first, 'only the first stage takes a mini-batch'
            microbatch.check(batch)
            pieces = microbatch.scatter(batch, self.chunks)
            m = len(pieces)
        else:
            assert not self.is_first, 'the first stage needs a mini-batch'
            pieces = None
            m = self.chunks
        stop = self._checkpoint_stop(m)
        tp = self.transport

        self._outputs = [None] * self.chunks
        self._flights = [None] * self.chunks
        self._live = m

        for mb in range(m):
            flight = _Flight()
            t0 = self._mark()
            if pieces is not None:
                received: TensorOrTensors = pieces[mb].value
            else:
                received = tp.recv(self.rank - 1, ('fwd', mb, 0))
                self._count_bytes(received)
            atomic_in = torch.is_tensor(received)
            raw_skips = []
            for slot, src, _ in self._routes_in:
                t = tp.recv(src, ('skip', mb, slot))
                self._count_bytes(t)
                raw_skips.append(t)
            with self.device_lock:
                self._forward_one(mb, m, stop, grad_mode, flight, received, atomic_in, raw_skips,
                                  t0)

        return self._outputs[:m] if m != self.chunks else self._outputs

    def _forward_one(self, mb: int, m: int, stop: int, grad_mode: bool, flight: '_Flight',
                     received: TensorOrTensors, atomic_in: bool, raw_skips: List[Tensor],
                     t0: Any) -> None:
        """Everything stage-local for micro-batch ``mb`` once its inputs have arrived."""
        tp = self.transport
        moved = to(self.device, received)
        boundary = tuple(self._as_leaf(t, grad_mode and not self.is_first)
                         for t in ((moved,) if atomic_in else moved))
        skips = [self._as_leaf(to(self.device, t), grad_mode) for t in raw_skips]
        t1 = self._mark()

        leaves = boundary + tuple(skips)
        flight.leaves = leaves
        flight.n_boundary = len(boundary)
        flight.atomic_in = atomic_in
        function = self._stage_function(len(boundary), atomic_in, flight)

        if self.cuda_graphs:
            staged = self._graphed_stage(len(boundary), atomic_in, flight)
            flight.staged = staged
            with nvtx.annotate(f'fwd mb{mb} stage{self.rank}'):
                produced, flight.rng = staged.forward(leaves)
            flight.n_outputs, flight.atomic = staged.meta['n_outputs'], staged.meta['atomic']
            if self.is_last and grad_mode:
                # the caller differentiates its loss w.r.t. these leaves; backward() feeds
                # their .grad to the backward graph
                for y in produced[:flight.n_outputs]:
                    if y.is_floating_point():
                        y.requires_grad_()
        elif mb < stop:
            with nvtx.annotate(f'fwd mb{mb} stage{self.rank} (checkpointed)'):
                chk = Checkpointing(function, Batch(leaves))
                out_batch = chk.checkpoint()
                chk.recompute(out_batch)
                produced = tuple(out_batch)
        else:
            with nvtx.annotate(f'fwd mb{mb} stage{self.rank}'):
                produced = function(leaves)

        if t0 is not None:
            self._marks.append(('fwd', t0, t1, self._mark()))
        outs = produced[:flight.n_outputs]
        stashes = produced[flight.n_outputs:]
        flight.outputs = produced
        self._flights[mb] = flight
        self._outputs[mb] = outs[0] if flight.atomic else outs

        if not self.is_last:
            tp.send(self.rank + 1, ('fwd', mb, 0),
                    outs[0].detach() if flight.atomic else tuple(y.detach() for y in outs))
        for (slot, dst, _), t in zip(self._routes_out, stashes):
            tp.send(dst, ('skip', mb, slot), t.detach())

    @staticmethod
    def _as_leaf(t: Tensor, grad_mode: bool) -> Tensor:
        t = t.detach()
        if grad_mode and t.is_floating_point():
            t.requires_grad_()
        return t

    # -- backward ------------------------------------------------------------------
    def backward(self, losses: Optional[List[Tensor]],
                 output_grads: Optional[List[TensorOrTensors]] = None) -> None:
        """Back-propagates every micro-batch (latest first) and passes gradients upstream.

        The last rank passes its per-micro-batch losses - or, instead, the gradients of its
        per-micro-batch outputs as ``output_grads`` (one tensor or tuple per micro-batch; used
        when the loss is computed outside, on the gathered output); other ranks pass ``None``
        and receive output gradients from the next stage.
        """
        tp = self.transport
        if losses is not None or output_grads is not None:
            assert self.is_last, 'only the last stage takes losses'
        with self.device_lock:
            for staged in self._graphed.values():
                staged.refresh_derived_weights()     # once per step, outside the graphs

        for mb in reversed(range(self._live)):
            flight = self._flights[mb]
            assert flight is not None, 'backward() called before forward()'
            produced = flight.outputs
            outs = produced[:flight.n_outputs]
            stashes = produced[flight.n_outputs:]

            if flight.staged is not None:
                self._graphed_backward(mb, flight, losses, output_grads)
                self._flights[mb] = None
                self._outputs[mb] = None
                continue

            roots: List[Tensor] = []
            grads: List[Optional[Tensor]] = []
            if losses is not None:
                roots.append(losses[mb])
                grads.append(None)
            elif output_grads is not None:
                given = output_grads[mb]
                for y, g in zip(outs, (given,) if torch.is_tensor(given) else given):
                    if y.requires_grad and g is not None:
                        roots.append(y)
                        grads.append(g)
            else:
                received = tp.recv(self.rank + 1, ('bwd', mb, 0),
                                   like=outs[0] if flight.atomic else outs)
                received = to(self.device, received)
                for y, g in zip(outs, (received,) if torch.is_tensor(received) else received):
                    if y.requires_grad:
                        roots.append(y)
                        grads.append(g)
            for (slot, dst, _), t in zip(self._routes_out, stashes):
                if not t.is_floating_point():
                    continue
                g = tp.recv(dst, ('skip_grad', mb, slot), like=t)
                if t.requires_grad:
                    roots.append(t)
                    grads.append(to(self.device, g))

            if roots:
                with nvtx.annotate(f'bwd mb{mb} stage{self.rank}'):
                    torch.autograd.backward(roots, grads)

            leaves = flight.leaves
            if not self.is_first:
                gin = tuple(self._grad_of(x) for x in leaves[:flight.n_boundary])
                tp.send(self.rank - 1, ('bwd', mb, 0),
                        gin[0] if flight.atomic_in else gin)
            for (slot, src, _), x in zip(self._routes_in, leaves[flight.n_boundary:]):
                if x.is_floating_point():
                    tp.send(src, ('skip_grad', mb, slot), self._grad_of(x))

            self._flights[mb] = None
            self._outputs[mb] = None

    def _graphed_stage(self, n_boundary: int, atomic_in: bool, flight: _Flight) -> Any:
        key = (n_boundary, atomic_in)
        staged = self._graphed.get(key)
        if staged is None:
            from torchgpipe_b200.runtime.graphs import GraphedStage
            meta: Dict[str, Any] = {}
            probe = _Flight()
            inner = self._stage_function(n_boundary, atomic_in, probe)

            def function(inputs: Tensors) -> Tensors:
                out = inner(inputs)
                meta['n_outputs'], meta['atomic'] = probe.n_outputs, probe.atomic
                return out

            staged = GraphedStage(function, self.module, torch.device(self.device))
            staged.meta = meta
            self._graphed[key] = staged
        return staged

    def _graphed_backward(self, mb: int, flight: _Flight, losses: Optional[List[Tensor]],
                          output_grads: Optional[List[TensorOrTensors]] = None) -> None:
        tp = self.transport
        produced = flight.outputs
        outs = produced[:flight.n_outputs]
        stashes = produced[flight.n_outputs:]
        gouts: List[Optional[Tensor]] = []
        t0 = self._mark()
        if losses is not None:
            if losses[mb].requires_grad:
                losses[mb].backward()
            gouts.extend(y.grad for y in outs)
        elif output_grads is not None:
            given = output_grads[mb]
            gouts.extend((given,) if torch.is_tensor(given) else given)
        else:
            received = tp.recv(self.rank + 1, ('bwd', mb, 0),
                               like=outs[0] if flight.atomic else outs)
            self._count_bytes(received)
            gouts.extend((received,) if torch.is_tensor(received) else received)
        for (slot, dst, _), t in zip(self._routes_out, stashes):
            if not t.is_floating_point():
                gouts.append(None)
                continue
            g = tp.recv(dst, ('skip_grad', mb, slot), like=t)
            self._count_bytes(g)
            gouts.append(g)

        with self.device_lock:          # every message is here: nothing below waits for a peer
            gouts = [None if g is None else to(self.device, g) for g in gouts]
            t1 = self._mark()
            with nvtx.annotate(f'bwd mb{mb} stage{self.rank}'):
                gin = flight.staged.backward(flight.leaves, flight.rng, gouts)
            if t0 is not None:
                self._marks.append(('bwd', t0, t1, self._mark()))

            leaves = flight.leaves
            nb = flight.n_boundary
            if not self.is_first:
                grads = tuple(self._like_leaf(x, g) for x, g in zip(leaves[:nb], gin[:nb]))
                tp.send(self.rank - 1, ('bwd', mb, 0), grads[0] if flight.atomic_in else grads)
            for (slot, src, _), x, g in zip(self._routes_in, leaves[nb:], gin[nb:]):
                if x.is_floating_point():
                    tp.send(src, ('skip_grad', mb, slot), self._like_leaf(x, g))

    @staticmethod
    def _like_leaf(leaf: Tensor, grad: Optional[Tensor]) -> Tensor:
        """The gradient to ship for ``leaf``: the peer's receive buffer has the (dense) layout of
        the tensor it sent forward and the transport moves raw bytes, so match that layout."""
        if grad is None:
            return torch.zeros_like(leaf)
        if grad.stride() != leaf.stride():
            fmt = torch.channels_last if (leaf.dim() == 4 and not leaf.is_contiguous()) \
                else torch.contiguous_format
            grad = grad.contiguous(memory_format=fmt)
        return grad

    @staticmethod
    def _grad_of(leaf: Tensor) -> Tensor:
        grad = leaf.grad
        if grad is None:
            return torch.zeros_like(leaf)
        leaf.grad = None
        if grad.stride() != leaf.stride():
            # the peer allocates its receive buffer with the layout of the tensor it sent
            # forward; ship the gradient in that same (dense) layout
            fmt = torch.channels_last if (leaf.dim() == 4 and not leaf.is_contiguous()) \
                else torch.contiguous_format
            grad = grad.contiguous(memory_format=fmt)
        return grad


class DistributedGPipeDataLoader:
    """Feeds ``(data, target)`` pairs to the right ranks.

    Rank 0 iterates the real loader, yields ``(data, None)`` and ships each
    target to the last stage; the last stage yields ``(None, target)``; middle
    stages yield ``(None, None)``. ``len()`` is ``num_iterations``.
    Contract: reference ``torchgpipe/distributed/gpipe.py:197-275``.

    With ``transport`` (and ``last_stage_rank``) targets travel through
    ``torch.distributed``; otherwise through the ``context`` target channel via
    the ``_put``/``_get`` hooks.
    """

    @staticmethod
    def _put(name: str, value: TensorOrTensors) -> None:
        context.put_target(name, value)

    @staticmethod
    def _get(name: str) -> TensorOrTensors:
        return context.get_target(name)

    def __init__(self,
                 data_loader: Optional[Iterable],
                 rank: int,
                 chunks: int,
                 num_iterations: int,
                 last_stage: bool,
                 last_stage_name: str,
                 *,
                 transport: Optional[Transport] = None,
                 last_stage_rank: Optional[int] = None,
                 ) -> None:
        self._data_loader = data_loader
        self._rank = rank
        self._chunks = chunks
        self._num_iterations = num_iterations
        self._last_stage = last_stage
        self._last_stage_name = last_stage_name
        self._transport = transport
        self._last_stage_rank = last_stage_rank
        self._step = 0

    def _send_target(self, target: TensorOrTensors) -> None:
        if self._transport is not None and self._last_stage_rank is not None:
            tensors = (target,) if torch.is_tensor(target) else tuple(target)
            device = getattr(self._transport, 'device', None)
            if device is not None:      # NCCL moves device memory only; loaders yield host tensors
                tensors = tuple(t.to(device, non_blocking=True) for t in tensors)
            self._transport.send(self._last_stage_rank, ('target', 0, 0), tensors)
        else:
            type(self)._put(self._last_stage_name, target)

    def _recv_target(self) -> TensorOrTensors:
        if self._transport is not None and self._last_stage_rank is not None:
            got = self._transport.recv(0, ('target', 0, 0))
            return got[0] if len(got) == 1 else got
        return type(self)._get(self._last_stage_name)

    def _first_stage_iter(self) -> Iterator[Tuple[Any, Any]]:
        assert self._data_loader is not None
        for (data, target), _ in zip(self._data_loader, range(self._num_iterations)):
            if self._last_stage:
                yield (data, target)        # single-stage pipeline
            else:
                self._send_target(target)
                yield (data, None)

    def _last_stage_iter(self) -> Iterator[Tuple[Any, Any]]:
        for _ in range(self._num_iterations):
            yield (None, self._recv_target())

    def _middle_stage_iter(self) -> Iterator[Tuple[Any, Any]]:
        for _ in range(self._num_iterations):
            yield (None, None)

    def __iter__(self) -> Iterator[Tuple[Any, Any]]:
        if self._rank == 0:
            return self._first_stage_iter()
        if self._last_stage:
            return self._last_stage_iter()
        return self._middle_stage_iter()

    def __len__(self) -> int:
        return self._num_iterations
