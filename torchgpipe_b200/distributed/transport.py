"""Point-to-point transports for the multi-process pipeline.

A transport moves tuples of tensors between stage ranks. Messages are
identified by ``(kind, micro-batch, slot)`` where ``kind`` is one of
``'fwd'`` (boundary activations), ``'bwd'`` (boundary gradients), ``'skip'`` /
``'skip_grad'`` (cross-stage ``@skippable`` tensors, ``slot`` = route index)
and ``'target'``.

* :class:`ChannelTransport` - in-process queues (the reference's
  ``context`` channels, ``torchgpipe/distributed/context.py``). Boundary
  traffic goes through ``DistributedGPipe._put/_get`` so tests can patch them
  exactly like the reference's tests do.
* :class:`ProcessGroupTransport` - ``torch.distributed`` send/recv. With the
  NCCL backend tensors go GPU->GPU over NVLink/NVSwitch and never touch host
  memory (the reference copies to CPU and pickles through RPC,
  ``torchgpipe/distributed/gpipe.py:94-96,176-177``). Shapes are announced by
  a small header the first time a message slot is used and cached afterwards
  (static shapes per micro-batch index), so steady-state traffic is payload
  only. For a pair of ranks, both ends issue their operations in the same
  program order, which is what NCCL point-to-point requires.

Failure detection (the reference's ``Queue.get`` waits forever,
``torchgpipe/distributed/context.py:141``): every receive has a deadline.
On CPU (gloo) the wait itself times out and raises ``TimeoutError``. On GPUs a
receive only *orders* the compute stream behind the transfer, so the host
cannot time out inside it; a watchdog thread follows a CUDA event recorded
after each receive and, when one stays pending past the deadline, reports
which message is missing, aborts the process group and ends the process so the
launcher tears the job down instead of leaving every rank spinning.
"""
import datetime
import os
import sys
import threading
import time
from collections import defaultdict
from queue import Queue
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

__all__ = ['Transport', 'ChannelTransport', 'ProcessGroupTransport']

Key = Tuple[str, int, int]
Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]

_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64,
           torch.int32, torch.int16, torch.int8, torch.uint8, torch.bool]
_MAX_TENSORS = 8
_MAX_DIMS = 6
_HEADER_LEN = 2 + _MAX_TENSORS * (2 + _MAX_DIMS)


class Transport:
    """Interface: ``send`` is asynchronous, ``recv`` returns usable tensors."""

    def send(self, dst: int, key: Key, value: TensorOrTensors) -> None:
        raise NotImplementedError

    def recv(self, src: int, key: Key, like: Optional[TensorOrTensors] = None,
             ) -> TensorOrTensors:
        """Receives a tensor or a tuple of tensors (whichever was sent).

        ``like`` gives the expected shapes/dtypes when the receiver already
        knows them (gradients mirror the forward tensors)."""
        raise NotImplementedError

    def flush(self) -> None:
        """Waits (on the host) until every send issued so far has left this rank."""

    def prepare_pairs(self, pairs: Sequence[Tuple[int, int]]) -> None:
        """Announces which (lower stage rank, higher stage rank) pairs will exchange messages.

        Called once by every rank with the same list, in the same order, before any traffic."""


class ChannelTransport(Transport):
    """Queues inside one process; ranks are threads (tests) or RPC peers."""

    _extra: Dict[Tuple[str, Key], Queue] = defaultdict(Queue)

    def __init__(self, workers: Dict[int, str], rank: int, owner: Any,
                 timeout: Optional[float] = None) -> None:
        self.workers = workers
        self.rank = rank
        self.owner = owner          # the DistributedGPipe class (for the _put/_get hooks)
        self.timeout = timeout

    def send(self, dst: int, key: Key, value: TensorOrTensors) -> None:
        kind, mb, _ = key
        name = self.workers[dst]
        if kind == 'fwd':
            self.owner._put(name, mb, value, False)
        elif kind == 'bwd':
            self.owner._put(name, mb, value, True)
        else:
            ChannelTransport._extra[(name, key)].put(value)

    def recv(self, src: int, key: Key, like: Optional[TensorOrTensors] = None,
             ) -> TensorOrTensors:
        kind, mb, _ = key
        name = self.workers[self.rank]
        if kind == 'fwd':
            return self.owner._get(name, mb, False)
        if kind == 'bwd':
            return self.owner._get(name, mb, True)
        from torchgpipe_b200.distributed.context import _take
        return _take(ChannelTransport._extra[(name, key)], self.timeout,
                     f'{kind} message of micro-batch {mb} for worker {name!r}')


class ProcessGroupTransport(Transport):
    """``torch.distributed`` point-to-point (NCCL on GPUs, gloo on CPU)."""

    def __init__(self, device: Optional[torch.device] = None, group: Any = None,
                 ranks: Optional[Dict[int, int]] = None, static_shapes: bool = True,
                 pair_groups: bool = False, timeout: Optional[float] = None) -> None:
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError('torch.distributed must be initialised before creating a '
                               'ProcessGroupTransport')
        self.dist = dist
        self.group = group
        self.device = device if device is not None else torch.device('cpu')
        self.ranks = ranks                  # stage rank -> global rank (identity by default)
        self.static_shapes = static_shapes
        self._pair_groups: Dict[Tuple[int, int], Any] = {}
        self._use_pair_groups = pair_groups
        self._known_send: Dict[Tuple[int, Key], Any] = {}
        self._known_recv: Dict[Tuple[int, Key], Any] = {}
        self._pending: List[Any] = []
        if timeout is None:
            timeout = float(os.environ.get('TGB_P2P_TIMEOUT', '1800'))
        #: seconds a receive may stay incomplete (<= 0: no deadline)
        self.timeout = timeout if timeout > 0 else None
        self._watch: List[Tuple[Any, float, str]] = []
        self._watch_lock = threading.Lock()
        self._watchdog: Optional[threading.Thread] = None

    # -- failure detection -----------------------------------------------------
    def _watch_receive(self, what: str) -> None:
        """GPU path: follow the receive that was just ordered into the current stream."""
        if self.timeout is None or self.device.type != 'cuda':
            return
        event = torch.cuda.Event()
        event.record(torch.cuda.current_stream(self.device))
        with self._watch_lock:
            self._watch.append((event, time.monotonic() + self.timeout, what))
            if self._watchdog is None:
                self._watchdog = threading.Thread(target=self._watchdog_loop, daemon=True,
                                                  name='tgb-p2p-watchdog')
                self._watchdog.start()

    def _watchdog_loop(self) -> None:
        period = min(1.0, max(0.05, (self.timeout or 1.0) / 20))
        while True:
            time.sleep(period)
            now = time.monotonic()
            with self._watch_lock:
                self._watch = [w for w in self._watch if not w[0].query()]
                expired = [w for w in self._watch if w[1] < now]
            if expired:
                self._abort(expired[0][2])

    def _abort(self, what: str) -> None:
        print(f'[torchgpipe_b200] rank {self.dist.get_rank()}: {what} did not arrive within '
              f'{self.timeout:g} s - a peer stage died or stalled; aborting the process group',
              file=sys.stderr, flush=True)
        try:
            from torch.distributed.distributed_c10d import _abort_process_group
            _abort_process_group(self.group)
        except Exception:       # older releases: nothing to abort with
            pass
        os._exit(70)

    def _global(self, stage_rank: int) -> int:
        return stage_rank if self.ranks is None else self.ranks[stage_rank]

    def prepare_pairs(self, pairs: Sequence[Tuple[int, int]]) -> None:
        """One communicator (and so one stream) per communicating pair of stages.

        NCCL point-to-point is a rendezvous and all operations of one communicator run in
        issue order on one stream. With a single communicator, a skip tensor sent from stage 1
        to stage 6 parks stage 1's stream until stage 6 gets around to receiving it, and every
        later send of stage 1 - including the next micro-batch for stage 2 - queues behind it:
        the pipeline degenerates to lock-step (measured: 8 stages slower than 4).

        ``torch.distributed`` already gives every rank pair its own NCCL communicator for
        point-to-point traffic when the process group is initialised *lazily* (no
        ``device_id=`` in ``init_process_group``); eager initialisation funnels everything
        through the one communicator. So the default here is to do nothing and to initialise
        lazily (see ``bench.py``); ``pair_groups=True`` builds explicit two-rank groups instead
        (used with gloo in the tests; with eagerly initialised NCCL it hung on 4 GPUs).
        """
        if not self._use_pair_groups or self.dist.get_world_size(self.group) <= 2:
            return
        me = self.dist.get_rank()
        for a, b in pairs:
            key = (min(a, b), max(a, b))
            if key in self._pair_groups:
                continue
            members = [self._global(key[0]), self._global(key[1])]
            # every rank must take part in every new_group call, member or not
            group = self.dist.new_group(ranks=members)
            if me in members:
                self._pair_groups[key] = group

    def _group_for(self, peer_stage: int, my_stage: Optional[int] = None) -> Any:
        if not self._pair_groups:
            return self.group
        me = self._stage_of_me()
        return self._pair_groups.get((min(me, peer_stage), max(me, peer_stage)), self.group)

    def _stage_of_me(self) -> int:
        if not hasattr(self, '_my_stage'):
            g = self.dist.get_rank()
            if self.ranks is None:
                self._my_stage = g
            else:
                self._my_stage = next(s for s, r in self.ranks.items() if r == g)
        return self._my_stage

    # -- headers -------------------------------------------------------------
    def _encode(self, tensors: Sequence[Tensor], atomic: bool) -> Tensor:
        if len(tensors) > _MAX_TENSORS:
            raise ValueError(f'at most {_MAX_TENSORS} tensors per message')
        header = [len(tensors), int(atomic)]
        for t in tensors:
            if t.dim() > _MAX_DIMS:
                raise ValueError(f'at most {_MAX_DIMS} dims per tensor')
            dims = list(t.shape) + [0] * (_MAX_DIMS - t.dim())
            cl = t.dim() == 4 and not t.is_contiguous() and \
                t.is_contiguous(memory_format=torch.channels_last)
            header += [_DTYPES.index(t.dtype) + (100 if cl else 0), t.dim()] + dims
        header += [0] * (_HEADER_LEN - len(header))
        return torch.tensor(header, dtype=torch.int64, device=self.device)

    @staticmethod
    def _decode(header: Tensor) -> Tuple[bool, List[Tuple[torch.dtype, Tuple[int, ...]]]]:
        values = header.tolist()
        out = []
        pos = 2
        for _ in range(values[0]):
            dtype = _DTYPES[values[pos] % 100]
            ndim = values[pos + 1]
            shape = tuple(values[pos + 2:pos + 2 + ndim])
            out.append((dtype, shape, values[pos] >= 100))
            pos += 2 + _MAX_DIMS
        return bool(values[1]), out

    # -- traffic ---------------------------------------------------------------
    def send(self, dst: int, key: Key, value: TensorOrTensors) -> None:
        peer = self._global(dst)
        group = self._group_for(dst)
        slot = (dst, key)
        atomic = torch.is_tensor(value)
        tensors = (value,) if atomic else tuple(value)
        if key[0] not in ('bwd', 'skip_grad'):
            # Gradient messages never need a header: the receiver passes `like`.
            meta = (atomic, tuple((t.dtype, tuple(t.shape), t.dim() == 4 and not t.is_contiguous()
                                   and t.is_contiguous(memory_format=torch.channels_last))
                                  for t in tensors))
            known = self._known_send.get(slot)
            if not self.static_shapes or known is None:
                self._pending.append(
                    self.dist.isend(self._encode(tensors, atomic), peer, group=group))
                self._known_send[slot] = meta
            elif known != meta:
                # the receiver allocates from the header it cached on the first message
                raise ValueError(
                    f'message {key} to stage {dst} changed shape, dtype or memory format '
                    f'({known[1]} -> {meta[1]}) but this transport was created with '
                    'static_shapes=True; use '
                    'ProcessGroupTransport(static_shapes=False) for varying batch sizes')
        # Completed sends are dropped right away: a pending handle keeps its tensor alive, and
        # activations are large (a leaked handle per message grows memory by GBs per step).
        self._reap()
        for t in tensors:
            t = t.detach()
            if not (t.is_contiguous() or t.is_contiguous(memory_format=torch.channels_last)):
                t = t.contiguous()
            self._pending.append(self.dist.isend(t, peer, group=group))

    def recv(self, src: int, key: Key, like: Optional[TensorOrTensors] = None,
             ) -> TensorOrTensors:
        peer = self._global(src)
        group = self._group_for(src)
        slot = (src, key)
        if like is not None:
            atomic = torch.is_tensor(like)
            meta = [(t.dtype, tuple(t.shape), t.dim() == 4 and not t.is_contiguous()
                     and t.is_contiguous(memory_format=torch.channels_last))
                    for t in ((like,) if atomic else like)]
        elif self.static_shapes and slot in self._known_recv:
            atomic, meta = self._known_recv[slot]
        else:
            header = torch.empty(_HEADER_LEN, dtype=torch.int64, device=self.device)
            self._wait(self.dist.irecv(header, peer, group=group),
                       f'header of message {key} from stage {src}')
            self._watch_receive(f'header of message {key} from stage {src}')
            atomic, meta = self._decode(header)
            self._known_recv[slot] = (atomic, meta)
        out = []
        works = []
        for dtype, shape, channels_last in meta:
            buf = torch.empty(shape, dtype=dtype, device=self.device,
                              memory_format=torch.channels_last if channels_last
                              else torch.contiguous_format)
            works.append(self.dist.irecv(buf, peer, group=group))
            out.append(buf)
        what = f'message {key} from stage {src}'
        for w in works:
            self._wait(w, what)
        self._watch_receive(what)
        return out[0] if atomic else tuple(out)

    def _wait(self, work: Any, what: str) -> None:
        if self.device.type == 'cuda' or self.timeout is None:
            work.wait()     # NCCL: orders the current stream behind the transfer
            return
        try:
            done = work.wait(datetime.timedelta(seconds=self.timeout))
        except RuntimeError as exc:     # gloo reports an expired wait as RuntimeError
            if 'time' in str(exc).lower():
                raise TimeoutError(f'{what} did not arrive within {self.timeout:g} s: the peer '
                                   'stage died or never sent it') from None
            raise
        if done is False:
            raise TimeoutError(f'{what} did not arrive within {self.timeout:g} s')

    def _reap(self) -> None:
        self._pending = [w for w in self._pending if not w.is_completed()]

    def flush(self) -> None:
        for w in self._pending:
            w.wait()
        self._pending.clear()
