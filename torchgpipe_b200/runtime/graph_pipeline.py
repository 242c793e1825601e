"""CUDA-graph engine behind ``GPipe(..., cuda_graphs=True)`` (single process, N devices).

The eager single-process engine (``runtime/pipeline.py``) issues every kernel of every cell from
Python; with the sm_100a operators a stage pass of one micro-batch takes less GPU time than the
host needs to launch it, so the reference's own API (``torchgpipe.GPipe(nn.Sequential, balance=,
chunks=, checkpoint=)``, reference ``torchgpipe/gpipe.py:211-380``) would be host-bound here.
This engine keeps that API and replaces what happens underneath:

* every partition is driven by the stage engine of the multi-process mode
  (:class:`~torchgpipe_b200.distributed.gpipe.DistributedGPipe` with
  :class:`~torchgpipe_b200.runtime.graphs.GraphedStage`): one ``cudaGraphLaunch`` per stage pass,
  recomputation by replaying the forward graph, ``except_last`` for free;
* the stages run on persistent threads of this process, one per partition, each feeding its own
  device; they hand micro-batches to each other through in-process mailboxes (tensors are passed
  by reference, the consumer's ``.to(device)`` is the peer copy over NVLink);
* the fill/drain order is not scheduled explicitly: each stage walks its micro-batches in order
  and blocks on its mailbox, which yields the same clock-cycle pipeline as
  ``runtime/schedule.py``;
* to the caller it is one autograd node: ``forward`` returns the gathered output, and
  ``output.backward()`` (through any loss) runs the backward pipeline, which adds the parameter
  gradients into ``param.grad``.

Limits of the mode (checked, with clear errors): CUDA devices only, no gradient with respect to
the pipeline *input*, no ``deferred_batch_norm``, layers must not synchronise with the host.
"""
import queue
import threading
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor, nn

from torchgpipe_b200 import microbatch
from torchgpipe_b200.distributed.transport import Transport

__all__ = ['GraphedPipeline', 'PipelineAborted']

Tensors = Tuple[Tensor, ...]
TensorOrTensors = Union[Tensor, Tensors]


class PipelineAborted(RuntimeError):
    """Another stage of the pipeline failed; this stage stopped waiting for it."""


class _Mailboxes(Transport):
    """Mailboxes between the stage threads of one process. ``recv`` wakes up regularly to see
    whether another stage has failed, so an exception anywhere ends the step everywhere."""

    def __init__(self, rank: int, boxes: Dict[Any, 'queue.Queue'], lock: threading.Lock,
                 failed: threading.Event) -> None:
        self.rank, self.boxes, self.lock, self.failed = rank, boxes, lock, failed

    def _box(self, dst: int, src: int, key: Any) -> 'queue.Queue':
        slot = (dst, src, key)
        with self.lock:
            box = self.boxes.get(slot)
            if box is None:
                box = self.boxes[slot] = queue.Queue()
            return box

    def send(self, dst: int, key: Any, value: TensorOrTensors) -> None:
        self._box(dst, self.rank, key).put(value)

    def recv(self, src: int, key: Any, like: Optional[TensorOrTensors] = None) -> TensorOrTensors:
        box = self._box(self.rank, src, key)
        while True:
            try:
                return box.get(timeout=0.1)
            except queue.Empty:
                if self.failed.is_set():
                    raise PipelineAborted(f'stage {self.rank} stopped: another stage failed')


class _StageThread(threading.Thread):
    """Persistent driver of one stage: executes ``(function, args)`` jobs in order."""

    def __init__(self, index: int, device: torch.device) -> None:
        super().__init__(daemon=True, name=f'tgb-graph-stage{index}')
        self.device = device
        self.jobs: 'queue.Queue' = queue.Queue()
        self.start()

    def run(self) -> None:
        torch.cuda.set_device(self.device)
        while True:
            job = self.jobs.get()
            if job is None:
                return
            fn, box = job
            try:
                box.append((True, fn()))
            except BaseException as exc:       # delivered to the caller, never lost
                box.append((False, exc))
            finally:
                box_done = box[0]
                box_done.set()

    def submit(self, fn) -> list:
        box: list = [threading.Event()]
        self.jobs.put((fn, box))
        return box


class _PipelineFunction(torch.autograd.Function):
    """The whole pipeline as one autograd node. ``anchor`` is a dummy input that requires grad so
    the outputs do, whatever the data input looks like."""

    @staticmethod
    def forward(ctx, engine: 'GraphedPipeline', anchor: Tensor, *inputs: Tensor):  # type: ignore[override]
        outputs, sizes = engine._forward(inputs, training_graph=True)
        ctx.engine = engine
        ctx.sizes = sizes
        ctx.n_inputs = len(inputs)
        ctx.mark_non_differentiable(*[o for o in outputs if not o.is_floating_point()])
        return tuple(outputs)

    @staticmethod
    def backward(ctx, *grads: Optional[Tensor]):  # type: ignore[override]
        ctx.engine._backward(grads, ctx.sizes)
        return (None, None) + (None,) * ctx.n_inputs


class GraphedPipeline:
    """Runs ``partitions`` (each on its device) as a CUDA-graph pipeline inside one process."""

    def __init__(self, partitions: Sequence[nn.Sequential], devices: Sequence[torch.device],
                 chunks: int, checkpoint: str) -> None:
        from torchgpipe_b200.distributed.gpipe import DistributedGPipe
        if any(torch.device(d).type != 'cuda' for d in devices):
            raise ValueError('cuda_graphs=True needs every partition on a CUDA device')
        self.devices = [torch.device(d) for d in devices]
        self.chunks = chunks
        n = len(partitions)
        self._failed = threading.Event()
        self._boxes: Dict[Any, 'queue.Queue'] = {}
        self._lock = threading.Lock()
        workers = {r: f'graph-stage{r}-{id(self)}' for r in range(n)}
        self.stages = [
            DistributedGPipe(None, r, workers, None, chunks, device=self.devices[r],  # type: ignore[arg-type]
                             checkpoint=checkpoint, cuda_graphs=True, partitions=list(partitions),
                             transport=_Mailboxes(r, self._boxes, self._lock, self._failed))
            for r in range(n)]
        # Stages that share a device (repeated entries in ``devices``) take turns: a graph capture
        # on a device must not overlap with work another thread enqueues on that device's
        # default stream (the capture would be invalidated), so there is one lock per device.
        locks: Dict[torch.device, Any] = {}
        for r, stage in enumerate(self.stages):
            if self.devices.count(self.devices[r]) > 1:
                stage.device_lock = locks.setdefault(self.devices[r], threading.RLock())
        # a single stage needs no thread: the caller's thread drives it
        self._threads = [_StageThread(r, self.devices[r]) for r in range(n)] if n > 1 else []
        self._anchor = torch.zeros((), device=self.devices[-1], requires_grad=True)

    def close(self) -> None:
        """Ends the stage threads (they are daemons, but a long-lived process that builds many
        pipelines should not collect idle threads)."""
        for t in self._threads:
            t.jobs.put(None)
        self._threads = []

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:       # interpreter shutdown: queues may already be gone
            pass

    def native_launches_replayed(self) -> int:
        return sum(s.native_launches_replayed() for s in self.stages)

    # -- running the stages ------------------------------------------------------------
    def _run_all(self, jobs: List[Any]) -> List[Any]:
        """One job per stage, concurrently; returns their results or raises the first failure."""
        if not self._threads:
            return [jobs[0]()]
        self._failed.clear()
        with self._lock:
            self._boxes.clear()

        def guarded(fn):
            def run():
                try:
                    return fn()
                except BaseException:
                    self._failed.set()          # wakes every stage that waits on a mailbox
                    raise
            return run

        boxes = [t.submit(guarded(fn)) for t, fn in zip(self._threads, jobs)]
        results, failure = [], None
        for box in boxes:
            box[0].wait()
            ok, payload = box[1]
            if not ok and not isinstance(payload, PipelineAborted) and failure is None:
                failure = payload
            results.append(payload if ok else None)
        if failure is None:
            for box in boxes:
                if not box[1][0]:
                    failure = box[1][1]
                    break
        if failure is not None:
            raise failure
        return results

    def _forward(self, inputs: Tuple[Tensor, ...], training_graph: bool,
                 ) -> Tuple[List[Tensor], List[int]]:
        grad_mode = training_graph
        value: TensorOrTensors = inputs[0] if len(inputs) == 1 and self._atomic_in else inputs
        n = len(self.stages)

        def job(r: int):
            def run():
                with torch.cuda.device(self.devices[r]), torch.set_grad_enabled(grad_mode):
                    return self.stages[r].forward(value if r == 0 else None)
            return run

        outs = self._run_all([job(r) for r in range(n)])[-1]
        batches = [microbatch.Batch(o if torch.is_tensor(o) else tuple(o)) for o in outs]
        gathered = microbatch.gather(batches)
        self._atomic_out = torch.is_tensor(gathered)
        tensors = [gathered] if self._atomic_out else list(gathered)
        sizes = [b[0].size(0) for b in batches]
        return tensors, sizes

    def _backward(self, grads: Sequence[Optional[Tensor]], sizes: List[int]) -> None:
        n = len(self.stages)
        pieces = [None if g is None else g.split(sizes) for g in grads]
        per_mb: List[TensorOrTensors] = []
        for k in range(len(sizes)):
            members = tuple(None if p is None else p[k] for p in pieces)
            per_mb.append(members[0] if self._atomic_out else members)  # type: ignore[arg-type]

        def job(r: int):
            def run():
                with torch.cuda.device(self.devices[r]), torch.enable_grad():
                    if r == n - 1:
                        self.stages[r].backward(None, output_grads=per_mb)
                    else:
                        self.stages[r].backward(None)
            return run

        self._run_all([job(r) for r in range(n)])

    # -- public --------------------------------------------------------------------------
    def __call__(self, input: TensorOrTensors) -> TensorOrTensors:
        self._atomic_in = torch.is_tensor(input)
        inputs = (input,) if self._atomic_in else tuple(input)
        for p in self.stages[0].module.parameters():
            if not p.is_cuda:
                raise ValueError('cuda_graphs=True needs every partition on a CUDA device')
            break
        if torch.is_grad_enabled() and any(t.requires_grad for t in inputs):
            raise NotImplementedError(
                'GPipe(cuda_graphs=True) does not back-propagate into the pipeline input; '
                'detach the input or use the eager engine (cuda_graphs=False)')
        if torch.is_grad_enabled():
            outputs = _PipelineFunction.apply(self, self._anchor, *inputs)
        else:
            outputs, _ = self._forward(inputs, training_graph=False)
        return outputs[0] if self._atomic_out else tuple(outputs)
