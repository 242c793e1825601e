"""CUDA-graph replay of one pipeline stage (forward, and recompute + backward).

Published GPipe configurations use tens to thousands of micro-batches per step
(reference ``benchmarks/unet-speed/main.py:41-68``: 16-40; ResNet-101: up to
1667). On a B200 a stage's kernels for one micro-batch take well under the time
the host needs to issue them one by one, so an eager stage is launch-bound and
extra GPUs buy nothing. ``GraphedStage`` records the kernels of one stage once
per input signature and afterwards issues a whole stage pass with one
``cudaGraphLaunch``:

* ``forward(inputs)``  - copies the inputs into static buffers, replays the
  forward graph and returns *copies* of the outputs (no autograd graph is kept:
  the static activation buffers belong to the next replay);
* ``backward(inputs, rng_state, grad_outputs)`` - restores the RNG state of that
  forward, replays the forward graph again (this is GPipe's checkpoint
  recomputation), replays the backward graph, adds the parameter gradients into
  ``param.grad`` with one fused multi-tensor add and returns the input gradients.

So a graphed stage checkpoints every micro-batch: it keeps only the stage inputs
between forward and backward, the memory behaviour of ``checkpoint='always'``
(reference ``torchgpipe/checkpoint.py``), and one set of static activations per
stage instead of one per micro-batch. The exception costs nothing: in the
fill/drain schedule a stage's last forward (micro-batch m-1) is directly followed
by the backward of that same micro-batch, whose activations are then still
resident in the static buffers - ``backward`` notices that and skips the
recomputation replay, which is ``checkpoint='except_last'`` without holding any
extra memory.

Everything the native kernels do on the host (tensor-map encoding, launch
configuration) happens at capture time; derived weights (the tap-flipped dgrad
weights) live in persistent buffers that are refreshed *outside* the graph once
per optimiser step (``refresh_derived_weights``).
"""
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor, nn

__all__ = ['GraphedStage']

Signature = Tuple[object, ...]


def _signature(inputs: Sequence[Tensor], module: Optional[nn.Module] = None,
               grad: Optional[bool] = None) -> Signature:
    """Key of one capture. Besides the tensors' shape / dtype / layout it holds everything that
    changes WHICH kernels the stage function issues: ``module.training`` (Dropout, BatchNorm
    batch-vs-running statistics) and the grad mode (a no-grad forward records no backward)."""
    mode = (bool(module.training) if module is not None else True,
            torch.is_grad_enabled() if grad is None else bool(grad))
    return (mode,) + tuple((tuple(t.shape), t.dtype, t.requires_grad,
                            t.dim() == 4 and not t.is_contiguous()) for t in inputs)


class _Captured:
    """Graphs and static buffers for one input signature."""

    def __init__(self) -> None:
        self.fwd = torch.cuda.CUDAGraph()
        self.bwd: Optional[torch.cuda.CUDAGraph] = None
        self.static_inputs: List[Tensor] = []
        self.static_outputs: Tuple[Tensor, ...] = ()
        self.static_grad_outputs: List[Optional[Tensor]] = []
        self.static_grad_inputs: List[Optional[Tensor]] = []
        self.static_grad_params: List[Optional[Tensor]] = []
        self.native_launches_fwd = 0
        self.native_launches_bwd = 0


class GraphedStage:
    """Replays ``function(tuple of tensors) -> tuple of tensors`` from CUDA graphs.

    Args:
        function: the stage as a pure tensor function (it may use parameters of
            ``module``; it must not synchronise with the host).
        module: owner of the parameters and buffers ``function`` uses.
        device: CUDA device of the stage.
        warmup: eager iterations on a side stream before capture.
    """

    def __init__(self, function: Callable[[Tuple[Tensor, ...]], Tuple[Tensor, ...]],
                 module: nn.Module, device: torch.device, warmup: int = 2) -> None:
        self.function = function
        self.module = module
        self.device = device
        self.warmup = warmup
        self.params = [p for p in module.parameters() if p.requires_grad]
        self._captured: Dict[Signature, _Captured] = {}
        self._pool = None
        self._capture_stream: Optional[torch.cuda.Stream] = None
        self._derived = [m for m in module.modules() if hasattr(m, 'refresh_derived_weights')]
        # Buffers the forward graph updates in place (BatchNorm running statistics): the
        # recomputation replay must not apply those updates a second time.
        self._buffers = [b for b in module.buffers() if b.is_cuda]
        self._shadow = [torch.empty_like(b) for b in self._buffers]
        self.replays = 0
        self.native_launches_replayed = 0
        # (capture, rng_state object) of the forward replay whose activations are still in the
        # static buffers; any other replay of this stage (all captures share one pool) clears it
        self._resident: Optional[Tuple[_Captured, Tensor]] = None
        self.recomputes_skipped = 0

    # -- capture ---------------------------------------------------------------
    def _capture(self, inputs: Sequence[Tensor], grad: bool = True) -> _Captured:
        from torchgpipe_b200 import ops
        cap = _Captured()
        with torch.cuda.device(self.device):
            cap.static_inputs = [torch.empty_like(t).requires_grad_(t.requires_grad)
                                 for t in inputs]
            for s, t in zip(cap.static_inputs, inputs):
                s.data.copy_(t)
            diff_inputs = [s for s in cap.static_inputs if s.requires_grad]

            def run_backward(outputs: Tuple[Tensor, ...], grad_outputs: List[Optional[Tensor]]):
                pairs = [(o, g) for o, g in zip(outputs, grad_outputs) if g is not None]
                if not pairs or not (diff_inputs or self.params):
                    return ()
                return torch.autograd.grad([o for o, _ in pairs], diff_inputs + self.params,
                                           [g for _, g in pairs], allow_unused=True)

            # Warm-up on a side stream: lazy initialisation (kernel attributes, caches, cuDNN
            # plans) must not happen inside a capture.
            if self._buffers:
                torch._foreach_copy_(self._shadow, self._buffers)   # warm-up must leave no trace
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                for _ in range(self.warmup):
                    with torch.set_grad_enabled(grad):
                        outs = self.function(tuple(cap.static_inputs))
                    gouts = [torch.ones_like(o) if o.requires_grad else None for o in outs]
                    run_backward(outs, gouts)
                    del outs, gouts
            torch.cuda.current_stream(self.device).wait_stream(side)
            if self._buffers:
                torch._foreach_copy_(self._buffers, self._shadow)
            torch.cuda.synchronize(self.device)

            if self._pool is None:
                self._pool = torch.cuda.graph_pool_handle()

            before = ops.launch_count()
            # relaxed: the stages of a single-process pipeline are threads that keep allocating,
            # freeing and querying events on their own devices while this one records, and this
            # thread may drop its last reference to a tensor that lives on a neighbour's device
            # (which makes that device's allocator record / query events) - calls the stricter
            # capture modes refuse. Work that truly cannot be captured (a host sync on the
            # capturing stream) still fails in this mode.
            # The capture stream must belong to this stage's device: torch's default capture
            # stream is one per process, created on whichever device captured first, and with
            # it a stage on another device records nothing (its kernels run eagerly, once).
            if self._capture_stream is None:
                self._capture_stream = torch.cuda.Stream(self.device)
            with torch.cuda.graph(cap.fwd, pool=self._pool, stream=self._capture_stream,
                                  capture_error_mode='relaxed'):
                with torch.set_grad_enabled(grad):
                    cap.static_outputs = tuple(self.function(tuple(cap.static_inputs)))
            cap.native_launches_fwd = ops.launch_count() - before

            cap.static_grad_outputs = [torch.zeros_like(o) if o.requires_grad else None
                                       for o in cap.static_outputs]
            if any(g is not None for g in cap.static_grad_outputs) and \
                    (diff_inputs or self.params):
                cap.bwd = torch.cuda.CUDAGraph()
                before = ops.launch_count()
                with torch.cuda.graph(cap.bwd, pool=self._pool, stream=self._capture_stream,
                                      capture_error_mode='relaxed'):
                    grads = run_backward(cap.static_outputs, cap.static_grad_outputs)
                cap.native_launches_bwd = ops.launch_count() - before
                n_in = len(diff_inputs)
                it = iter(grads[:n_in])
                cap.static_grad_inputs = [next(it) if s.requires_grad else None
                                          for s in cap.static_inputs]
                cap.static_grad_params = list(grads[n_in:])
        return cap

    def _get(self, inputs: Sequence[Tensor], grad: Optional[bool] = None) -> _Captured:
        if grad is None:
            grad = torch.is_grad_enabled()
        sig = _signature(inputs, self.module, grad)
        cap = self._captured.get(sig)
        if cap is None:
            cap = self._capture(inputs, grad)
            self._captured[sig] = cap
        return cap

    # -- replay ----------------------------------------------------------------
    @staticmethod
    def _load(static: Sequence[Tensor], values: Sequence[Tensor]) -> None:
        for s, v in zip(static, values):
            if s.data_ptr() != v.data_ptr():
                s.data.copy_(v, non_blocking=True)

    def forward(self, inputs: Sequence[Tensor]) -> Tuple[Tuple[Tensor, ...], Tensor]:
        """Runs the stage without keeping activations.

        Returns ``(fresh output tensors, rng_state)``; hand ``rng_state`` back to
        :meth:`backward` so the recomputation draws the same random numbers. (The state is
        taken here, after a possible first-use capture, which itself consumes random numbers.)
        """
        cap = self._get(inputs)
        rng_state = torch.cuda.get_rng_state(self.device)
        self._load(cap.static_inputs, inputs)
        cap.fwd.replay()
        self.replays += 1
        self.native_launches_replayed += cap.native_launches_fwd
        self._resident = (cap, rng_state)
        return tuple(o.detach().clone() for o in cap.static_outputs), rng_state

    def refresh_derived_weights(self) -> None:
        """Recomputes weights derived from parameters (call once per optimiser step)."""
        for m in self._derived:
            m.refresh_derived_weights()

    def backward(self, inputs: Sequence[Tensor], rng_state: Optional[Tensor],
                 grad_outputs: Sequence[Optional[Tensor]]) -> List[Optional[Tensor]]:
        """Recomputes the stage for ``inputs`` and back-propagates ``grad_outputs``.

        Parameter gradients are accumulated into ``param.grad``; the gradients of the inputs
        that require grad are returned (``None`` for the others).
        """
        cap = self._get(inputs, grad=True)      # only a capture made with grad on has a backward
        if cap.bwd is None:
            return [None for _ in inputs]
        resident = self._resident
        self._resident = None
        if resident is not None and resident[0] is cap and resident[1] is rng_state \
                and rng_state is not None:
            # the forward replay of exactly this micro-batch was the last thing this stage ran:
            # its activations (and static inputs) are still in place
            self.recomputes_skipped += 1
            self.replays -= 1
            self.native_launches_replayed -= cap.native_launches_fwd
        else:
            if rng_state is not None:
                torch.cuda.set_rng_state(rng_state, self.device)
            self._load(cap.static_inputs, inputs)
            if self._buffers:
                torch._foreach_copy_(self._shadow, self._buffers)
            cap.fwd.replay()                         # checkpoint recomputation
            if self._buffers:
                torch._foreach_copy_(self._buffers, self._shadow)
        for s, g in zip(cap.static_grad_outputs, grad_outputs):
            if s is not None:
                if g is None:
                    s.zero_()
                else:
                    s.copy_(g, non_blocking=True)
        cap.bwd.replay()
        self.replays += 2
        self.native_launches_replayed += cap.native_launches_fwd + cap.native_launches_bwd

        have, add_to, add_from = [], [], []
        for p, g in zip(self.params, cap.static_grad_params):
            if g is None:
                continue
            if p.grad is None:
                p.grad = g.detach().clone()
                have.append(p)
            else:
                add_to.append(p.grad)
                add_from.append(g)
        if add_to:
            torch._foreach_add_(add_to, add_from)
        return [None if g is None else g.detach().clone() for g in cap.static_grad_inputs]
