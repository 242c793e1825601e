"""nn.Module front-ends of the fused sm_100a operators."""
from typing import Optional

import torch
from torch import Tensor, nn
import torch.nn.functional as F

from torchgpipe_b200.checkpoint import is_recomputing
from torchgpipe_b200.batchnorm import MiniBatchMoments
from torchgpipe_b200.ops import functional as TF

__all__ = ['Absorbed', 'ConvInstNormLeaky', 'Conv2dTC', 'ConvBatchNormAct', 'AbsorbedBatchNorm2d',
           'MaxPool2x2', 'UpsamplePopCat', 'LinearTC', 'Pool2dNHWC', 'GlobalAvgPoolNHWC',
           'BatchNormResidualAct', 'AmoebaStem', 'AmoebaFactorizedReduce', 'AmoebaCell']

_CL = torch.channels_last


class Absorbed(nn.Module):
    """Placeholder for a layer whose work now happens inside a preceding fused layer.

    Keeps the position and name of the original layer so ``balance`` lists and layer counts are
    unchanged by :func:`~torchgpipe_b200.ops.convert_to_sm100`.
    """

    def __init__(self, original: nn.Module, into: str) -> None:
        super().__init__()
        self.original_repr = repr(original)
        self.into = into

    def extra_repr(self) -> str:
        return f'{self.original_repr} -> fused into {self.into!r}'

    def forward(self, input):  # type: ignore[override]
        return input


def _conv_any(x: Tensor, weight: Tensor, stride, padding, stats_mode: int, cache: dict):
    """``(y, stats)`` of a bias-free convolution on whichever native path covers it
    (``TF.conv_kind``): implicit GEMM, packed-window stem, or - statistics by the moments
    kernel - the streaming head. ``None`` when no native path applies."""
    K, C, R, S = weight.shape
    kind = TF.conv_kind(C, K, (R, S), stride, 1, 1)
    if kind == 'tc':
        return TF.conv2d(x, weight, padding, stats_mode=stats_mode, cache=cache, stride=stride)
    if kind == 'stem' and padding[0] >= 0 and padding[1] >= 0:
        return TF.stem_conv2d(x, weight, padding, stride, stats_mode=stats_mode, cache=cache)
    if kind == 'head' and tuple(padding) == (0, 0):
        y = TF.head_conv2d(x, weight, None)
        stats = None
        if stats_mode:
            stats = TF.channel_moments(y, per_image=stats_mode == 1) if K % 4 == 0 else None
        return y, stats
    return None


class Conv2dTC(nn.Module):
    """``nn.Conv2d`` (groups 1, no dilation, strides up to 2x2) on the native kernels; same
    ``weight`` / ``bias`` parameters. Wide layers run the tcgen05 implicit GEMM, a 1-4 channel
    input the packed-window stem path, a 1x1 layer with at most 4 outputs the streaming head."""

    def __init__(self, conv: nn.Conv2d) -> None:
        super().__init__()
        self.in_channels, self.out_channels = conv.in_channels, conv.out_channels
        self.kernel_size, self.padding, self.stride = conv.kernel_size, conv.padding, conv.stride
        self.weight = conv.weight
        self.weight.data = self.weight.data.contiguous(memory_format=_CL)
        self.bias = conv.bias
        self.kind = TF.conv_kind(self.in_channels, self.out_channels, self.kernel_size,
                                 self.stride, 1, 1)
        self._cache: dict = {}

    def refresh_derived_weights(self) -> None:
        """Updates the cached derived weights in place after an optimiser step (graph mode)."""
        TF.refresh_dgrad_weight(self.weight, self._cache)
        TF.refresh_stem_weight(self.weight, self._cache)

    def extra_repr(self) -> str:
        return (f'{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, '
                f'stride={self.stride}, padding={self.padding}, bias={self.bias is not None}, '
                f'native={self.kind}')

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if input.is_cuda and input.dtype == torch.float32 and self.kind is not None:
            if self.kind == 'head' and tuple(self.padding) == (0, 0):
                return TF.head_conv2d(input, self.weight, self.bias)
            got = _conv_any(input, self.weight, self.stride, self.padding, 0, self._cache)
            if got is not None:
                y = got[0]
                if self.bias is not None:
                    y = TF.add_channel_bias(y, self.bias)
                return y
        return F.conv2d(input, self.weight, self.bias, self.stride, self.padding)


class ConvInstNormLeaky(nn.Module):
    """``Conv2d(bias=False) -> Dropout2d(p) -> InstanceNorm2d -> LeakyReLU`` as two kernels.

    The convolution's epilogue accumulates the per-(image, channel) sum and sum of squares, so
    the normalisation needs one read and one write of the activation; the channel dropout is a
    per-(n,c) factor folded into the normalisation coefficient (its mask comes from PyTorch's
    CUDA generator, so checkpoint recomputation reproduces it). Reference: the unfused cell of
    ``benchmarks/models/unet/__init__.py:42-48``.
    """

    def __init__(self, conv: nn.Conv2d, p: float, eps: float, slope: float) -> None:
        super().__init__()
        self.in_channels, self.out_channels = conv.in_channels, conv.out_channels
        self.kernel_size, self.padding, self.stride = conv.kernel_size, conv.padding, conv.stride
        self.weight = conv.weight
        self.weight.data = self.weight.data.contiguous(memory_format=_CL)
        self.p, self.eps, self.slope = float(p), float(eps), float(slope)
        self._cache: dict = {}

    def extra_repr(self) -> str:
        return (f'{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, '
                f'padding={self.padding}, dropout2d={self.p}, eps={self.eps}, '
                f'leaky={self.slope}, tcgen05/tf32')

    def refresh_derived_weights(self) -> None:
        """Updates the cached derived weights in place after an optimiser step (graph mode)."""
        TF.refresh_dgrad_weight(self.weight, self._cache)
        TF.refresh_stem_weight(self.weight, self._cache)

    def _drop_noise(self, n: int, device: torch.device) -> Optional[Tensor]:
        """Uniform noise for the channel dropout (one launch, PyTorch's CUDA generator, so
        checkpoint recomputation and graph replays reproduce it); the keep / scale decision is
        made inside the normalisation's prepare kernel."""
        if not self.training or self.p == 0.0:
            return None
        return torch.rand(n, self.out_channels, device=device)

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if not (input.is_cuda and input.dtype == torch.float32):
            y = F.conv2d(input, self.weight, None, self.stride, self.padding)
            y = F.dropout2d(y, self.p, self.training)
            return F.leaky_relu(F.instance_norm(y, eps=self.eps), self.slope)

        noise = self._drop_noise(input.size(0), input.device)
        got = _conv_any(input, self.weight, self.stride, self.padding, 1, self._cache)
        if got is None:         # no native path (odd output channels): library convolution
            x = input.contiguous(memory_format=_CL)
            got = F.conv2d(x, self.weight, None, self.stride, self.padding), None
        y, stats = got
        return TF.instnorm_leaky(y, stats, noise, self.eps, self.slope,
                                 drop_p=self.p if noise is not None else 0.0)


class AbsorbedBatchNorm2d(nn.BatchNorm2d):
    """A ``BatchNorm2d`` whose arithmetic has moved into the preceding :class:`ConvBatchNormAct`.

    The instance keeps owning its parameters and running statistics (so ``state_dict`` keys do
    not change) but its ``forward`` is the identity. ``deferred_chunks > 0`` switches the owner
    to DeferredBatchNorm semantics: statistics are accumulated over that many micro-batches
    (buffers ``sum`` / ``sum_squares``, as in reference ``torchgpipe/batchnorm.py:33-34``) and the
    running estimates are committed once per mini-batch.
    """

    deferred_chunks = 0
    moments: Optional[MiniBatchMoments] = None

    @property
    def tracked(self) -> int:
        """Micro-batches of the current mini-batch accumulated so far (DeferredBatchNorm's name)."""
        return self.moments.seen if self.moments is not None else 0

    @property
    def counter(self) -> int:
        """Elements per channel accumulated so far."""
        return self.moments.elements if self.moments is not None else 0

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        return input

    def extra_repr(self) -> str:
        extra = super().extra_repr() + ', absorbed'
        if self.deferred_chunks:
            extra += f', deferred over {self.deferred_chunks} micro-batches'
        return extra

    def make_deferred(self, chunks: int) -> None:
        if 'sum' not in self._buffers:
            self.register_buffer('sum', torch.zeros_like(self.running_mean))
            self.register_buffer('sum_squares', torch.zeros_like(self.running_var))
        self.deferred_chunks = chunks
        self.moments = MiniBatchMoments(self, chunks)

    def track(self, stats: Tensor, count: int) -> None:
        """Adds one micro-batch's (sum, sum of squares) as produced by the convolution epilogue
        (``stats[:, 0]``, ``stats[:, 1]``); the running estimates move once per mini-batch."""
        assert self.moments is not None
        self.moments.absorb(stats[:, 0], stats[:, 1], count)


def _bn_running_args(bn: AbsorbedBatchNorm2d, stats: Tensor, count: int):
    """What the fused kernels need to keep ``bn``'s running estimates: ``(running_mean,
    running_var, momentum, num_batches_tracked)`` - ``(None, None, 0, None)`` while recomputing,
    without tracking, or under DeferredBatchNorm (whose accumulation happens here as a side
    effect)."""
    if bn.track_running_stats and not is_recomputing():
        if bn.deferred_chunks:
            bn.track(stats, count)
        else:
            # the prepare kernel increments num_batches_tracked itself; momentum None (the
            # cumulative average) is passed as -1 and resolved on the device from that counter
            momentum = -1.0 if bn.momentum is None else float(bn.momentum)
            return bn.running_mean, bn.running_var, momentum, bn.num_batches_tracked
    return None, None, 0.0, None


def _bn_torch(bn: AbsorbedBatchNorm2d, y: Tensor) -> Tensor:
    """ATen path (CPU tensors, eval mode, other dtypes) with the same bookkeeping."""
    if bn.training and (is_recomputing() or bn.deferred_chunks):
        if bn.deferred_chunks and bn.track_running_stats and not is_recomputing():
            with torch.no_grad():
                stats = torch.stack([y.sum((0, 2, 3)), (y * y).sum((0, 2, 3))], dim=1)
            bn.track(stats, y.numel() // y.size(1))
        return F.batch_norm(y, None, None, bn.weight, bn.bias, True, 0.0, bn.eps)
    return nn.BatchNorm2d.forward(bn, y)


class ConvBatchNormAct(nn.Module):
    """``Conv2d(bias=False) -> BatchNorm2d [-> ReLU]`` with the statistics taken in the
    convolution's epilogue and one normalise(+ReLU) pass (reference: separate cuDNN conv, BN and
    ATen ReLU kernels in ``benchmarks/models/resnet`` / ``amoebanet``).

    The BatchNorm layer stays in the model as an :class:`AbsorbedBatchNorm2d` that owns the affine
    parameters and running statistics; this module only *uses* them.
    """

    def __init__(self, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool,
                 defer: bool = False) -> None:
        super().__init__()
        # defer=True: output (raw convolution, statistics) and leave the normalisation to a
        # following BatchNormResidualAct, which fuses it with the shortcut add and the ReLU
        self.defer = bool(defer)
        self.in_channels, self.out_channels = conv.in_channels, conv.out_channels
        self.kernel_size, self.padding, self.stride = conv.kernel_size, conv.padding, conv.stride
        self.weight = conv.weight
        self.weight.data = self.weight.data.contiguous(memory_format=_CL)
        self.relu = bool(relu)
        self._bn = (bn,)            # a tuple hides it from nn.Module registration
        self._cache: dict = {}

    @property
    def bn(self) -> AbsorbedBatchNorm2d:
        return self._bn[0]  # type: ignore[return-value]

    def refresh_derived_weights(self) -> None:
        TF.refresh_dgrad_weight(self.weight, self._cache)
        TF.refresh_stem_weight(self.weight, self._cache)

    def extra_repr(self) -> str:
        return (f'{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, '
                f'stride={self.stride}, padding={self.padding}, bn(eps={self.bn.eps}), '
                f'relu={self.relu}, defer={self.defer}, tcgen05/tf32')

    def fusion_partners(self):
        # a deferring layer only hands (convolution, statistics) on: it may sit in another stage
        return [] if self.defer else [self.bn]

    def forward(self, input: Tensor, add: Optional[Tensor] = None):  # type: ignore[override]
        """``add`` (optional, not with ``defer``): a tensor of the output's shape that is added
        after the normalisation and before the ReLU, inside the normalisation kernel."""
        bn = self.bn
        native = (input.is_cuda and input.dtype == torch.float32
                  and self.out_channels % 4 == 0)
        training = self.training if self.defer else bn.training
        if not native or not training:
            y = F.conv2d(input, self.weight, None, self.stride, self.padding)
            if self.defer:
                return y, y.new_empty(0)
            y = _bn_torch(bn, y)
            if add is not None:
                y = y + add
            return F.relu(y) if self.relu else y

        got = _conv_any(input, self.weight, self.stride, self.padding, 2, self._cache)
        if got is None:         # no native path: library kernel, statistics by tgb_moments
            x = input.contiguous(memory_format=_CL)
            got = F.conv2d(x, self.weight, None, self.stride, self.padding), None
        y, stats = got
        if stats is None:
            stats = TF.channel_moments(y)
        if self.defer:
            return y, stats

        running_mean, running_var, momentum, tracked = _bn_running_args(bn, stats,
                                                                        y.numel() // y.size(1))
        if add is not None:
            # relu?(BatchNorm(conv) + add) in the one normalisation pass (a sum of two branches)
            return TF.batchnorm_add_act(y, stats, bn.weight, bn.bias, add, running_mean,
                                        running_var, bn.eps, momentum, self.relu, tracked)
        return TF.batchnorm_act(y, stats, bn.weight, bn.bias, running_mean, running_var, bn.eps,
                                momentum, self.relu, tracked)


def _bn_apply(bn: AbsorbedBatchNorm2d, y: Tensor, stats: Optional[Tensor], relu: bool) -> Tensor:
    """Training-mode BatchNorm(+ReLU) of ``y`` from epilogue statistics, with ``bn``'s running
    estimates (or DeferredBatchNorm accumulators) kept up to date."""
    if stats is None:
        stats = TF.channel_moments(y)
    running_mean, running_var, momentum, tracked = _bn_running_args(bn, stats,
                                                                    y.numel() // y.size(1))
    return TF.batchnorm_act(y, stats, bn.weight, bn.bias, running_mean, running_var, bn.eps,
                            momentum, relu, tracked)


class AmoebaStem(nn.Module):
    """AmoebaNet's ``Stem`` (``relu -> conv3x3/2 (3 channels in) -> bn``, reference
    ``benchmarks/models/amoebanet/__init__.py:47-58``) on the packed-window stem kernels: the
    ReLU of the image is folded into the packing pass, the BatchNorm statistics come from the
    convolution's epilogue. Children keep their names (``relu``, ``conv``, ``bn``)."""

    def __init__(self, stem: nn.Module) -> None:
        super().__init__()
        self.relu, self.conv, self.bn = stem.relu, stem.conv, stem.bn
        self.bn.__class__ = AbsorbedBatchNorm2d
        self._cache: dict = {}

    def refresh_derived_weights(self) -> None:
        TF.refresh_stem_weight(self.conv.weight, self._cache)

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        conv, bn = self.conv, self.bn
        native = (input.is_cuda and input.dtype == torch.float32 and bn.training
                  and TF.conv_kind(conv.in_channels, conv.out_channels, conv.kernel_size,
                                   conv.stride, conv.dilation, conv.groups) == 'stem')
        if not native:
            y = F.conv2d(F.relu(input), conv.weight, None, conv.stride, conv.padding)
            return _bn_torch(bn, y)
        y, stats = TF.stem_conv2d(input, conv.weight, conv.padding, conv.stride, stats_mode=2,
                                  cache=self._cache, relu_input=True)
        return _bn_apply(bn, y, stats, relu=False)


class AmoebaFactorizedReduce(nn.Module):
    """AmoebaNet's ``FactorizedReduce`` (reference ``amoebanet/operations.py:24-40``): ReLU, two
    offset 1x1 stride-2 convolutions writing the two channel halves of one tensor (no shifted
    copy, no concatenation), BatchNorm from the epilogue statistics. Children keep their names
    (``relu``, ``pad``, ``conv1``, ``conv2``, ``bn``)."""

    def __init__(self, reduce: nn.Module) -> None:
        super().__init__()
        self.relu, self.pad = reduce.relu, reduce.pad
        self.conv1, self.conv2, self.bn = reduce.conv1, reduce.conv2, reduce.bn
        self.bn.__class__ = AbsorbedBatchNorm2d
        for conv in (self.conv1, self.conv2):
            conv.weight.data = conv.weight.data.contiguous(memory_format=_CL)
        self._cache1: dict = {}
        self._cache2: dict = {}

    def refresh_derived_weights(self) -> None:
        TF.refresh_dgrad_weight(self.conv1.weight, self._cache1)
        TF.refresh_dgrad_weight(self.conv2.weight, self._cache2)

    def forward(self, input: Tensor, relu_done: bool = False) -> Tensor:  # type: ignore[override]
        bn = self.bn
        x = input if relu_done else F.relu(input)
        c_in, k1, k2 = self.conv1.in_channels, self.conv1.out_channels, self.conv2.out_channels
        native = (x.is_cuda and x.dtype == torch.float32 and bn.training
                  and c_in % 4 == 0 and k1 % 4 == 0 and k2 % 4 == 0)
        if not native:
            shifted = self.pad(x[:, :, 1:, 1:])
            y = torch.cat([F.conv2d(x, self.conv1.weight, None, 2),
                           F.conv2d(shifted, self.conv2.weight, None, 2)], dim=1)
            return _bn_torch(bn, y)
        y, stats = TF.factorized_reduce_conv(x, self.conv1.weight, self.conv2.weight,
                                             self._cache1, self._cache2)
        return _bn_apply(bn, y, stats, relu=False)


class AmoebaCell(nn.Module):
    """An AmoebaNet-D ``Cell`` (reference ``benchmarks/models/amoebanet/__init__.py:97-136``) with
    the work between its operations fused:

    * every genotype operation starts with a ReLU of one of the cell's states, and several
      operations read the same state - the ReLU of a state is computed once and shared
      (the reference applies it once per operation: ``__init__.py:27``, 275 ReLUs in (18,256));
    * ``ops[k](a) + ops[k+1](b)`` (``__init__.py:132``): when one of the two branches ends in
      convolution + BatchNorm, the sum is folded into that BatchNorm's normalisation kernel
      (``relu?(bn(y) + other)`` in one pass) instead of a separate elementwise add.

    Children, their names and parameters are those of the original cell (``reduce1``,
    ``reduce2``, ``operations``), so ``state_dict`` keys and layer counts do not change. The
    final ``torch.cat`` of the chosen states stays (one copy per cell).
    """

    def __init__(self, cell: nn.Module) -> None:
        super().__init__()
        self.reduce1, self.reduce2, self.operations = cell.reduce1, cell.reduce2, cell.operations
        self.indices, self.concat = tuple(cell.indices), list(cell.concat)

    def extra_repr(self) -> str:
        return f'indices: {self.indices}, fused'

    # an operation's body as (needs_relu_input, callable(relu'd or raw input, add) -> output)
    @staticmethod
    def _leading_relu(body: nn.Module) -> bool:
        return (isinstance(body, AmoebaFactorizedReduce)
                or (type(body) is nn.Sequential and len(body) > 0 and type(body[0]) is nn.ReLU))

    @staticmethod
    def _can_join(body: nn.Module) -> bool:
        """Ends in a (non-deferring) convolution + BatchNorm without ReLU: a sum can be folded in."""
        if type(body) is not nn.Sequential:
            return False
        convs = [m for m in body if isinstance(m, ConvBatchNormAct)]
        return bool(convs) and not convs[-1].relu and not convs[-1].defer \
            and isinstance(body[len(body) - 1], AbsorbedBatchNorm2d) \
            and body[len(body) - 2] is convs[-1]

    @staticmethod
    def _run(body: nn.Module, x: Tensor, relu_done: bool, add: Optional[Tensor]) -> Tensor:
        if isinstance(body, AmoebaFactorizedReduce):
            out = body(x, relu_done=relu_done)
            return out if add is None else out + add
        if type(body) is nn.Sequential:
            layers = list(body)
            if relu_done and layers and type(layers[0]) is nn.ReLU:
                layers = layers[1:]
            last_conv = None
            if add is not None:
                last_conv = len(layers) - 2         # [..., ConvBatchNormAct, AbsorbedBatchNorm2d]
            for i, layer in enumerate(layers):
                x = layer(x, add=add) if i == last_conv else layer(x)
            return x
        out = body(x)
        return out if add is None else out + add

    def forward(self, input_or_states):  # type: ignore[override]
        if isinstance(input_or_states, tuple):
            s1, s2 = input_or_states
        else:
            s1 = s2 = input_or_states
        skip = s1
        relu_of: dict = {}

        def relu(t: Tensor) -> Tensor:
            got = relu_of.get(id(t))
            if got is None:
                got = relu_of[id(t)] = (t, F.relu(t))       # keep t alive: ids are reused
            return got[1]

        def feed(body: nn.Module, t: Tensor):
            return (relu(t), True) if self._leading_relu(body) else (t, False)

        def unwrap(op: nn.Module) -> nn.Module:
            return op.module if hasattr(op, 'module') and hasattr(op, 'name') else op

        x1, d1 = feed(self.reduce1, s1)
        x2, d2 = feed(self.reduce2, s2)
        states = [self._run(self.reduce1, x1, d1, None), self._run(self.reduce2, x2, d2, None)]

        ops, idx = self.operations, self.indices
        for k in range(0, len(ops), 2):
            a, b = unwrap(ops[k]), unwrap(ops[k + 1])
            ia, ib = states[idx[k]], states[idx[k + 1]]
            if self._can_join(b):               # a first, folded into b's last BatchNorm
                xa, da = feed(a, ia)
                out_a = self._run(a, xa, da, None)
                xb, db = feed(b, ib)
                states.append(self._run(b, xb, db, out_a))
            elif self._can_join(a):
                xb, db = feed(b, ib)
                out_b = self._run(b, xb, db, None)
                xa, da = feed(a, ia)
                states.append(self._run(a, xa, da, out_b))
            else:
                xa, da = feed(a, ia)
                xb, db = feed(b, ib)
                states.append(self._run(a, xa, da, None) + self._run(b, xb, db, None))
        return torch.cat([states[i] for i in self.concat], dim=1), skip


class MaxPool2x2(nn.Module):
    """``nn.MaxPool2d(2, stride=2)`` on the NHWC kernel (index-free backward)."""

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if (input.is_cuda and input.dtype == torch.float32 and input.dim() == 4
                and input.size(1) % 4 == 0):
            return TF.maxpool2x2(input)
        return F.max_pool2d(input, 2, 2)

    def extra_repr(self) -> str:
        return 'kernel_size=2, stride=2, nhwc'


def _make_upsample_popcat():
    from torchgpipe_b200.skip import pop, skippable

    @skippable(pop=['skip'])
    class UpsamplePopCat(nn.Module):
        """Nearest x2 upsample of the input, pad to the popped skip tensor's size and concat
        (input channels first) - ``nn.Upsample(scale_factor=2)`` followed by the reference's
        ``PopCat`` - as one kernel. The preceding ``nn.Upsample`` layer becomes a placeholder."""

        def forward(self, input: Tensor):  # type: ignore[override]
            skipped = yield pop('skip')
            if (input.is_cuda and input.dtype == torch.float32 and input.size(1) % 4 == 0
                    and skipped.size(1) % 4 == 0):
                return TF.upsample2x_cat(input, skipped)
            up = F.interpolate(input, scale_factor=2.0, mode='nearest')
            dh, dw = skipped.size(2) - up.size(2), skipped.size(3) - up.size(3)
            if dh or dw:
                up = F.pad(up, [0, dw, 0, dh])
            return torch.cat((up, skipped), dim=1)

    return UpsamplePopCat


UpsamplePopCat = _make_upsample_popcat()


class Pool2dNHWC(nn.Module):
    """``nn.MaxPool2d`` / ``nn.AvgPool2d`` (square window, floor mode) on the NHWC kernels with an
    index-free, deterministic backward. ``mode``: ``'max'``, ``'avg'`` (count_include_pad=False)
    or ``'avg_include_pad'``."""

    def __init__(self, kernel_size: int, stride: int, padding: int, mode: str) -> None:
        super().__init__()
        self.kernel_size, self.stride, self.padding, self.mode = kernel_size, stride, padding, mode

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if TF.pool_supported(input, self.kernel_size, self.padding):
            return TF.pool2d(input, self.kernel_size, self.stride, self.padding, self.mode)
        if self.mode == 'max':
            return F.max_pool2d(input, self.kernel_size, self.stride, self.padding)
        return F.avg_pool2d(input, self.kernel_size, self.stride, self.padding,
                            count_include_pad=self.mode == 'avg_include_pad')

    def extra_repr(self) -> str:
        return (f'kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}, '
                f'mode={self.mode}, nhwc')


class GlobalAvgPoolNHWC(nn.Module):
    """``nn.AdaptiveAvgPool2d(1)`` on the NHWC kernel."""

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if (input.is_cuda and input.dtype == torch.float32 and input.dim() == 4
                and input.size(1) % 4 == 0):
            return TF.global_avgpool(input)
        return F.adaptive_avg_pool2d(input, 1)


def _make_residual():
    from torchgpipe_b200.skip import pop, skippable

    @skippable(pop=['identity'])
    class BatchNormResidualAct(nn.Module):
        """``bn3 -> residual (pop shortcut, optional projection, add) -> relu3`` of a bottleneck
        (reference ``benchmarks/models/resnet/bottleneck.py:31-79``) as one kernel each way.

        Sits at the position of the reference's ``Residual`` layer (same skip route, same
        ``downsample`` child and keys). Its input is the ``(raw convolution, statistics)`` pair a
        deferring :class:`ConvBatchNormAct` produced; the BatchNorm layer in between is an
        :class:`AbsorbedBatchNorm2d` whose parameters this layer uses. Without a BatchNorm partner
        (``bn=None``) it is the plain fused ``relu?(input + shortcut)``.
        """

        def __init__(self, bn: Optional[nn.BatchNorm2d], downsample: Optional[nn.Module],
                     relu: bool) -> None:
            super().__init__()
            self._bn = (bn,)
            self.downsample = downsample
            self.relu = bool(relu)

        @property
        def bn(self) -> Optional[AbsorbedBatchNorm2d]:
            return self._bn[0]  # type: ignore[return-value]

        def fusion_partners(self):
            return [] if self.bn is None else [self.bn]

        def extra_repr(self) -> str:
            return f'bn={"absorbed" if self.bn is not None else None}, relu={self.relu}'

        def forward(self, input):  # type: ignore[override]
            identity = yield pop('identity')
            if self.downsample is not None:
                identity = self.downsample(identity)
            bn = self.bn
            if bn is None:
                y = input
                if (y.is_cuda and y.dtype == torch.float32 and y.dim() == 4
                        and y.size(1) % 4 == 0):
                    return TF.add_act(y, identity, self.relu)
                y = y + identity
                return F.relu(y) if self.relu else y
            y, stats = input
            if stats.numel() == 0 or not bn.training:       # ATen path (CPU / eval)
                z = _bn_torch(bn, y) + identity
                return F.relu(z) if self.relu else z
            running_mean, running_var, momentum, tracked = _bn_running_args(
                bn, stats, y.numel() // y.size(1))
            return TF.batchnorm_add_act(y, stats, bn.weight, bn.bias, identity, running_mean,
                                        running_var, bn.eps, momentum, self.relu, tracked)

    return BatchNormResidualAct


BatchNormResidualAct = _make_residual()


class LinearTC(nn.Module):
    """``nn.Linear`` on the tcgen05 kernels (same ``weight`` / ``bias`` parameters and keys)."""

    def __init__(self, linear: nn.Linear) -> None:
        super().__init__()
        self.in_features, self.out_features = linear.in_features, linear.out_features
        self.weight, self.bias = linear.weight, linear.bias
        self._cache: dict = {}

    def refresh_derived_weights(self) -> None:
        w4 = self.weight.view(self.out_features, self.in_features, 1, 1)
        TF.refresh_dgrad_weight(w4, self._cache)

    def extra_repr(self) -> str:
        return (f'in_features={self.in_features}, out_features={self.out_features}, '
                f'bias={self.bias is not None}, tcgen05/tf32')

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        if input.is_cuda and input.dtype == torch.float32:
            return TF.linear(input, self.weight, self.bias, cache=self._cache)
        return F.linear(input, self.weight, self.bias)
