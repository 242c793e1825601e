"""Autograd functions over the native sm_100a kernels (NHWC / channels-last, fp32 + TF32 MMA).

Every function here takes and returns ordinary ``torch`` tensors of logical shape ``[N,C,H,W]``
in ``torch.channels_last`` memory format (weights ``[K,C,R,S]`` channels-last = KRSC memory).
The kernels are enqueued on the current CUDA stream of the input's device, so they compose with
the pipeline's stream discipline like any ATen op.
"""
import ctypes
from typing import Optional, Tuple

import torch
from torch import Tensor

from torchgpipe_b200.ops import _lib

__all__ = ['linear', 'conv2d', 'conv2d_forward', 'instnorm_leaky', 'channel_moments', 'conv_supported',
           'refresh_dgrad_weight', 'batchnorm_act', 'maxpool2x2', 'upsample2x_cat', 'pool2d',
           'pool_supported', 'global_avgpool', 'add_act', 'batchnorm_add_act', 'conv2d_plan']

_CL = torch.channels_last
_c_int = ctypes.c_int
_c_float = ctypes.c_float
_c_ll = ctypes.c_longlong


def _cl(t: Tensor) -> Tensor:
    return t if t.is_contiguous(memory_format=_CL) else t.contiguous(memory_format=_CL)


def conv_supported(c_in: int, c_out: int, stride, dilation, groups: int) -> bool:
    """Whether the tcgen05 kernels cover this convolution (else callers use ATen/cuDNN)."""
    stride = tuple(stride) if not isinstance(stride, int) else (stride, stride)
    dilation = tuple(dilation) if not isinstance(dilation, int) else (dilation, dilation)
    return (stride == (1, 1) and dilation == (1, 1) and groups == 1
            and c_in % 4 == 0 and c_out % 4 == 0)


def conv2d_forward(x: Tensor, w: Tensor, padding: Tuple[int, int], stats_mode: int = 0,
                   ) -> Tuple[Tensor, Optional[Tensor]]:
    """Raw forward: returns ``(y, stats)``; ``stats`` is ``[N,K,2]`` (mode 1), ``[K,2]`` (mode 2)
    or ``None``. No autograd."""
    lib = _lib.require()
    x, w = _cl(x), _cl(w)
    N, C, H, W = x.shape
    K, Cw, R, S = w.shape
    assert C == Cw, f'channel mismatch: input {C}, weight {Cw}'
    ph, pw = padding
    Ho, Wo = H + 2 * ph - R + 1, W + 2 * pw - S + 1
    y = torch.empty((N, K, Ho, Wo), device=x.device, dtype=torch.float32, memory_format=_CL)
    stats = None
    if stats_mode == 1:
        stats = torch.zeros((N, K, 2), device=x.device, dtype=torch.float32)
    elif stats_mode == 2:
        stats = torch.zeros((K, 2), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        rc = lib.tgb_conv2d_fprop_tf32(_lib.ptr(x), _lib.ptr(w), _lib.ptr(y), _lib.ptr(stats),
                                       _c_int(N), _c_int(H), _c_int(W), _c_int(C), _c_int(K),
                                       _c_int(R), _c_int(S), _c_int(ph), _c_int(pw),
                                       _c_int(stats_mode), _lib.stream_ptr(x.device))
    _lib.check(rc, 'tgb_conv2d_fprop_tf32')
    return y, stats


def _conv2d_wgrad(dy: Tensor, x: Tensor, w_shape, padding: Tuple[int, int]) -> Tensor:
    lib = _lib.require()
    lib.tgb_conv2d_wgrad_scratch_bytes.restype = _c_ll
    N, C, H, W = x.shape
    K, _, R, S = w_shape
    ph, pw = padding
    args = (_c_int(N), _c_int(H), _c_int(W), _c_int(C), _c_int(K), _c_int(R), _c_int(S),
            _c_int(ph), _c_int(pw))
    need = lib.tgb_conv2d_wgrad_scratch_bytes(*args)
    if need < 0:
        _lib.check(1, 'tgb_conv2d_wgrad_scratch_bytes')
    scratch = torch.empty(need // 4, device=x.device, dtype=torch.float32) if need else None
    dw = torch.empty(w_shape, device=x.device, dtype=torch.float32, memory_format=_CL)
    with torch.cuda.device(x.device):
        rc = lib.tgb_conv2d_wgrad_tf32(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(dw), _lib.ptr(scratch),
                                       _c_ll(need), *args, _c_int(0),
                                       _lib.stream_ptr(x.device))
    _lib.check(rc, 'tgb_conv2d_wgrad_tf32')
    return dw


def _dgrad_weight(w: Tensor, cache: Optional[dict]) -> Tensor:
    """Tap-flipped, channel-transposed weights (KRSC memory) for the data-gradient pass.

    ``cache`` is a dict owned by the calling module: with several micro-batches per optimiser
    step the transform then runs once per parameter version instead of once per backward call.
    """
    if cache is not None:
        if cache.get('version') == w._version and cache.get('ptr') == w.data_ptr():
            return cache['wt']
    K, C, R, S = w.shape
    old = cache.get('wt') if cache is not None else None
    if old is not None and old.shape == (C, K, R, S) and old.device == w.device:
        wt = old                # keep the address: captured CUDA graphs read this buffer
    else:
        wt = torch.empty((C, K, R, S), device=w.device, dtype=torch.float32, memory_format=_CL)
    src = _cl(w.detach())
    with torch.cuda.device(w.device):
        rc = _lib.require().tgb_dgrad_weight(_lib.ptr(src), _lib.ptr(wt), _c_int(K), _c_int(C),
                                             _c_int(R), _c_int(S), _lib.stream_ptr(w.device))
    _lib.check(rc, 'tgb_dgrad_weight')
    if cache is not None:
        cache['version'], cache['ptr'], cache['wt'] = w._version, w.data_ptr(), wt
    return wt


def refresh_dgrad_weight(w: Tensor, cache: dict) -> None:
    """Brings a cached dgrad weight up to date in place (no-op if never built or current)."""
    if 'wt' in cache and (cache.get('version') != w._version or cache.get('ptr') != w.data_ptr()):
        _dgrad_weight(w, cache)


class _Conv2d(torch.autograd.Function):
    """Stride-1 convolution; backward = dgrad (same tcgen05 kernel, flipped weights) + wgrad."""

    @staticmethod
    def forward(ctx, x: Tensor, w: Tensor, ph: int, pw: int, stats_mode: int,  # type: ignore[override]
                cache: Optional[dict]):
        x, w = _cl(x), _cl(w)
        y, stats = conv2d_forward(x, w, (ph, pw), stats_mode)
        ctx.save_for_backward(x, w)
        ctx.padding = (ph, pw)
        ctx.cache = cache
        if stats is None:
            stats = x.new_empty(0)
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, dy: Tensor, _dstats):  # type: ignore[override]
        x, w = ctx.saved_tensors
        ph, pw = ctx.padding
        K, C, R, S = w.shape
        dy = _cl(dy)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if conv_supported(K, C, 1, 1, 1):
                # dX = conv(dY, W^T flipped), pad' = R-1-pad
                dx, _ = conv2d_forward(dy, _dgrad_weight(w, ctx.cache), (R - 1 - ph, S - 1 - pw), 0)
            else:
                dx = torch.nn.grad.conv2d_input(x.shape, w, dy, padding=(ph, pw))
        if ctx.needs_input_grad[1]:
            dw = _conv2d_wgrad(dy, x, w.shape, (ph, pw))
        return dx, dw, None, None, None, None


def conv2d(x: Tensor, w: Tensor, padding: Tuple[int, int], stats_mode: int = 0,
           cache: Optional[dict] = None) -> Tuple[Tensor, Optional[Tensor]]:
    """Differentiable stride-1 convolution on tcgen05. Returns ``(y, stats or None)``.

    ``cache``: optional dict owned by the caller, reused across calls to keep derived weights.
    """
    y, stats = _Conv2d.apply(x, w, int(padding[0]), int(padding[1]), int(stats_mode), cache)
    return y, (stats if stats.numel() else None)


def channel_moments(x: Tensor, per_image: bool = False) -> Tensor:
    """(sum, sum of squares) per channel (``[C,2]``) or per (image, channel) (``[N,C,2]``)."""
    lib = _lib.require()
    x = _cl(x)
    N, C, H, W = x.shape
    G, P = (N, H * W) if per_image else (1, N * H * W)
    out = torch.zeros((G, C, 2), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        rc = lib.tgb_moments(_lib.ptr(x), _lib.ptr(out), _c_int(G), _c_int(P), _c_int(C),
                             _lib.stream_ptr(x.device))
    _lib.check(rc, 'tgb_moments')
    return out if per_image else out[0]


class _InstNormLeaky(torch.autograd.Function):
    """z = LeakyReLU(InstanceNorm(Dropout2d(y))) with the dropout folded into the norm.

    Saves only the output ``z`` and the per-(n,c) ``(mean, coef)`` pair: the pre-activation and
    the convolution output are re-derived from ``z`` in backward.
    """

    @staticmethod
    def forward(ctx, y: Tensor, stats: Optional[Tensor], drop_scale: Optional[Tensor],  # type: ignore[override]
                eps: float, slope: float):
        lib = _lib.require()
        y = _cl(y)
        N, C, H, W = y.shape
        P = H * W
        if stats is None or stats.numel() == 0:
            stats = channel_moments(y, per_image=True)
        mean_coef = torch.empty((N, C, 2), device=y.device, dtype=torch.float32)
        z = torch.empty_like(y, memory_format=_CL)
        if drop_scale is not None:
            drop_scale = drop_scale.reshape(N, C).contiguous()
        with torch.cuda.device(y.device):
            st = _lib.stream_ptr(y.device)
            rc = lib.tgb_in_prepare(_lib.ptr(stats), _lib.ptr(drop_scale), _lib.ptr(mean_coef),
                                    _c_int(N), _c_int(P), _c_int(C), _c_float(eps), st)
            _lib.check(rc, 'tgb_in_prepare')
            rc = lib.tgb_in_leaky_fwd(_lib.ptr(y), _lib.ptr(mean_coef), _lib.ptr(z), _c_int(N),
                                      _c_int(P), _c_int(C), _c_float(slope), st)
            _lib.check(rc, 'tgb_in_leaky_fwd')
        ctx.save_for_backward(z, mean_coef)
        ctx.slope = slope
        return z

    @staticmethod
    def backward(ctx, dz: Tensor):  # type: ignore[override]
        lib = _lib.require()
        z, mean_coef = ctx.saved_tensors
        dz = _cl(dz)
        N, C, H, W = z.shape
        P = H * W
        sums = torch.zeros((N, C, 2), device=z.device, dtype=torch.float32)
        dy = torch.empty_like(z, memory_format=_CL)
        with torch.cuda.device(z.device):
            st = _lib.stream_ptr(z.device)
            rc = lib.tgb_in_leaky_bwd_sums(_lib.ptr(dz), _lib.ptr(z), _lib.ptr(sums), _c_int(N),
                                           _c_int(P), _c_int(C), _c_float(ctx.slope), st)
            _lib.check(rc, 'tgb_in_leaky_bwd_sums')
            rc = lib.tgb_in_leaky_bwd(_lib.ptr(dz), _lib.ptr(z), _lib.ptr(mean_coef),
                                      _lib.ptr(sums), _lib.ptr(dy), _c_int(N), _c_int(P),
                                      _c_int(C), _c_float(ctx.slope), st)
            _lib.check(rc, 'tgb_in_leaky_bwd')
        return dy, None, None, None, None


def instnorm_leaky(y: Tensor, stats: Optional[Tensor] = None, drop_scale: Optional[Tensor] = None,
                   eps: float = 1e-5, slope: float = 0.01) -> Tensor:
    """Fused ``LeakyReLU(InstanceNorm2d(Dropout2d(y)))`` (no affine, batch statistics).

    ``stats`` are the per-(n,c) (sum, sum of squares) from the convolution epilogue (computed
    here if absent); ``drop_scale`` is ``[N,C]`` with entries 0 or ``1/(1-p)`` (``None`` = no
    dropout). ``slope`` must be positive.
    """
    return _InstNormLeaky.apply(y, stats, drop_scale, float(eps), float(slope))


class _BatchNormAct(torch.autograd.Function):
    """z = act(BatchNorm(y)) in training mode from precomputed per-channel (sum, sum of squares).

    Saves the convolution output ``y`` and two ``[C,2]`` tables; the ReLU mask and the
    normalised activation are re-derived from ``y`` in backward.
    """

    @staticmethod
    def forward(ctx, y: Tensor, stats: Tensor, weight: Optional[Tensor],  # type: ignore[override]
                bias: Optional[Tensor], running_mean: Optional[Tensor],
                running_var: Optional[Tensor], eps: float, momentum: float, relu: bool):
        lib = _lib.require()
        y = _cl(y)
        N, C, H, W = y.shape
        P = N * H * W
        coef = torch.empty((C, 2), device=y.device, dtype=torch.float32)
        saved = torch.empty((C, 2), device=y.device, dtype=torch.float32)
        z = torch.empty_like(y, memory_format=_CL)
        with torch.cuda.device(y.device):
            st = _lib.stream_ptr(y.device)
            rc = lib.tgb_bn_prepare(_lib.ptr(stats), _lib.ptr(weight), _lib.ptr(bias),
                                    _lib.ptr(coef), _lib.ptr(saved), _lib.ptr(running_mean),
                                    _lib.ptr(running_var), _c_int(P), _c_int(C), _c_float(eps),
                                    _c_float(momentum), st)
            _lib.check(rc, 'tgb_bn_prepare')
            rc = lib.tgb_bn_act_fwd(_lib.ptr(y), _lib.ptr(coef), _lib.ptr(z), _c_int(P),
                                    _c_int(C), _c_int(int(relu)), st)
            _lib.check(rc, 'tgb_bn_act_fwd')
        ctx.save_for_backward(y, coef, saved)
        ctx.relu = relu
        ctx.affine = (weight is not None, bias is not None)
        return z

    @staticmethod
    def backward(ctx, dz: Tensor):  # type: ignore[override]
        lib = _lib.require()
        y, coef, saved = ctx.saved_tensors
        dz = _cl(dz)
        N, C, H, W = y.shape
        P = N * H * W
        sums = torch.zeros((C, 2), device=y.device, dtype=torch.float32)
        dy = torch.empty_like(y, memory_format=_CL)
        with torch.cuda.device(y.device):
            rc = lib.tgb_bn_act_bwd(_lib.ptr(dz), _lib.ptr(y), _lib.ptr(coef), _lib.ptr(saved),
                                    _lib.ptr(sums), _lib.ptr(dy), _c_int(P), _c_int(C),
                                    _c_int(int(ctx.relu)), _lib.stream_ptr(y.device))
        _lib.check(rc, 'tgb_bn_act_bwd')
        dweight = sums[:, 1].clone() if ctx.affine[0] else None
        dbias = sums[:, 0].clone() if ctx.affine[1] else None
        return dy, None, dweight, dbias, None, None, None, None, None


def batchnorm_act(y: Tensor, stats: Optional[Tensor], weight: Optional[Tensor],
                  bias: Optional[Tensor], running_mean: Optional[Tensor] = None,
                  running_var: Optional[Tensor] = None, eps: float = 1e-5,
                  momentum: float = 0.1, relu: bool = False) -> Tensor:
    """Fused training-mode ``ReLU?(BatchNorm2d(y))``.

    ``stats``: per-channel ``[C,2]`` (sum, sum of squares) from the convolution epilogue
    (``stats_mode=2``); computed here when ``None``. When ``running_mean``/``running_var`` are
    given they are updated in place like ``nn.BatchNorm2d`` does (pass ``None`` to skip, e.g.
    during checkpoint recomputation or under DeferredBatchNorm).
    """
    if stats is None or stats.numel() == 0:
        stats = channel_moments(y)
    return _BatchNormAct.apply(y, stats, weight, bias, running_mean, running_var, float(eps),
                               float(momentum), bool(relu))


class _MaxPool2x2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor):  # type: ignore[override]
        lib = _lib.require()
        x = _cl(x)
        N, C, H, W = x.shape
        y = torch.empty((N, C, H // 2, W // 2), device=x.device, dtype=torch.float32,
                        memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_maxpool2x2_fwd(_lib.ptr(x), _lib.ptr(y), _c_int(N), _c_int(H), _c_int(W),
                                        _c_int(C), _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_maxpool2x2_fwd')
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        lib = _lib.require()
        x, = ctx.saved_tensors
        dy = _cl(dy)
        N, C, H, W = x.shape
        dx = torch.empty_like(x, memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_maxpool2x2_bwd(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(dx), _c_int(N),
                                        _c_int(H), _c_int(W), _c_int(C),
                                        _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_maxpool2x2_bwd')
        return dx


def maxpool2x2(x: Tensor) -> Tensor:
    """``MaxPool2d(2, stride=2)`` on a channels-last fp32 CUDA tensor (C % 4 == 0)."""
    return _MaxPool2x2.apply(x)


class _UpsampleCat(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor, skip: Tensor):  # type: ignore[override]
        lib = _lib.require()
        x, skip = _cl(x), _cl(skip)
        N, C1, Hx, Wx = x.shape
        _, C2, Hs, Ws = skip.shape
        out = torch.empty((N, C1 + C2, Hs, Ws), device=x.device, dtype=torch.float32,
                          memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_upsample2x_cat_fwd(_lib.ptr(x), _lib.ptr(skip), _lib.ptr(out), _c_int(N),
                                            _c_int(Hx), _c_int(Wx), _c_int(C1), _c_int(Hs),
                                            _c_int(Ws), _c_int(C2), _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_upsample2x_cat_fwd')
        ctx.dims = (N, Hx, Wx, C1, Hs, Ws, C2)
        return out

    @staticmethod
    def backward(ctx, dout: Tensor):  # type: ignore[override]
        lib = _lib.require()
        N, Hx, Wx, C1, Hs, Ws, C2 = ctx.dims
        dout = _cl(dout)
        dx = torch.empty((N, C1, Hx, Wx), device=dout.device, dtype=torch.float32,
                         memory_format=_CL)
        dskip = torch.empty((N, C2, Hs, Ws), device=dout.device, dtype=torch.float32,
                            memory_format=_CL)
        with torch.cuda.device(dout.device):
            rc = lib.tgb_upsample2x_cat_bwd(_lib.ptr(dout), _lib.ptr(dx), _lib.ptr(dskip),
                                            _c_int(N), _c_int(Hx), _c_int(Wx), _c_int(C1),
                                            _c_int(Hs), _c_int(Ws), _c_int(C2),
                                            _lib.stream_ptr(dout.device))
        _lib.check(rc, 'tgb_upsample2x_cat_bwd')
        return dx, dskip


def upsample2x_cat(x: Tensor, skip: Tensor) -> Tensor:
    """``cat(pad_to(upsample_nearest_x2(x), skip), skip, dim=1)`` in one pass (channels-last)."""
    return _UpsampleCat.apply(x, skip)


def linear(x: Tensor, weight: Tensor, bias: Optional[Tensor] = None,
           cache: Optional[dict] = None) -> Tensor:
    """``F.linear`` on the tcgen05 convolution kernels: an ``[N, in]`` matrix is an NHWC image
    batch with 1x1 pixels, the weight a 1x1 filter bank (forward, dgrad and wgrad all reuse the
    convolution kernels). Needs ``in % 4 == 0`` and ``out % 4 == 0``."""
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    y, _ = conv2d(x2.view(x2.size(0), x2.size(1), 1, 1), weight.view(*weight.shape, 1, 1),
                  (0, 0), cache=cache)
    y = y.view(x2.size(0), weight.size(0))
    if bias is not None:
        y = y + bias
    return y.view(*lead, weight.size(0))


# ---------------------------------------------------------------- pooling family / residual joins
_POOL_MODES = {'max': 0, 'avg': 1, 'avg_include_pad': 2}


class _Pool2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor, k: int, s: int, p: int, mode: int):  # type: ignore[override]
        lib = _lib.require()
        x = _cl(x)
        N, C, H, W = x.shape
        Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        y = torch.empty((N, C, Ho, Wo), device=x.device, dtype=torch.float32, memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_pool2d_fwd(_lib.ptr(x), _lib.ptr(y), _c_int(N), _c_int(H), _c_int(W),
                                    _c_int(C), _c_int(k), _c_int(s), _c_int(p), _c_int(mode),
                                    _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_pool2d_fwd')
        ctx.geom = (N, C, H, W, k, s, p, mode)
        if mode == 0:
            ctx.save_for_backward(x)        # average pooling needs no input in backward
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        lib = _lib.require()
        N, C, H, W, k, s, p, mode = ctx.geom
        x = ctx.saved_tensors[0] if mode == 0 else None
        dy = _cl(dy)
        dx = torch.empty((N, C, H, W), device=dy.device, dtype=torch.float32, memory_format=_CL)
        with torch.cuda.device(dy.device):
            rc = lib.tgb_pool2d_bwd(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(dx), _c_int(N), _c_int(H),
                                    _c_int(W), _c_int(C), _c_int(k), _c_int(s), _c_int(p),
                                    _c_int(mode), _lib.stream_ptr(dy.device))
        _lib.check(rc, 'tgb_pool2d_bwd')
        return dx, None, None, None, None


def pool2d(x: Tensor, kernel_size: int, stride: int, padding: int = 0, mode: str = 'max') -> Tensor:
    """``MaxPool2d`` / ``AvgPool2d`` (square window, floor output size) on a channels-last fp32
    CUDA tensor with ``C % 4 == 0``. ``mode``: ``'max'``, ``'avg'`` (count_include_pad=False) or
    ``'avg_include_pad'``."""
    return _Pool2d.apply(x, int(kernel_size), int(stride), int(padding), _POOL_MODES[mode])


def pool_supported(x: Tensor, kernel_size: int, padding: int) -> bool:
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.size(1) % 4 == 0
            and 1 <= kernel_size <= 7 and 2 * padding <= kernel_size)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor):  # type: ignore[override]
        lib = _lib.require()
        x = _cl(x)
        N, C, H, W = x.shape
        y = torch.empty((N, C, 1, 1), device=x.device, dtype=torch.float32, memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_global_avgpool_fwd(_lib.ptr(x), _lib.ptr(y), _c_int(N), _c_int(H * W),
                                            _c_int(C), _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_global_avgpool_fwd')
        ctx.geom = (N, C, H, W)
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        lib = _lib.require()
        N, C, H, W = ctx.geom
        dy = dy.reshape(N, C).contiguous()
        dx = torch.empty((N, C, H, W), device=dy.device, dtype=torch.float32, memory_format=_CL)
        with torch.cuda.device(dy.device):
            rc = lib.tgb_global_avgpool_bwd(_lib.ptr(dy), _lib.ptr(dx), _c_int(N), _c_int(H * W),
                                            _c_int(C), _lib.stream_ptr(dy.device))
        _lib.check(rc, 'tgb_global_avgpool_bwd')
        return dx


def global_avgpool(x: Tensor) -> Tensor:
    """``AdaptiveAvgPool2d(1)``: ``[N,C,H,W]`` (channels-last fp32, C % 4 == 0) -> ``[N,C,1,1]``."""
    return _GlobalAvgPool.apply(x)


class _AddAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a: Tensor, b: Tensor, relu: bool):  # type: ignore[override]
        lib = _lib.require()
        a, b = _cl(a), _cl(b)
        z = torch.empty_like(a, memory_format=_CL)
        with torch.cuda.device(a.device):
            rc = lib.tgb_add_act_fwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(z), _c_ll(a.numel()),
                                     _c_int(int(relu)), _lib.stream_ptr(a.device))
        _lib.check(rc, 'tgb_add_act_fwd')
        ctx.relu = relu
        if relu:
            ctx.save_for_backward(z)
        return z

    @staticmethod
    def backward(ctx, dz: Tensor):  # type: ignore[override]
        if not ctx.relu:
            return dz, dz, None
        lib = _lib.require()
        z, = ctx.saved_tensors
        dz = _cl(dz)
        d = torch.empty_like(z, memory_format=_CL)
        with torch.cuda.device(z.device):
            rc = lib.tgb_relu_mask(_lib.ptr(dz), _lib.ptr(z), _lib.ptr(d), _c_ll(z.numel()),
                                   _lib.stream_ptr(z.device))
        _lib.check(rc, 'tgb_relu_mask')
        return d, d, None


def add_act(a: Tensor, b: Tensor, relu: bool = True) -> Tensor:
    """``relu?(a + b)`` for two same-shape channels-last fp32 CUDA tensors (numel % 4 == 0); the
    backward is one masked copy shared by both summands."""
    return _AddAct.apply(a, b, bool(relu))


class _BatchNormAddAct(torch.autograd.Function):
    """z = relu?(BatchNorm(y) + res), training mode, statistics given as per-channel
    (sum, sum of squares). Saves ``y``, ``z`` (alive anyway as the next layer's input) and two
    ``[C,2]`` tables; the ReLU mask comes from ``z > 0``."""

    @staticmethod
    def forward(ctx, y: Tensor, stats: Tensor, weight: Optional[Tensor],  # type: ignore[override]
                bias: Optional[Tensor], res: Tensor, running_mean: Optional[Tensor],
                running_var: Optional[Tensor], eps: float, momentum: float, relu: bool):
        lib = _lib.require()
        y, res = _cl(y), _cl(res)
        N, C, H, W = y.shape
        P = N * H * W
        coef = torch.empty((C, 2), device=y.device, dtype=torch.float32)
        saved = torch.empty((C, 2), device=y.device, dtype=torch.float32)
        z = torch.empty_like(y, memory_format=_CL)
        with torch.cuda.device(y.device):
            st = _lib.stream_ptr(y.device)
            rc = lib.tgb_bn_prepare(_lib.ptr(stats), _lib.ptr(weight), _lib.ptr(bias),
                                    _lib.ptr(coef), _lib.ptr(saved), _lib.ptr(running_mean),
                                    _lib.ptr(running_var), _c_int(P), _c_int(C), _c_float(eps),
                                    _c_float(momentum), st)
            _lib.check(rc, 'tgb_bn_prepare')
            rc = lib.tgb_bn_add_act_fwd(_lib.ptr(y), _lib.ptr(coef), _lib.ptr(res), _lib.ptr(z),
                                        _c_int(P), _c_int(C), _c_int(int(relu)), st)
            _lib.check(rc, 'tgb_bn_add_act_fwd')
        ctx.save_for_backward(y, z, coef, saved)
        ctx.relu = relu
        ctx.affine = (weight is not None, bias is not None)
        return z

    @staticmethod
    def backward(ctx, dz: Tensor):  # type: ignore[override]
        lib = _lib.require()
        y, z, coef, saved = ctx.saved_tensors
        dz = _cl(dz)
        N, C, H, W = y.shape
        P = N * H * W
        sums = torch.zeros((C, 2), device=y.device, dtype=torch.float32)
        dy = torch.empty_like(y, memory_format=_CL)
        dres = torch.empty_like(y, memory_format=_CL)
        with torch.cuda.device(y.device):
            rc = lib.tgb_bn_add_act_bwd(_lib.ptr(dz), _lib.ptr(y), _lib.ptr(z), _lib.ptr(coef),
                                        _lib.ptr(saved), _lib.ptr(sums), _lib.ptr(dy),
                                        _lib.ptr(dres), _c_int(P), _c_int(C),
                                        _c_int(int(ctx.relu)), _lib.stream_ptr(y.device))
        _lib.check(rc, 'tgb_bn_add_act_bwd')
        dweight = sums[:, 1].clone() if ctx.affine[0] else None
        dbias = sums[:, 0].clone() if ctx.affine[1] else None
        return dy, None, dweight, dbias, dres, None, None, None, None, None


def batchnorm_add_act(y: Tensor, stats: Optional[Tensor], weight: Optional[Tensor],
                      bias: Optional[Tensor], res: Tensor,
                      running_mean: Optional[Tensor] = None,
                      running_var: Optional[Tensor] = None, eps: float = 1e-5,
                      momentum: float = 0.1, relu: bool = True) -> Tensor:
    """Fused training-mode ``ReLU?(BatchNorm2d(y) + res)`` (``bn3 -> residual -> relu3`` of a
    bottleneck, reference ``benchmarks/models/resnet/bottleneck.py:31-79``). Arguments as
    :func:`batchnorm_act`."""
    if stats is None or stats.numel() == 0:
        stats = channel_moments(y)
    return _BatchNormAddAct.apply(y, stats, weight, bias, res, running_mean, running_var,
                                  float(eps), float(momentum), bool(relu))


_PLAN_FIELDS = ('BN', 'BW', 'BH', 'BNB', 'tiles_w', 'tiles_h', 'tiles_n', 'tiles_k', 'splits',
                'iters_per_split', 'halo', 'resident_weights', 'stages', 'iters_per_tile', 'Ho',
                'Wo')


def conv2d_plan(N: int, C: int, H: int, W: int, K: int, R: int = 3, S: int = 3,
                padding: Tuple[int, int] = (1, 1), stats_mode: int = 0) -> dict:
    """Host-side decisions of the forward convolution for one problem (tile shape, kernel
    variant, split of the reduction) as a dict - no launch, no GPU needed. For tuning, for
    explaining a layer's speed, and pinned by the CPU test-suite."""
    out = (ctypes.c_int * 16)()
    rc = _lib.require().tgb_conv2d_fprop_plan(_c_int(N), _c_int(H), _c_int(W), _c_int(C), _c_int(K),
                                              _c_int(R), _c_int(S), _c_int(padding[0]),
                                              _c_int(padding[1]), _c_int(stats_mode), out)
    _lib.check(rc, 'tgb_conv2d_fprop_plan')
    plan = dict(zip(_PLAN_FIELDS, list(out)))
    plan['halo'], plan['resident_weights'] = bool(plan['halo']), bool(plan['resident_weights'])
    return plan
