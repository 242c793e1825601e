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

__all__ = ['linear', 'conv2d', 'conv2d_forward', 'conv2d_dgrad', 'conv_kind', 'stem_conv2d',
           'head_conv2d', 'factorized_reduce_conv', 'refresh_stem_weight', 'add_channel_bias', 'instnorm_leaky', 'channel_moments', 'conv_supported',
           'refresh_dgrad_weight', 'batchnorm_act', 'maxpool2x2', 'upsample2x_cat', 'pool2d',
           'pool_supported', 'global_avgpool', 'add_act', 'batchnorm_add_act', 'conv2d_plan']

_CL = torch.channels_last
_c_int = ctypes.c_int
_c_float = ctypes.c_float
_c_ll = ctypes.c_longlong


def _cl(t: Tensor) -> Tensor:
    return t if t.is_contiguous(memory_format=_CL) else t.contiguous(memory_format=_CL)


def _pair(v) -> Tuple[int, int]:
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


def conv_kind(c_in: int, c_out: int, kernel_size, stride, dilation, groups: int) -> Optional[str]:
    """Which native path covers a convolution: ``'tc'`` (implicit GEMM on tcgen05, strides up to
    2x2), ``'stem'`` (1-4 input channels, filter rows of at most 8 taps: packed-window tcgen05
    path), ``'head'`` (1x1, stride 1, at most 4 output channels: streaming dot product) or
    ``None`` (not covered)."""
    sh, sw = _pair(stride)
    R, S = _pair(kernel_size)
    if _pair(dilation) != (1, 1) or groups != 1 or sh < 1 or sw < 1 or sh * sw > 4:
        return None
    if c_in % 4 == 0 and c_out % 4 == 0:
        return 'tc'
    if c_in <= 4 and c_out % 4 == 0 and S <= 8:
        return 'stem'
    if c_in % 4 == 0 and c_out <= 4 and (R, S, sh, sw) == (1, 1, 1, 1):
        return 'head'
    return None


def conv_supported(c_in: int, c_out: int, stride, dilation, groups: int,
                   kernel_size=3) -> bool:
    """Whether the tcgen05 implicit-GEMM kernels cover this convolution directly."""
    return conv_kind(c_in, c_out, kernel_size, stride, dilation, groups) == 'tc'


def _out_hw(H: int, W: int, R: int, S: int, padding, stride) -> Tuple[int, int]:
    return ((H + 2 * padding[0] - R) // stride[0] + 1, (W + 2 * padding[1] - S) // stride[1] + 1)


def conv2d_forward(x: Tensor, w: Tensor, padding: Tuple[int, int], stats_mode: int = 0,
                   stride: Tuple[int, int] = (1, 1), out: Optional[Tensor] = None,
                   out_hw: Optional[Tuple[int, int]] = None,
                   ) -> Tuple[Tensor, Optional[Tensor]]:
    """Raw forward: returns ``(y, stats)``; ``stats`` is ``[N,K,2]`` (mode 1), ``[K,2]`` (mode 2)
    or ``None``. No autograd. ``out``: write into this ``[N,K,Ho,Wo]`` tensor, which may be a
    channel slice of a wider channels-last tensor (a concatenation written in place).
    ``out_hw``: fix the output extent instead of deriving it (padding that differs between the
    near and the far edge: the far edge is zero-filled as far as the extent asks)."""
    lib = _lib.require()
    x, w = _cl(x), _cl(w)
    N, C, H, W = x.shape
    K, Cw, R, S = w.shape
    assert C == Cw, f'channel mismatch: input {C}, weight {Cw}'
    ph, pw = padding
    sh, sw = stride
    Ho, Wo = out_hw if out_hw is not None else _out_hw(H, W, R, S, padding, stride)
    if out is None:
        y = torch.empty((N, K, Ho, Wo), device=x.device, dtype=torch.float32, memory_format=_CL)
        y_pixel = K
    else:
        y = out
        assert y.shape == (N, K, Ho, Wo) and _is_channel_slice(y), \
            'out must be a channel slice of an NHWC tensor'
        y_pixel = _pixel_stride(y)
    stats = None            # zeroed by the launch itself (negative stats_mode)
    if stats_mode == 1:
        stats = torch.empty((N, K, 2), device=x.device, dtype=torch.float32)
    elif stats_mode == 2:
        stats = torch.empty((K, 2), device=x.device, dtype=torch.float32)
    stats_mode = -stats_mode
    with torch.cuda.device(x.device):
        if (sh, sw) == (1, 1) and y_pixel == K and out_hw is None:
            rc = lib.tgb_conv2d_fprop_tf32(_lib.ptr(x), _lib.ptr(w), _lib.ptr(y), _lib.ptr(stats),
                                           _c_int(N), _c_int(H), _c_int(W), _c_int(C), _c_int(K),
                                           _c_int(R), _c_int(S), _c_int(ph), _c_int(pw),
                                           _c_int(stats_mode), _lib.stream_ptr(x.device))
        else:
            rc = lib.tgb_conv2d_fprop_strided_tf32(
                _lib.ptr(x), _lib.ptr(w), _lib.ptr(y), _lib.ptr(stats), _c_int(N), _c_int(H),
                _c_int(W), _c_int(C), _c_int(K), _c_int(R), _c_int(S), _c_int(ph), _c_int(pw),
                _c_int(sh), _c_int(sw), _c_int(Ho if out_hw else 0), _c_int(Wo if out_hw else 0),
                _c_ll(y_pixel), _c_int(stats_mode), _lib.stream_ptr(x.device))
    _lib.check(rc, 'tgb_conv2d_fprop_tf32')
    return y, stats


def _conv2d_wgrad(dy: Tensor, x: Tensor, w_shape, padding: Tuple[int, int],
                  stride: Tuple[int, int] = (1, 1), fixed_hw: bool = False) -> Tensor:
    """``dy`` may be a channel slice of a wider NHWC tensor (gradient of a concatenation)."""
    lib = _lib.require()
    lib.tgb_conv2d_wgrad_scratch_bytes.restype = _c_ll
    lib.tgb_conv2d_wgrad_strided_scratch_bytes.restype = _c_ll
    N, C, H, W = x.shape
    K, _, R, S = w_shape
    ph, pw = padding
    sh, sw = stride
    dy_pixel = _pixel_stride(dy)
    plain = (sh, sw) == (1, 1) and dy_pixel == K and not fixed_hw
    ho, wo = (dy.size(2), dy.size(3)) if fixed_hw else (0, 0)
    args = (_c_int(N), _c_int(H), _c_int(W), _c_int(C), _c_int(K), _c_int(R), _c_int(S),
            _c_int(ph), _c_int(pw))
    if plain:
        need = lib.tgb_conv2d_wgrad_scratch_bytes(*args)
    else:
        need = lib.tgb_conv2d_wgrad_strided_scratch_bytes(*args, _c_int(sh), _c_int(sw),
                                                          _c_int(ho), _c_int(wo), _c_ll(dy_pixel))
    if need < 0:
        _lib.check(1, 'tgb_conv2d_wgrad_scratch_bytes')
    scratch = torch.empty(need // 4, device=x.device, dtype=torch.float32) if need else None
    dw = torch.empty(w_shape, device=x.device, dtype=torch.float32, memory_format=_CL)
    with torch.cuda.device(x.device):
        if plain:
            rc = lib.tgb_conv2d_wgrad_tf32(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(dw),
                                           _lib.ptr(scratch), _c_ll(need), *args, _c_int(0),
                                           _lib.stream_ptr(x.device))
        else:
            rc = lib.tgb_conv2d_wgrad_strided_tf32(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(dw),
                                                   _lib.ptr(scratch), _c_ll(need), *args,
                                                   _c_int(sh), _c_int(sw), _c_int(ho), _c_int(wo),
                                                   _c_ll(dy_pixel), _c_int(0),
                                                   _lib.stream_ptr(x.device))
    _lib.check(rc, 'tgb_conv2d_wgrad_tf32')
    return dw


def _dgrad_weight(w: Tensor, cache: Optional[dict], padding: Tuple[int, int] = (0, 0),
                  stride: Tuple[int, int] = (1, 1)) -> Tensor:
    """Weights of the data-gradient pass (KRSC memory): tap-flipped and channel-transposed; for a
    strided convolution one such bank per input-pixel parity class, concatenated (flat buffer).

    ``cache`` is a dict owned by the calling module: with several micro-batches per optimiser
    step the transform then runs once per parameter version instead of once per backward call.
    """
    if cache is not None:
        if cache.get('version') == w._version and cache.get('ptr') == w.data_ptr():
            return cache['wt']
    lib = _lib.require()
    K, C, R, S = w.shape
    old = cache.get('wt') if cache is not None else None
    src = _cl(w.detach())
    if stride == (1, 1):
        if old is not None and old.shape == (C, K, R, S) and old.device == w.device:
            wt = old                # keep the address: captured CUDA graphs read this buffer
        else:
            wt = torch.empty((C, K, R, S), device=w.device, dtype=torch.float32,
                             memory_format=_CL)
        with torch.cuda.device(w.device):
            rc = lib.tgb_dgrad_weight(_lib.ptr(src), _lib.ptr(wt), _c_int(K), _c_int(C),
                                      _c_int(R), _c_int(S), _lib.stream_ptr(w.device))
        _lib.check(rc, 'tgb_dgrad_weight')
    else:
        lib.tgb_dgrad_weight_strided_elems.restype = _c_ll
        geom = (_c_int(K), _c_int(C), _c_int(R), _c_int(S), _c_int(padding[0]),
                _c_int(padding[1]), _c_int(stride[0]), _c_int(stride[1]))
        elems = max(4, int(lib.tgb_dgrad_weight_strided_elems(*geom)))
        if old is not None and old.dim() == 1 and old.numel() == elems \
                and old.device == w.device:
            wt = old
        else:
            wt = torch.empty(elems, device=w.device, dtype=torch.float32)
        with torch.cuda.device(w.device):
            rc = lib.tgb_dgrad_weight_strided(_lib.ptr(src), _lib.ptr(wt), *geom,
                                              _lib.stream_ptr(w.device))
        _lib.check(rc, 'tgb_dgrad_weight_strided')
    if cache is not None:
        cache['version'], cache['ptr'], cache['wt'] = w._version, w.data_ptr(), wt
        cache['geom'] = (padding, stride)
    return wt


def refresh_dgrad_weight(w: Tensor, cache: dict) -> None:
    """Brings a cached dgrad weight up to date in place (no-op if never built or current)."""
    if 'wt' in cache and (cache.get('version') != w._version or cache.get('ptr') != w.data_ptr()):
        padding, stride = cache.get('geom', ((0, 0), (1, 1)))
        _dgrad_weight(w, cache, padding, stride)


def refresh_stem_weight(w: Tensor, cache: dict) -> None:
    """Brings a cached packed stem weight up to date in place (no-op if never built)."""
    if 'stem_w' in cache and (cache.get('stem_version') != w._version
                              or cache.get('stem_ptr') != w.data_ptr()):
        cache['stem_w'].copy_(_stem_weight(w))
        cache['stem_version'], cache['stem_ptr'] = w._version, w.data_ptr()


def add_channel_bias(y: Tensor, bias: Tensor) -> Tensor:
    """``y + bias[None, :, None, None]`` (one ATen broadcast add; convolution biases are rare in
    the benchmark models - every conv is followed by a normalisation)."""
    return y + bias.view(1, -1, 1, 1)


def _pixel_stride(t: Tensor) -> int:
    """Elements between consecutive pixels of an NHWC(-slice) tensor; strides of size-1
    dimensions carry no information, so take the innermost dimension that has extent."""
    N, K, H, W = t.shape
    if W > 1:
        return t.stride(3)
    if H > 1:
        return t.stride(2)
    if N > 1:
        return t.stride(0)
    return K


def _is_channel_slice(t: Tensor) -> bool:
    """A ``[N,K,H,W]`` view whose pixels are K consecutive floats inside a wider (or equally
    wide) NHWC tensor: pixel (n,h,w) starts ``((n*H + h)*W + w) * pixel_stride`` elements in."""
    if t.dim() != 4 or (t.size(1) > 1 and t.stride(1) != 1) or t.data_ptr() % 16:
        return False
    N, K, H, W = t.shape
    ps = _pixel_stride(t)
    return (ps % 4 == 0 and ps >= K and (W == 1 or t.stride(3) == ps)
            and (H == 1 or t.stride(2) == W * ps) and (N == 1 or t.stride(0) == H * W * ps))


def conv2d_dgrad(dy: Tensor, w: Tensor, x_shape, padding: Tuple[int, int],
                 stride: Tuple[int, int] = (1, 1), cache: Optional[dict] = None,
                 out: Optional[Tensor] = None, keep_untouched: bool = False,
                 fixed_hw: bool = False) -> Tensor:
    """Data gradient of ``conv2d`` on the tcgen05 kernels (one stride-1 launch per input-pixel
    parity class for a strided convolution). ``dy`` may be a channel slice of a wider NHWC
    gradient. ``out`` / ``keep_untouched``: write into an existing dense dX and leave the pixels
    this convolution never reads alone."""
    K, C, R, S = w.shape
    N, _, H, W = x_shape
    ph, pw = padding
    wt = _dgrad_weight(w, cache, padding, stride)
    if stride == (1, 1) and out is None:
        dx, _ = conv2d_forward(_cl(dy), wt, (R - 1 - ph, S - 1 - pw), 0)
        return dx
    if not _is_channel_slice(dy):
        dy = _cl(dy)
    dx = out if out is not None else torch.empty((N, C, H, W), device=dy.device,
                                                 dtype=torch.float32, memory_format=_CL)
    assert dx.is_contiguous(memory_format=_CL)
    with torch.cuda.device(dy.device):
        rc = _lib.require().tgb_conv2d_dgrad_tf32(
            _lib.ptr(dy), _lib.ptr(wt), _lib.ptr(dx), _c_int(N), _c_int(H), _c_int(W), _c_int(C),
            _c_int(K), _c_int(R), _c_int(S), _c_int(ph), _c_int(pw), _c_int(stride[0]),
            _c_int(stride[1]), _c_int(dy.size(2) if fixed_hw else 0),
            _c_int(dy.size(3) if fixed_hw else 0), _c_ll(_pixel_stride(dy)),
            _c_int(int(keep_untouched)),
            _lib.stream_ptr(dy.device))
    _lib.check(rc, 'tgb_conv2d_dgrad_tf32')
    return dx


def _dense_nhwc(t: Tensor) -> bool:
    return t.is_contiguous(memory_format=_CL)


class _Conv2d(torch.autograd.Function):
    """Convolution on tcgen05; backward = dgrad (same kernel, derived weights) + wgrad."""

    @staticmethod
    def forward(ctx, x: Tensor, w: Tensor, ph: int, pw: int, stats_mode: int,  # type: ignore[override]
                cache: Optional[dict], sh: int = 1, sw: int = 1):
        x, w = _cl(x), _cl(w)
        y, stats = conv2d_forward(x, w, (ph, pw), stats_mode, (sh, sw))
        ctx.save_for_backward(x, w)
        ctx.padding = (ph, pw)
        ctx.stride = (sh, sw)
        ctx.cache = cache
        if stats is None:
            stats = x.new_empty(0)
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, dy: Tensor, _dstats):  # type: ignore[override]
        x, w = ctx.saved_tensors
        # a channel slice of an NHWC gradient (this output was one part of a concatenation) is
        # read in place; anything else is made dense NHWC
        if not _is_channel_slice(dy):
            dy = _cl(dy)
        dx = dw = None
        if ctx.needs_input_grad[1]:
            dw = _conv2d_wgrad(dy, x, w.shape, ctx.padding, ctx.stride)
        if ctx.needs_input_grad[0]:
            dx = conv2d_dgrad(dy, w, x.shape, ctx.padding, ctx.stride, ctx.cache)
        return dx, dw, None, None, None, None, None, None


def conv2d(x: Tensor, w: Tensor, padding: Tuple[int, int], stats_mode: int = 0,
           cache: Optional[dict] = None, stride: Tuple[int, int] = (1, 1),
           ) -> Tuple[Tensor, Optional[Tensor]]:
    """Differentiable convolution on tcgen05 (strides up to 2x2, any integer padding including
    negative = crop). Returns ``(y, stats or None)``.

    ``cache``: optional dict owned by the caller, reused across calls to keep derived weights.
    """
    y, stats = _Conv2d.apply(x, w, int(padding[0]), int(padding[1]), int(stats_mode), cache,
                             int(stride[0]), int(stride[1]))
    return y, (stats if stats.numel() else None)


class _FactorizedReduceConv(torch.autograd.Function):
    """The two offset 1x1 stride-2 convolutions of AmoebaNet's ``FactorizedReduce`` (reference
    ``benchmarks/models/amoebanet/operations.py:24-40``) and their channel concatenation.

    ``conv1`` reads the pixels (2i, 2j); ``conv2`` - applied in the reference to the input
    shifted by one pixel and zero-padded at the far edge - reads (2i+1, 2j+1), which is a 1x1
    stride-2 convolution with padding -1. Each writes its half of the channels of one output
    tensor, so there is no shifted copy and no ``torch.cat``; backward reads the halves of dY
    in place, and the two data gradients fill disjoint pixel classes of one dX.
    Returns ``(out, stats [K,2])`` (per-channel sum / sum of squares for the following BN)."""

    @staticmethod
    def forward(ctx, x: Tensor, w1: Tensor, w2: Tensor, cache1: dict, cache2: dict):  # type: ignore[override]
        x, w1, w2 = _cl(x), _cl(w1), _cl(w2)
        N, C, H, W = x.shape
        k1, k2 = w1.size(0), w2.size(0)
        Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        out = torch.empty((N, k1 + k2, Ho, Wo), device=x.device, dtype=torch.float32,
                          memory_format=_CL)
        if H < 2 or W < 2:
            out[:, k1:].zero_()         # conv2 sees only padding
        _, s1 = conv2d_forward(x, w1, (0, 0), 2, (2, 2), out=out[:, :k1])
        if H >= 2 and W >= 2:
            _, s2 = conv2d_forward(x, w2, (-1, -1), 2, (2, 2), out=out[:, k1:],
                                   out_hw=(Ho, Wo))
        else:
            s2 = torch.zeros((k2, 2), device=x.device, dtype=torch.float32)
        ctx.save_for_backward(x, w1, w2)
        ctx.caches = (cache1, cache2)
        stats = torch.cat([s1, s2], dim=0)
        ctx.mark_non_differentiable(stats)
        return out, stats

    @staticmethod
    def backward(ctx, dout: Tensor, _dstats):  # type: ignore[override]
        x, w1, w2 = ctx.saved_tensors
        k1 = w1.size(0)
        dout = _cl(dout)
        dy1, dy2 = dout[:, :k1], dout[:, k1:]
        dx = dw1 = dw2 = None
        if ctx.needs_input_grad[1]:
            dw1 = _conv2d_wgrad(dy1, x, w1.shape, (0, 0), (2, 2))
        if ctx.needs_input_grad[2]:
            dw2 = _conv2d_wgrad(dy2, x, w2.shape, (-1, -1), (2, 2), fixed_hw=True)
        if ctx.needs_input_grad[0]:
            dx = torch.zeros_like(x, memory_format=_CL)
            conv2d_dgrad(dy1, w1, x.shape, (0, 0), (2, 2), ctx.caches[0], out=dx,
                         keep_untouched=True)
            if x.size(2) >= 2 and x.size(3) >= 2:
                conv2d_dgrad(dy2, w2, x.shape, (-1, -1), (2, 2), ctx.caches[1], out=dx,
                             keep_untouched=True, fixed_hw=True)
        return dx, dw1, dw2, None, None


def factorized_reduce_conv(x: Tensor, w1: Tensor, w2: Tensor, cache1: dict, cache2: dict,
                           ) -> Tuple[Tensor, Tensor]:
    """``cat([conv1x1_s2(x, w1), conv1x1_s2(shift(x), w2)], 1)`` and its BN statistics."""
    return _FactorizedReduceConv.apply(x, w1, w2, cache1, cache2)


# ---------------------------------------------------------------- small-channel stem and head
def _stem_weight(w: Tensor) -> Tensor:
    """``[K,C,R,S]`` -> ``[K,R,32]`` with entry ``4*s + c`` (zero padded)."""
    K, C, R, S = w.shape
    wv = torch.zeros((K, R, 8, 4), device=w.device, dtype=torch.float32)
    wv[:, :, :S, :C] = w.detach().permute(0, 2, 3, 1)
    return wv.view(K, R, 32)


class _StemConv2d(torch.autograd.Function):
    """Convolution of an image with 1-4 channels (any layout) on the packed-window tcgen05 path
    (``csrc/conv_small.cu``). Saves the packed image (4 channels, zero halo) for the weight
    gradient; the data gradient (rarely needed: the input is data) runs the general tcgen05
    data-gradient kernels on the filter bank padded to 4 input channels."""

    @staticmethod
    def forward(ctx, x: Tensor, w: Tensor, ph: int, pw: int, sh: int, sw: int,  # type: ignore[override]
                stats_mode: int, cache: Optional[dict], relu_input: bool = False):
        lib = _lib.require()
        N, C, H, W = x.shape
        K, _, R, S = w.shape
        Ho, Wo = _out_hw(H, W, R, S, (ph, pw), (sh, sw))

        Hp, Wp = H + 2 * ph, W + 2 * pw - S + 8
        xp = torch.empty((N, Hp, Wp, 4), device=x.device, dtype=torch.float32)
        wv = None
        if cache is not None and cache.get('stem_version') == w._version \
                and cache.get('stem_ptr') == w.data_ptr():
            wv = cache['stem_w']
        if wv is None:
            fresh = _stem_weight(w)
            wv = cache.get('stem_w') if cache is not None else None
            if wv is not None and wv.shape == fresh.shape and wv.device == fresh.device:
                wv.copy_(fresh)         # keep the address (captured graphs read it)
            else:
                wv = fresh
            if cache is not None:
                cache['stem_version'], cache['stem_ptr'], cache['stem_w'] = \
                    w._version, w.data_ptr(), wv
        y = torch.empty((N, K, Ho, Wo), device=x.device, dtype=torch.float32, memory_format=_CL)
        stats = None        # zeroed by the launch itself (negative stats_mode)
        if stats_mode == 1:
            stats = torch.empty((N, K, 2), device=x.device, dtype=torch.float32)
        elif stats_mode == 2:
            stats = torch.empty((K, 2), device=x.device, dtype=torch.float32)
        stats_mode = -stats_mode
        with torch.cuda.device(x.device):
            st = _lib.stream_ptr(x.device)
            rc = lib.tgb_stem_pack(_lib.ptr(x), _lib.ptr(xp), _c_int(N), _c_int(C), _c_int(H),
                                   _c_int(W), _c_ll(x.stride(0)), _c_ll(x.stride(1)),
                                   _c_ll(x.stride(2)), _c_ll(x.stride(3)), _c_int(ph), _c_int(pw),
                                   _c_int(Hp), _c_int(Wp), _c_int(int(relu_input)), st)
            _lib.check(rc, 'tgb_stem_pack')
            rc = lib.tgb_conv2d_stem_fprop_tf32(_lib.ptr(xp), _lib.ptr(wv), _lib.ptr(y),
                                                _lib.ptr(stats), _c_int(N), _c_int(Hp),
                                                _c_int(Wp), _c_int(K), _c_int(R), _c_int(sh),
                                                _c_int(sw), _c_int(stats_mode), st)
            _lib.check(rc, 'tgb_conv2d_stem_fprop_tf32')
        ctx.save_for_backward(xp, w, x if x.requires_grad else x.new_empty(0))
        ctx.geom = (N, C, H, W, K, R, S, ph, pw, sh, sw, Hp, Wp)
        ctx.relu_input = bool(relu_input)
        if stats is None:
            stats = x.new_empty(0)
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, dy: Tensor, _dstats):  # type: ignore[override]
        lib = _lib.require()
        lib.tgb_conv2d_stem_wgrad_scratch_bytes.restype = _c_ll
        xp, w, x = ctx.saved_tensors
        N, C, H, W, K, R, S, ph, pw, sh, sw, Hp, Wp = ctx.geom
        dy = _cl(dy)
        dx = dw = None
        if ctx.needs_input_grad[1]:
            geom = (_c_int(N), _c_int(Hp), _c_int(Wp), _c_int(K), _c_int(R), _c_int(sh),
                    _c_int(sw))
            need = lib.tgb_conv2d_stem_wgrad_scratch_bytes(*geom)
            if need < 0:
                _lib.check(1, 'tgb_conv2d_stem_wgrad_scratch_bytes')
            scratch = torch.empty(need // 4, device=dy.device, dtype=torch.float32) \
                if need else None
            dwv = torch.empty((K, R, 32), device=dy.device, dtype=torch.float32)
            with torch.cuda.device(dy.device):
                rc = lib.tgb_conv2d_stem_wgrad_tf32(_lib.ptr(dy), _lib.ptr(xp), _lib.ptr(dwv),
                                                    _lib.ptr(scratch), _c_ll(need), *geom,
                                                    _lib.stream_ptr(dy.device))
            _lib.check(rc, 'tgb_conv2d_stem_wgrad_tf32')
            dw = dwv.view(K, R, 8, 4)[:, :, :S, :C].permute(0, 3, 1, 2).contiguous(
                memory_format=_CL)
        if ctx.needs_input_grad[0]:
            # Rarely wanted (the input is data), but the eager pipeline's ordering edges make
            # the first stage's micro-batches "require grad": the general data-gradient kernels
            # on the filter bank padded to 4 input channels, then drop the padding channels.
            w4 = torch.zeros((K, 4, R, S), device=w.device, dtype=torch.float32).contiguous(
                memory_format=_CL)
            w4[:, :C] = w.detach()
            dx = conv2d_dgrad(dy, w4, (N, 4, H, W), (ph, pw), (sh, sw), None)[:, :C]
            if ctx.relu_input:
                dx = dx * (x > 0)
        return dx, dw, None, None, None, None, None, None, None


def stem_conv2d(x: Tensor, w: Tensor, padding: Tuple[int, int], stride: Tuple[int, int] = (1, 1),
                stats_mode: int = 0, cache: Optional[dict] = None, relu_input: bool = False,
                ) -> Tuple[Tensor, Optional[Tensor]]:
    """Convolution of a 1-4 channel image (NCHW or channels-last, fp32 CUDA) with filter rows of
    at most 8 taps, on the packed-window tcgen05 path. Returns ``(y, stats or None)``.
    ``relu_input``: convolve ``relu(x)`` (the ReLU is folded into the packing pass)."""
    assert x.size(1) <= 4 and w.size(3) <= 8 and padding[0] >= 0 and padding[1] >= 0
    y, stats = _StemConv2d.apply(x, w, int(padding[0]), int(padding[1]), int(stride[0]),
                                 int(stride[1]), int(stats_mode), cache, bool(relu_input))
    return y, (stats if stats.numel() else None)


class _HeadConv(torch.autograd.Function):
    """1x1 convolution to at most 4 output channels (+ bias): a streaming dot product."""

    @staticmethod
    def forward(ctx, x: Tensor, w: Tensor, bias: Optional[Tensor]):  # type: ignore[override]
        lib = _lib.require()
        x = _cl(x)
        N, C, H, W = x.shape
        K = w.size(0)
        w2 = w.detach().reshape(K, C).contiguous()
        y = torch.empty((N, K, H, W), device=x.device, dtype=torch.float32, memory_format=_CL)
        with torch.cuda.device(x.device):
            rc = lib.tgb_head_fwd(_lib.ptr(x), _lib.ptr(w2), _lib.ptr(bias), _lib.ptr(y),
                                  _c_ll(N * H * W), _c_int(C), _c_int(K),
                                  _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_head_fwd')
        ctx.save_for_backward(x, w2)
        ctx.w_shape = tuple(w.shape)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        lib = _lib.require()
        x, w2 = ctx.saved_tensors
        N, C, H, W = x.shape
        K = w2.size(0)
        dy = _cl(dy)
        need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        need_db = ctx.has_bias and ctx.needs_input_grad[2]
        dx = torch.empty_like(x, memory_format=_CL) if need_dx else None
        dw = torch.empty((K, C), device=x.device, dtype=torch.float32) if need_dw else None
        db = torch.empty(K, device=x.device, dtype=torch.float32) if need_db else None
        with torch.cuda.device(x.device):
            rc = lib.tgb_head_bwd(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(w2), _lib.ptr(dx),
                                  _lib.ptr(dw), _lib.ptr(db), _c_ll(N * H * W), _c_int(C),
                                  _c_int(K), _lib.stream_ptr(x.device))
        _lib.check(rc, 'tgb_head_bwd')
        if dw is not None:
            dw = dw.view(ctx.w_shape)
        return dx, dw, db


def head_conv2d(x: Tensor, w: Tensor, bias: Optional[Tensor] = None) -> Tensor:
    """``F.conv2d(x, w, bias)`` for a 1x1 filter with 1-4 output channels (``C % 4 == 0``)."""
    return _HeadConv.apply(x, w, bias)


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
                eps: float, slope: float, drop_p: float = 0.0):
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
            if drop_p > 0.0:        # drop_scale holds uniform noise: the mask is made in-kernel
                rc = lib.tgb_in_prepare_dropout(_lib.ptr(stats), _lib.ptr(drop_scale),
                                                _c_float(drop_p), _lib.ptr(mean_coef), _c_int(N),
                                                _c_int(P), _c_int(C), _c_float(eps), st)
            else:
                rc = lib.tgb_in_prepare(_lib.ptr(stats), _lib.ptr(drop_scale),
                                        _lib.ptr(mean_coef), _c_int(N), _c_int(P), _c_int(C),
                                        _c_float(eps), st)
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
        return dy, None, None, None, None, None


def instnorm_leaky(y: Tensor, stats: Optional[Tensor] = None, drop_scale: Optional[Tensor] = None,
                   eps: float = 1e-5, slope: float = 0.01, drop_p: float = 0.0) -> Tensor:
    """Fused ``LeakyReLU(InstanceNorm2d(Dropout2d(y)))`` (no affine, batch statistics).

    ``stats`` are the per-(n,c) (sum, sum of squares) from the convolution epilogue (computed
    here if absent); ``drop_scale`` is ``[N,C]`` with entries 0 or ``1/(1-p)`` (``None`` = no
    dropout). With ``drop_p > 0``, ``drop_scale`` instead holds uniform random numbers in
    ``[0,1)`` and the keep / scale decision is made inside the prepare kernel (channel kept
    when ``u >= drop_p``). ``slope`` must be positive.
    """
    return _InstNormLeaky.apply(y, stats, drop_scale, float(eps), float(slope), float(drop_p))


def _bn_prepare(stats: Tensor, weight: Optional[Tensor], bias: Optional[Tensor],
                running_mean: Optional[Tensor], running_var: Optional[Tensor],
                tracked: Optional[Tensor], P: int, C: int, eps: float, momentum: float, st):
    """``(coef, saved)`` tables of a training-mode BatchNorm; updates the running estimates (and
    ``tracked`` = num_batches_tracked, inside the same kernel) when they are given."""
    lib = _lib.require()
    coef = torch.empty((C, 2), device=stats.device, dtype=torch.float32)
    saved = torch.empty((C, 2), device=stats.device, dtype=torch.float32)
    if tracked is not None and running_mean is not None:
        rc = lib.tgb_bn_prepare_tracked(_lib.ptr(stats), _lib.ptr(weight), _lib.ptr(bias),
                                        _lib.ptr(coef), _lib.ptr(saved), _lib.ptr(running_mean),
                                        _lib.ptr(running_var), _lib.ptr(tracked), _c_int(P),
                                        _c_int(C), _c_float(eps), _c_float(momentum), st)
    else:
        rc = lib.tgb_bn_prepare(_lib.ptr(stats), _lib.ptr(weight), _lib.ptr(bias),
                                _lib.ptr(coef), _lib.ptr(saved), _lib.ptr(running_mean),
                                _lib.ptr(running_var), _c_int(P), _c_int(C), _c_float(eps),
                                _c_float(momentum), st)
    _lib.check(rc, 'tgb_bn_prepare')
    return coef, saved


def _affine_grads(sums: Tensor, affine: Tuple[bool, bool]):
    """(dgamma, dbeta) as contiguous rows of ONE transposed copy of the ``[C,2]`` sums table
    (one small kernel instead of two strided clones per BatchNorm layer and micro-batch)."""
    if not (affine[0] or affine[1]):
        return None, None
    rows = sums.t().contiguous()        # [2, C]: row 0 = sum dn = dbeta, row 1 = dgamma
    return (rows[1] if affine[0] else None), (rows[0] if affine[1] else None)


class _BatchNormAct(torch.autograd.Function):
    """z = act(BatchNorm(y)) in training mode from precomputed per-channel (sum, sum of squares).

    Saves the convolution output ``y`` and two ``[C,2]`` tables; the ReLU mask and the
    normalised activation are re-derived from ``y`` in backward.
    """

    @staticmethod
    def forward(ctx, y: Tensor, stats: Tensor, weight: Optional[Tensor],  # type: ignore[override]
                bias: Optional[Tensor], running_mean: Optional[Tensor],
                running_var: Optional[Tensor], eps: float, momentum: float, relu: bool,
                tracked: Optional[Tensor] = None):
        y = _cl(y)
        N, C, H, W = y.shape
        P = N * H * W
        z = torch.empty_like(y, memory_format=_CL)
        lib = _lib.require()
        with torch.cuda.device(y.device):
            st = _lib.stream_ptr(y.device)
            coef, saved = _bn_prepare(stats, weight, bias, running_mean, running_var, tracked, P,
                                      C, eps, momentum, st)
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
        dweight, dbias = _affine_grads(sums, ctx.affine)
        return dy, None, dweight, dbias, None, None, None, None, None, None


def batchnorm_act(y: Tensor, stats: Optional[Tensor], weight: Optional[Tensor],
                  bias: Optional[Tensor], running_mean: Optional[Tensor] = None,
                  running_var: Optional[Tensor] = None, eps: float = 1e-5,
                  momentum: float = 0.1, relu: bool = False,
                  tracked: Optional[Tensor] = None) -> Tensor:
    """Fused training-mode ``ReLU?(BatchNorm2d(y))``.

    ``tracked``: the layer's ``num_batches_tracked`` (int64 CUDA tensor); the prepare kernel then
    increments it, and a negative ``momentum`` selects the cumulative average 1/count.

    ``stats``: per-channel ``[C,2]`` (sum, sum of squares) from the convolution epilogue
    (``stats_mode=2``); computed here when ``None``. When ``running_mean``/``running_var`` are
    given they are updated in place like ``nn.BatchNorm2d`` does (pass ``None`` to skip, e.g.
    during checkpoint recomputation or under DeferredBatchNorm).
    """
    if stats is None or stats.numel() == 0:
        stats = channel_moments(y)
    return _BatchNormAct.apply(y, stats, weight, bias, running_mean, running_var, float(eps),
                               float(momentum), bool(relu), tracked)


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
                running_var: Optional[Tensor], eps: float, momentum: float, relu: bool,
                tracked: Optional[Tensor] = None):
        lib = _lib.require()
        y, res = _cl(y), _cl(res)
        N, C, H, W = y.shape
        P = N * H * W
        z = torch.empty_like(y, memory_format=_CL)
        with torch.cuda.device(y.device):
            st = _lib.stream_ptr(y.device)
            coef, saved = _bn_prepare(stats, weight, bias, running_mean, running_var, tracked, P,
                                      C, eps, momentum, st)
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
        dweight, dbias = _affine_grads(sums, ctx.affine)
        return dy, None, dweight, dbias, dres, None, None, None, None, None, None


def batchnorm_add_act(y: Tensor, stats: Optional[Tensor], weight: Optional[Tensor],
                      bias: Optional[Tensor], res: Tensor,
                      running_mean: Optional[Tensor] = None,
                      running_var: Optional[Tensor] = None, eps: float = 1e-5,
                      momentum: float = 0.1, relu: bool = True,
                      tracked: Optional[Tensor] = None) -> Tensor:
    """Fused training-mode ``ReLU?(BatchNorm2d(y) + res)`` (``bn3 -> residual -> relu3`` of a
    bottleneck, reference ``benchmarks/models/resnet/bottleneck.py:31-79``). Arguments as
    :func:`batchnorm_act`."""
    if stats is None or stats.numel() == 0:
        stats = channel_moments(y)
    return _BatchNormAddAct.apply(y, stats, weight, bias, res, running_mean, running_var,
                                  float(eps), float(momentum), bool(relu), tracked)


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
