"""The allocation-free weight-gradient entry point, called the way a non-Python host would:
straight through the C ABI of ``libtgb200.so``.

Who needs it: callers that may not allocate device memory while they run - an inference /
real-time loop living inside a fixed memory reservation, a C or C++ program with no caching
allocator (every ``cudaMalloc`` would stall the GPU), or code recording a step for replay.
``tgb_conv2d_wgrad_tf32`` asks them for a scratch buffer; ``tgb_conv2d_wgrad_tf32_noalloc``
keeps the partial sums in a 32 MiB buffer that is part of the library instead.

The test drives both entry points with the same inputs and requires the same gradients, over
several calls in a row (the built-in buffer is reused and still holds the previous call's
partials), checks that free device memory does not move across the calls, and that a problem
too large for the buffer is refused with a message naming the limit.

This file sorts last on purpose: if it fails under ``pytest -x`` every other test has already run.
It does not depend on any other test having run (or not run) before it.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

CL = torch.channels_last
LIMIT = 32 << 20


def _lib():
    from torchgpipe_b200.ops import _lib
    lib = _lib.require()
    lib.tgb_conv2d_wgrad_scratch_bytes.restype = ctypes.c_longlong
    lib.tgb_last_error.restype = ctypes.c_char_p
    return lib


def _ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _problem(N, C, H, W, K, R, S, ph, pw, seed):
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.randn(N, C, H, W, device='cuda', generator=g).contiguous(memory_format=CL)
    Ho, Wo = H + 2 * ph - R + 1, W + 2 * pw - S + 1
    dy = torch.randn(N, K, Ho, Wo, device='cuda', generator=g).contiguous(memory_format=CL)
    dims = [ctypes.c_int(v) for v in (N, H, W, C, K, R, S, ph, pw)]
    return x, dy, dims, (K, C, R, S)


def _with_caller_scratch(lib, x, dy, dims, wshape):
    need = lib.tgb_conv2d_wgrad_scratch_bytes(*dims)
    assert need >= 0, lib.tgb_last_error()
    scratch = torch.empty(max(need // 4, 1), device='cuda')
    dw = torch.full(wshape, float('nan'), device='cuda').contiguous(memory_format=CL)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.tgb_conv2d_wgrad_tf32(_ptr(dy), _ptr(x), _ptr(dw), _ptr(scratch),
                                   ctypes.c_longlong(need), *dims, ctypes.c_int(0), stream)
    assert rc == 0, lib.tgb_last_error()
    return dw, need


def _no_alloc(lib, x, dy, dims, dw):
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.tgb_conv2d_wgrad_tf32_noalloc(_ptr(dy), _ptr(x), _ptr(dw), *dims,
                                             ctypes.c_int(0), stream)


def rel_err(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def test_no_alloc_wgrad_matches_regular_entry_point_on_consecutive_calls():
    lib = _lib()
    # many pixel splits (scratch in use), few splits, and the first one again: the second and
    # third calls find the previous call's partials still sitting in the built-in buffer
    shapes = [(4, 64, 96, 96, 64, 3, 3, 1, 1), (2, 32, 24, 40, 96, 3, 3, 1, 1),
              (4, 64, 96, 96, 64, 3, 3, 1, 1), (3, 256, 12, 12, 128, 3, 3, 1, 1)]
    used_scratch = False
    for k, shape in enumerate(shapes):
        x, dy, dims, wshape = _problem(*shape, seed=100 + k)
        want, need = _with_caller_scratch(lib, x, dy, dims, wshape)
        used_scratch |= need > 0
        assert need <= LIMIT
        got = torch.full(wshape, float('nan'), device='cuda').contiguous(memory_format=CL)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        rc = _no_alloc(lib, x, dy, dims, got)
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0, lib.tgb_last_error()
        if k > 0:       # the very first call may load the kernel image
            assert free_after == free_before, 'the call changed the amount of free device memory'
        assert torch.isfinite(got).all()
        assert rel_err(got, want) < 2e-3        # tolerance of the convolution tests
        assert torch.equal(got, want)           # and in fact the same fixed-order sums
    assert used_scratch, 'no case exercised the scratch path'


def test_no_alloc_wgrad_agrees_with_fp32_reference():
    import torch.nn.functional as F
    lib = _lib()
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        x, dy, dims, wshape = _problem(2, 64, 48, 48, 64, 3, 3, 1, 1, seed=7)
        got = torch.empty(wshape, device='cuda').contiguous(memory_format=CL)
        for _ in range(2):
            assert _no_alloc(lib, x, dy, dims, got) == 0, lib.tgb_last_error()
        w = torch.zeros(wshape, device='cuda', requires_grad=True)
        F.conv2d(x, w, padding=1).backward(dy)
        assert rel_err(got, w.grad) < 2e-3
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_no_alloc_wgrad_refuses_what_does_not_fit():
    lib = _lib()
    # 5x5 filter, 64 -> 128 channels on a large image: dozens of pixel splits x 819 KB of
    # partials each is more than the 32 MiB the library carries
    x, dy, dims, wshape = _problem(8, 64, 64, 64, 128, 5, 5, 2, 2, seed=3)
    need = lib.tgb_conv2d_wgrad_scratch_bytes(*dims)
    if need <= LIMIT:
        pytest.skip(f'plan needs only {need} bytes on this GPU')
    dw = torch.zeros(wshape, device='cuda').contiguous(memory_format=CL)
    rc = _no_alloc(lib, x, dy, dims, dw)
    assert rc != 0
    msg = lib.tgb_last_error().decode()
    assert str(LIMIT) in msg and str(need) in msg
    torch.cuda.synchronize()
    assert dw.abs().sum().item() == 0       # nothing was launched
