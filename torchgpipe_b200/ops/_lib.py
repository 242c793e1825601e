"""Loader for the native sm_100a library (``libtgb200.so``, built in-tree by ``csrc/Makefile``).

The library has a plain C ABI (``include/tgb200.h``); it is bound with ``ctypes`` so no
PyTorch C++ extension build is needed and the same entry points serve non-Python callers.
On a machine with a GPU a missing library is an error for every op that needs it (no silent
fallback to PyTorch kernels).
"""
import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB_PATH = os.path.join(_HERE, 'libtgb200.so')
_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None


def lib_path() -> str:
    return _LIB_PATH


def load() -> Optional[ctypes.CDLL]:
    """Returns the loaded library, or ``None`` (with :func:`load_error` set) if unavailable."""
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        _load_error = f'{_LIB_PATH} not found: run `make -C csrc` (or __graft_entry__.build())'
        return None
    try:
        lib = ctypes.CDLL(_LIB_PATH)
    except OSError as exc:      # e.g. libcudart missing on a CPU-only box
        _load_error = f'cannot load {_LIB_PATH}: {exc}'
        return None
    lib.tgb_last_error.restype = ctypes.c_char_p
    lib.tgb_launch_count.restype = ctypes.c_longlong
    _lib = lib
    return _lib


def load_error() -> Optional[str]:
    return _load_error


def require() -> ctypes.CDLL:
    lib = load()
    if lib is None:
        raise RuntimeError(f'torchgpipe_b200 native library unavailable: {_load_error}')
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = require().tgb_last_error().decode()
        raise RuntimeError(f'{what} failed: {msg}')


def stream_ptr(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())
