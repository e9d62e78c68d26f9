"""ctypes binding to ``libb200ddp.so`` (C ABI declared in ``include/b200ddp.h``).

The library is built in-tree by ``python -m b200ddp._native.build`` (plain
``nvcc -gencode arch=compute_100a,code=sm_100a``; no torch headers), so the
``.so`` travels with the source tree.  Every entry point takes raw device
pointers and a ``cudaStream_t``; tensors are passed as ``data_ptr()`` and the
current torch stream, which keeps the calls capturable in CUDA graphs.

On a box with a GPU the library is mandatory: a missing or unloadable ``.so``
raises instead of silently falling back to eager PyTorch.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ddp.so")

_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None
_inited = False
launch_count = 0  # kernels launched through this binding (bench.py reports it)


def lib() -> ctypes.CDLL:
    global _lib, _load_error
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m b200ddp._native.build`")
        try:
            _lib = ctypes.CDLL(LIB_PATH)
        except OSError as e:  # pragma: no cover
            _load_error = str(e)
            raise RuntimeError(f"cannot load {LIB_PATH}: {e}") from e
        from . import bindings
        bindings.declare(_lib)
    return _lib


def available() -> bool:
    """True when the CUDA kernels can run.  Raises on a GPU box without the library."""
    global _inited
    if not torch.cuda.is_available():
        return False
    L = lib()
    if not _inited:
        L.b200_init.restype = ctypes.c_int
        check(L.b200_init(), "b200_init")
        _inited = True
    return True


def stream_ptr(device=None) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def optimizer_step(opt, start: int, end: int, grad_scale: float, *, device_step: bool = False,
                   advance: bool = False) -> None:
    from . import bindings
    bindings.optimizer_step(opt, start, end, grad_scale, device_step=device_step, advance=advance)
