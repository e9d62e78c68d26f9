"""ctypes signatures for the peer-memory collectives (include/b200ddp.h)."""
from __future__ import annotations

import ctypes as C

MAX_RANKS = 8
FLAG_WORDS = 2 * MAX_RANKS + 2


class Comm(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int),
                ("grad", C.c_void_p * MAX_RANKS), ("push", C.c_void_p * MAX_RANKS),
                ("flags", C.c_void_p * MAX_RANKS),
                ("grad_mc", C.c_void_p), ("push_mc", C.c_void_p),
                ("timeout_ms", C.c_int), ("max_blocks", C.c_int),
                ("master_peer", C.c_void_p * MAX_RANKS), ("f32_mask", C.c_void_p),
                ("parity_stride", C.c_int64), ("no_exit_wait", C.c_int)]


def declare(L: C.CDLL) -> None:
    from .bindings import OptHParams
    L.b200_symm_alloc.argtypes = [C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]
    L.b200_symm_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    L.b200_symm_close.argtypes = [C.c_void_p]
    L.b200_symm_free.argtypes = [C.c_void_p]
    L.b200_allreduce_optimizer.argtypes = [
        C.POINTER(Comm), C.c_int, C.POINTER(OptHParams), C.c_int64, C.c_int64, C.c_int,
        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
        C.c_void_p]
    L.b200_allreduce.argtypes = [C.POINTER(Comm), C.c_int, C.c_int64, C.c_int64, C.c_int,
                                 C.c_void_p, C.c_float, C.c_void_p]
    L.b200_broadcast.argtypes = [C.POINTER(Comm), C.c_int, C.c_int64, C.c_int64, C.c_void_p]
    L.b200_broadcast.restype = C.c_int
    L.b200_vmm_support.argtypes = [C.POINTER(C.c_int)]
    L.b200_vmm_padded_size.argtypes = [C.c_int64, C.c_int, C.POINTER(C.c_int64)]
    L.b200_vmm_create.argtypes = [C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.b200_vmm_export.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    L.b200_vmm_import.argtypes = [C.c_int, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.b200_vmm_release.argtypes = [C.c_void_p]
    L.b200_mc_create.argtypes = [C.c_int64, C.c_int, C.POINTER(C.c_void_p)]
    L.b200_mc_import.argtypes = [C.c_int, C.c_int64, C.POINTER(C.c_void_p)]
    L.b200_mc_add_device.argtypes = [C.c_void_p]
    L.b200_mc_bind_map.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
    L.b200_close_fd.argtypes = [C.c_int]
    for name in ("b200_vmm_support", "b200_vmm_padded_size", "b200_vmm_create", "b200_vmm_export",
                 "b200_vmm_import", "b200_vmm_release", "b200_mc_create", "b200_mc_import",
                 "b200_mc_add_device", "b200_mc_bind_map", "b200_close_fd"):
        getattr(L, name).restype = C.c_int
    for name in ("b200_symm_alloc", "b200_symm_open", "b200_symm_close", "b200_symm_free",
                 "b200_allreduce_optimizer", "b200_allreduce"):
        getattr(L, name).restype = C.c_int
