"""ctypes binding of ``lib/libpsb200_kernels.so`` (C ABI: ``include/psb200.h``).

The library is built in-tree by ``psb200.build`` (``__graft_entry__.build()``); on a GPU
machine a missing library is a hard error — there is no PyTorch fallback behind the
kernels.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_uint64, c_void_p

from .. import PACKAGE_DIR

LIB_PATH = os.path.join(PACKAGE_DIR, "lib", "libpsb200_kernels.so")


class Epilogue(Structure):
    """Mirror of ``psb_epilogue_t``."""
    _fields_ = [
        ("alpha", c_float), ("act", c_int), ("dact", c_int),
        ("bias", c_void_p),
        ("yprev", c_void_p), ("ld_yprev", c_int),
        ("out", c_void_p), ("ld_out", c_int),
        ("out_t", c_void_p), ("ld_out_t", c_int),
        ("out_f32", c_void_p), ("ld_f32", c_int),
        ("colsum", c_void_p), ("split_k", c_int), ("split_stride", ctypes.c_longlong), ("b_static", c_int),
        ("ready", c_void_p), ("ready_step", c_void_p), ("ready_rows", c_int), ("ready_dim", c_int),
        ("ready_nblk", c_int), ("ready_first", c_int), ("bias_ready", c_void_p), ("ready_status", c_void_p),
        ("ready_timeout_ns", c_uint64), ("pull_mode", c_int),
    ]


class KernelLibraryError(RuntimeError):
    pass


_lib = None


def _declare(lib) -> None:
    lib.psb_last_error.restype = c_char_p
    lib.psb_launch_count.restype = c_uint64
    lib.psb_version.restype = c_int
    lib.psb_gemm_bf16.restype = c_int
    lib.psb_gemm_bf16.argtypes = [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int,
                                  POINTER(Epilogue), c_int, c_void_p]
    lib.psb_gemm_bf16_pull.restype = c_int
    lib.psb_gemm_bf16_pull.argtypes = lib.psb_gemm_bf16.argtypes
    for name, sig in _EXTRA_SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise KernelLibraryError("%s is missing from %s — rebuild with `python -m psb200.build`"
                                     % (name, LIB_PATH))
        fn.restype = c_int
        fn.argtypes = sig


# name -> argtypes for entry points declared by other modules of this package
_EXTRA_SIGNATURES = {}


def register_signature(name: str, argtypes) -> None:
    _EXTRA_SIGNATURES[name] = list(argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype = c_int
        fn.argtypes = list(argtypes)


def lib():
    """Load (once) and return the kernel library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            # not built yet (fresh checkout): build in-tree now — nvcc is part of the image; if that
            # fails there is no fallback path, the error below is final
            try:
                from ..build import build_kernels
                build_kernels()
            except Exception as e:
                raise KernelLibraryError(
                    "kernel library %s not found and building it failed (%s) — run "
                    "`python -m psb200.build`" % (LIB_PATH, e))
        handle = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        _declare(handle)
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().psb_last_error().decode("utf-8", "replace")
        raise KernelLibraryError("%s failed (%d): %s" % (what or "kernel call", rc, msg))


def launch_count() -> int:
    return int(lib().psb_launch_count())


def current_stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
