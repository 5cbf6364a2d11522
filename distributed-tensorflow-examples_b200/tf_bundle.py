"""TensorFlow V2 checkpoint bundles (``.index`` string table + ``.data-00000-of-00001``).

``tf.train.Saver()`` (``encoder:138``, ``gan:160``, ``lstm:101``) stores every variable in this
format; ``Saver(bundle_format="tf")`` / ``--checkpoint_format=tf`` makes this framework write it
too, so the files carry the reference's on-disk layout as well as its variable names (SURVEY
§2.4, §5.4).  The table builder, the reader and CRC32C are native (``csrc/tf_bundle.cpp`` in
``libpsb200_ps.so``); this module is the ctypes face.  TensorFlow is not available in this
image: the format is checked by our own reader and by byte-level tests, not by TF.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List

import numpy as np
import torch

from .control.client import ps_lib

# TensorFlow DataType enum values for the dtypes a checkpoint here can hold
_TF_DTYPE = {torch.float32: 1, torch.float64: 2, torch.int32: 3, torch.int64: 9, torch.bfloat16: 14, torch.float16: 19}
_NP_OF_TF = {1: np.float32, 2: np.float64, 3: np.int32, 9: np.int64, 19: np.float16}
_signatures_done = False


def _lib():
    global _signatures_done
    lib = ps_lib()
    if not _signatures_done:
        c = ctypes
        lib.psb_bundle_error.restype = c.c_char_p
        lib.psb_crc32c.restype = c.c_uint32
        lib.psb_crc32c.argtypes = [c.c_void_p, c.c_uint64]
        lib.psb_crc32c_sw.restype = c.c_uint32
        lib.psb_crc32c_sw.argtypes = [c.c_void_p, c.c_uint64]
        lib.psb_crc32c_mask.restype = c.c_uint32
        lib.psb_crc32c_mask.argtypes = [c.c_uint32]
        lib.psb_bundle_writer_open.restype = c.c_void_p
        lib.psb_bundle_writer_open.argtypes = [c.c_char_p]
        lib.psb_bundle_writer_add.restype = c.c_int
        lib.psb_bundle_writer_add.argtypes = [c.c_void_p, c.c_char_p, c.c_int, c.c_int, c.POINTER(c.c_int64),
                                              c.c_void_p, c.c_uint64]
        lib.psb_bundle_writer_finish.restype = c.c_int
        lib.psb_bundle_writer_finish.argtypes = [c.c_void_p]
        lib.psb_bundle_reader_open.restype = c.c_void_p
        lib.psb_bundle_reader_open.argtypes = [c.c_char_p]
        lib.psb_bundle_reader_count.restype = c.c_int
        lib.psb_bundle_reader_count.argtypes = [c.c_void_p]
        lib.psb_bundle_reader_entry.restype = c.c_int
        lib.psb_bundle_reader_entry.argtypes = [c.c_void_p, c.c_int, c.c_char_p, c.c_int, c.POINTER(c.c_int),
                                                c.POINTER(c.c_int), c.POINTER(c.c_int64), c.POINTER(c.c_int64)]
        lib.psb_bundle_reader_read.restype = c.c_int
        lib.psb_bundle_reader_read.argtypes = [c.c_void_p, c.c_int, c.c_void_p, c.c_uint64]
        lib.psb_bundle_reader_close.restype = None
        lib.psb_bundle_reader_close.argtypes = [c.c_void_p]
        _signatures_done = True
    return lib


def _err(lib) -> str:
    return (lib.psb_bundle_error() or b"").decode()


def crc32c(data: bytes, software: bool = False) -> int:
    """CRC-32C (Castagnoli) of ``data``; ``software=True`` forces the table implementation."""
    lib = _lib()
    buf = ctypes.create_string_buffer(data, len(data)) if data else None
    f = lib.psb_crc32c_sw if software else lib.psb_crc32c
    return int(f(ctypes.cast(buf, ctypes.c_void_p) if buf is not None else None, len(data)))


def masked_crc32c(data: bytes) -> int:
    """The rotated+offset form in which the table and the bundle entries store checksums."""
    return int(_lib().psb_crc32c_mask(crc32c(data)))


def is_tf_bundle(prefix: str) -> bool:
    """True if ``<prefix>.index`` ends with the string-table magic number."""
    try:
        with open(prefix + ".index", "rb") as f:
            f.seek(-8, 2)
            return f.read(8) == (0xdb4775248b80fb57).to_bytes(8, "little")
    except OSError:
        return False


def write_bundle(prefix: str, values: Dict[str, torch.Tensor]) -> None:
    """Write ``<prefix>.index`` and ``<prefix>.data-00000-of-00001`` holding ``values``."""
    lib = _lib()
    w = lib.psb_bundle_writer_open(prefix.encode())
    if not w:
        raise OSError(_err(lib))
    try:
        for name in sorted(values):
            t = values[name].detach().cpu().contiguous()
            if t.dtype not in _TF_DTYPE:
                raise TypeError("%s: dtype %s has no TensorFlow counterpart here" % (name, t.dtype))
            dims = (ctypes.c_int64 * max(1, t.dim()))(*t.shape)
            raw = t.view(torch.int16).numpy() if t.dtype == torch.bfloat16 else t.numpy()
            if lib.psb_bundle_writer_add(w, name.encode(), _TF_DTYPE[t.dtype], t.dim(), dims,
                                         raw.ctypes.data if raw.size else None, raw.nbytes) != 0:
                raise OSError(_err(lib))
    except BaseException:
        lib.psb_bundle_writer_finish(w)
        raise
    if lib.psb_bundle_writer_finish(w) != 0:
        raise OSError(_err(lib))


class _Reader:
    def __init__(self, prefix: str):
        self.lib = _lib()
        self.h = self.lib.psb_bundle_reader_open(prefix.encode())
        if not self.h:
            raise OSError("%s: %s" % (prefix, _err(self.lib)))

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.lib.psb_bundle_reader_close(self.h)
        self.h = None

    def entries(self):
        name = ctypes.create_string_buffer(1024)
        dtype, ndim, nbytes = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
        dims = (ctypes.c_int64 * 8)()
        for i in range(self.lib.psb_bundle_reader_count(self.h)):
            if self.lib.psb_bundle_reader_entry(self.h, i, name, 1024, ctypes.byref(dtype), ctypes.byref(ndim),
                                                dims, ctypes.byref(nbytes)) != 0:
                raise OSError(_err(self.lib))
            yield i, name.value.decode(), dtype.value, [int(dims[k]) for k in range(ndim.value)], int(nbytes.value)

    def read(self, i: int, dtype: int, shape: List[int], nbytes: int) -> torch.Tensor:
        raw = np.empty(nbytes, dtype=np.uint8)
        if self.lib.psb_bundle_reader_read(self.h, i, raw.ctypes.data if nbytes else None, nbytes) != 0:
            raise OSError(_err(self.lib))
        if dtype == 14:
            return torch.from_numpy(raw.view(np.int16).reshape(shape).copy()).view(torch.bfloat16)
        if dtype not in _NP_OF_TF:
            raise TypeError("TensorFlow dtype %d is not supported" % dtype)
        return torch.from_numpy(raw.view(_NP_OF_TF[dtype]).reshape(shape).copy())


def read_bundle(prefix: str) -> Dict[str, torch.Tensor]:
    """All tensors of a bundle; block and tensor checksums are verified on the way."""
    with _Reader(prefix) as r:
        return {name: r.read(i, dt, shape, nb) for i, name, dt, shape, nb in r.entries()}


def list_variables(prefix: str) -> Dict[str, List[int]]:
    with _Reader(prefix) as r:
        return {name: shape for _, name, _, shape, _ in r.entries()}
