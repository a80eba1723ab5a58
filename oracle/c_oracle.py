"""ctypes access to oracle/splice_oracle.c (TEST INFRASTRUCTURE)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "liboracle_splice.so")


def build():
    src = os.path.join(_HERE, "splice_oracle.c")
    if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", SO, src])
    return SO


def load():
    lib = C.CDLL(build())
    lib.oracle_attach_positions.restype = C.c_int64
    lib.oracle_attach_positions.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    lib.oracle_embed_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64]
    lib.oracle_splice_rows.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
    return lib


def attach_positions(ids: np.ndarray, attach: int):
    lib = load()
    ids = np.ascontiguousarray(ids.reshape(-1), dtype=np.int64)
    dest = np.empty(ids.shape[0], np.int32)
    sop = np.empty(ids.shape[0], np.int32)
    k = lib.oracle_attach_positions(ids.ctypes.data, ids.shape[0], attach, dest.ctypes.data, sop.ctypes.data)
    return dest[:k].copy(), sop


def embed_splice(table: np.ndarray, ids: np.ndarray, src: np.ndarray, rows: np.ndarray) -> np.ndarray:
    """table/src/out as 2-D byte-compatible arrays of one dtype: out = table[ids]; out[rows[k]] = src[k]."""
    lib = load()
    ids = np.ascontiguousarray(ids.reshape(-1), dtype=np.int64)
    rb = table.strides[0]
    out = np.empty((ids.shape[0], table.shape[1]), dtype=table.dtype)
    assert lib.oracle_embed_rows(out.ctypes.data, table.ctypes.data, table.shape[0], ids.ctypes.data, ids.shape[0], rb) == 0
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    assert lib.oracle_splice_rows(out.ctypes.data, out.shape[0], src.ctypes.data, rows.ctypes.data, rows.shape[0], rb) == 0
    return out
