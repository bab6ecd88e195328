"""oracle/c_lib.py -- TEST INFRASTRUCTURE.  ctypes access to the plain-C restatement
(oracle/r2iq_oracle.c -> oracle/_ref/libr2iq_oracle_c.so)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libr2iq_oracle_c.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "r2iq_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "port"], check=True, stdout=subprocess.DEVNULL)
    return LIB


def get():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
        _lib.oracle_r2iq.restype = ctypes.c_long
        _lib.oracle_r2iq.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_float,
                                     ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_design_filter.restype = ctypes.c_int
        _lib.oracle_design_filter.argtypes = [ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        _lib.oracle_convert.restype = None
        _lib.oracle_convert.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_int]
    return _lib


def r2iq(blocks, d, tunebin, *, gain=1.0, rand=False, sideband=False, fftn=2048, history=None) -> np.ndarray:
    L = get()
    blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
    nb = blocks.shape[0] // 65536
    out = np.zeros(2 * nb * (32768 >> d), dtype=np.float32)
    hp = None
    if history is not None:
        history = np.ascontiguousarray(history, dtype=np.int16)
        hp = history.ctypes.data
    n = L.oracle_r2iq(blocks.ctypes.data, nb, hp, None, gain, d, tunebin, int(rand), int(sideband), fftn, out.ctypes.data)
    assert n == nb * (32768 >> d)
    return out.view(np.complex64)


def design_filter(gain, d, fftn=2048) -> np.ndarray:
    out = np.zeros(fftn, dtype=np.float32)
    assert get().oracle_design_filter(gain, d, fftn, out.ctypes.data) == fftn // 2
    return out.view(np.complex64)


def convert(x, rand) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.int16)
    out = np.zeros(x.shape[0], dtype=np.float32)
    get().oracle_convert(x.ctypes.data, out.ctypes.data, x.shape[0], int(rand))
    return out
