"""ctypes binding of the C ABI in include/r2iq_cuda.h (the drop-in boundary).

Loads ``extio_sddc_b200/lib/libr2iq_cuda.so``.  There is no fallback of any kind: if the library
is missing or no sm_100 device is present the calls raise.
"""
from __future__ import annotations

import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libr2iq_cuda.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "r2iq_cuda.h")

TRANSFER_SAMPLES = 65536
NDECIDX = 7
FFTN = 2048
HALF_FFT = FFTN // 2

OK = 0
E_INVALID, E_CUDA, E_NOMEM, E_STATE, E_NODEVICE = -1, -2, -3, -4, -5
NCO_REFERENCE, NCO_DIRECT = 0, 1
FFT_R2C, FFT_C2C_FORWARD, FFT_C2C_BACKWARD = 0, 1, 2


class Config(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int), ("gain", ctypes.c_float), ("chunk_blocks", ctypes.c_int),
                ("fftn", ctypes.c_int)]


class Stats(ctypes.Structure):
    _fields_ = [("kernel_launches", ctypes.c_uint64), ("blocks_processed", ctypes.c_uint64),
                ("h2d_bytes", ctypes.c_uint64), ("d2h_bytes", ctypes.c_uint64),
                ("sm_count", ctypes.c_int), ("cc_major", ctypes.c_int), ("cc_minor", ctypes.c_int)]


class R2iqError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"r2iq_cuda error {code}: {msg}")
        self.code = code


_lib = None

_VP, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float
_SIGNATURES = {
    "r2iq_cuda_last_error": (ctypes.c_char_p, []),
    "r2iq_cuda_version": (ctypes.c_char_p, []),
    "r2iq_cuda_create": (_I, [ctypes.POINTER(Config), ctypes.POINTER(_VP)]),
    "r2iq_cuda_destroy": (_I, [_VP]),
    "r2iq_cuda_set_filter": (_I, [_VP, _I, _VP, _I]),
    "r2iq_cuda_get_filter": (_I, [_VP, _I, _VP, _I]),
    "r2iq_cuda_set_decimate": (_I, [_VP, _I]),
    "r2iq_cuda_get_decimate": (_I, [_VP]),
    "r2iq_cuda_get_ratio": (_I, [_VP]),
    "r2iq_cuda_set_freq_offset": (_F, [_VP, _F]),
    "r2iq_cuda_set_tunebin": (_I, [_VP, _I]),
    "r2iq_cuda_get_tunebin": (_I, [_VP]),
    "r2iq_cuda_set_rand": (_I, [_VP, _I]),
    "r2iq_cuda_get_rand": (_I, [_VP]),
    "r2iq_cuda_set_sideband": (_I, [_VP, _I]),
    "r2iq_cuda_get_sideband": (_I, [_VP]),
    "r2iq_cuda_turn_on": (_I, [_VP]),
    "r2iq_cuda_turn_off": (_I, [_VP]),
    "r2iq_cuda_is_on": (_I, [_VP]),
    "r2iq_cuda_out_per_block": (_L, [_VP]),
    "r2iq_cuda_process": (_I, [_VP, _VP, _L, _VP, _VP]),
    "r2iq_cuda_process_device": (_I, [_VP, _VP, _L, _VP, _VP]),
    "r2iq_cuda_process_device_static": (_I, [_VP, _VP, _L, _VP, _VP]),
    "r2iq_cuda_process_rates": (_I, [_VP, _VP, _L, _VP, _VP, _I, _VP]),
    "r2iq_cuda_process_channels": (_I, [_VP, _VP, _L, _VP, _VP, _I, _VP]),
    "r2iq_cuda_process_channels_device": (_I, [_VP, _VP, _L, _VP, _I, _VP, _VP]),
    "r2iq_cuda_set_fine_tune": (_F, [_VP, _F]),
    "r2iq_cuda_set_fine_tune_mode": (_I, [_VP, _I]),
    "r2iq_cuda_get_fine_tune_mode": (_I, [_VP]),
    "r2iq_cuda_process_async": (_I, [_VP, _VP, _L, _VP, _VP, ctypes.POINTER(ctypes.c_uint64)]),
    "r2iq_cuda_wait_input": (_I, [_VP, ctypes.c_uint64]),
    "r2iq_cuda_wait": (_I, [_VP, ctypes.c_uint64]),
    "r2iq_cuda_read": (_I, [_VP, ctypes.c_uint64, _L, _L, _VP]),
    "r2iq_cuda_release": (_I, [_VP, ctypes.c_uint64]),
    "r2iq_cuda_host_register": (_I, [_VP, ctypes.c_size_t]),
    "r2iq_cuda_host_unregister": (_I, [_VP]),
    "r2iq_cuda_sync": (_I, [_VP]),
    "r2iq_cuda_host_alloc": (_VP, [ctypes.c_size_t]),
    "r2iq_cuda_host_alloc_on": (_VP, [ctypes.c_size_t, _I]),
    "r2iq_cuda_device_numa_node": (_I, [_I]),
    "r2iq_cuda_host_free": (_I, [_VP]),
    "r2iq_cuda_device_alloc": (_VP, [_I, ctypes.c_size_t]),
    "r2iq_cuda_device_free": (_I, [_VP]),
    "r2iq_cuda_ipc_export": (_I, [_VP, _VP]),
    "r2iq_cuda_ipc_open": (_I, [_I, _VP, ctypes.POINTER(_VP)]),
    "r2iq_cuda_ipc_close": (_I, [_VP]),
    "r2iq_cuda_convert": (_I, [_VP, _VP, _L, _I, _VP]),
    "r2iq_cuda_get_stats": (_I, [_VP, ctypes.POINTER(Stats)]),
    "r2iq_cuda_get_fftn": (_I, [_VP]),
    "r2iq_cuda_kernel_name": (_I, [_VP, _I, ctypes.c_char_p, _I]),
    "r2iq_cuda_design_filter": (_I, [_F, _I, _VP, _I]),
    "r2iq_cuda_kaiser": (_I, [_I, _F, _F, _F, _VP]),
    "r2iq_cuda_freq_to_tunebin": (_F, [_F, _I, ctypes.POINTER(_I)]),
    "r2iq_cuda_fft_plan": (_I, [_I, _I, _I, _L, ctypes.POINTER(_VP)]),
    "r2iq_cuda_fft_execute": (_I, [_VP, _VP, _VP]),
    "r2iq_cuda_fft_execute_device": (_I, [_VP, _VP, _VP, _L, _VP]),
    "r2iq_cuda_fft_batch": (_L, [_VP]),
    "r2iq_cuda_fft_destroy": (_I, [_VP]),
}


def declared_symbols() -> list[str]:
    """Every entry point include/r2iq_cuda.h declares (parsed from the header itself)."""
    with open(HEADER_PATH) as f:
        text = f.read()
    return sorted(set(re.findall(r"R2IQ_API\s+[\w\s\*]+?\b(r2iq_cuda_\w+)\s*\(", text)))


def load() -> ctypes.CDLL:
    """Load the shared library and bind every declared symbol.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'`. "
            "The r2iq backend is CUDA-only: there is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().r2iq_cuda_last_error().decode(errors="replace")


def check(rc: int) -> None:
    if rc != OK:
        raise R2iqError(rc, last_error())
