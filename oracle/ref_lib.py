"""oracle/ref_lib.py -- TEST INFRASTRUCTURE, not product code.

ctypes access to ``oracle/_ref/libref_r2iq_<FFTN>.so``: the UNMODIFIED reference operator
(``Core/fft_mt_r2iq*.cpp`` + ``Core/fir.cpp`` compiled in place by ``oracle/Makefile``) behind the
harness in ``oracle/ref_driver.cpp``.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs may import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
REFERENCE_ROOT = "/root/reference"

BACKEND_DOUBLE = 0
BACKEND_MKL = 1
VARIANTS = {"dispatch": 0, "def": 1, "avx": 2, "avx2": 3, "avx512": 4}

_libs: dict[int, "RefLib"] = {}


# RadioHandlerClass-level drop-in check (oracle/ref_radio_driver.cpp), run by tests/test_host_adapter.py on the GPU
RADIO_DROPIN = os.path.join(REF_DIR, "ref_radio_dropin")


def lib_path(fftn: int) -> str:
    return os.path.join(REF_DIR, f"libref_r2iq_{fftn}.so")


def build(force: bool = False) -> bool:
    """Run ``make ref`` when the reference tree is present (this container).  On the GPU box the
    prebuilt libraries travel with the snapshot and /root/reference does not exist."""
    have = all(os.path.exists(lib_path(n)) for n in (2048, 8192)) and mixer_available() and os.path.exists(RADIO_DROPIN)
    if os.path.isdir(os.path.join(REFERENCE_ROOT, "Core")):
        # always: make is incremental, and ref_radio_dropin links the adapter sources of this repo
        subprocess.run(["make", "-C", _HERE, "ref"] + (["-B"] if force else []), check=True, stdout=subprocess.DEVNULL)
        have = True
    return have


def available(fftn: int = 2048) -> bool:
    return os.path.exists(lib_path(fftn))


def mkl_library_path() -> str | None:
    """torch's libtorch_cpu.so exports the MKL DFTI entry points the shim's MKL backend binds."""
    try:
        import torch  # noqa: F401  (only to locate the file; no tensors are created)
        p = os.path.join(os.path.dirname(torch.__file__), "lib", "libtorch_cpu.so")
        return p if os.path.exists(p) else None
    except Exception:
        return None


class RefLib:
    def __init__(self, fftn: int):
        self.fftn = fftn
        self.lib = ctypes.CDLL(lib_path(fftn))
        L = self.lib
        L.refr2iq_fftn.restype = ctypes.c_int
        L.refr2iq_half_fft.restype = ctypes.c_int
        L.refr2iq_fft_per_buf.restype = ctypes.c_int
        L.refr2iq_max_threads.restype = ctypes.c_int
        L.refr2iq_set_fft_backend.argtypes = [ctypes.c_int, ctypes.c_char_p]
        L.refr2iq_set_fft_backend.restype = ctypes.c_int
        L.refr2iq_get_fft_backend.restype = ctypes.c_int
        L.refr2iq_kaiser.argtypes = [ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_void_p]
        L.refr2iq_kaiser.restype = ctypes.c_int
        L.refr2iq_filter.argtypes = [ctypes.c_float, ctypes.c_int, ctypes.c_void_p]
        L.refr2iq_filter.restype = ctypes.c_int
        L.refr2iq_set_freq_offset.argtypes = [ctypes.c_float, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.refr2iq_set_freq_offset.restype = ctypes.c_float
        L.refr2iq_convert.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.refr2iq_convert.restype = None
        L.refr2iq_run.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_float, ctypes.c_int,
                                  ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.refr2iq_run.restype = ctypes.c_long
        L.refr2iq_bench.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_long, ctypes.c_long, ctypes.c_float,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        L.refr2iq_bench.restype = ctypes.c_double
        assert L.refr2iq_fftn() == fftn

    # -- controls --------------------------------------------------------------------------
    def set_fft_backend(self, backend: int) -> int:
        path = mkl_library_path() if backend == BACKEND_MKL else None
        if backend == BACKEND_MKL and path is None:
            return -2
        return self.lib.refr2iq_set_fft_backend(backend, path.encode() if path else None)

    def fft_backend(self) -> int:
        return self.lib.refr2iq_get_fft_backend()

    # -- pieces of the path ----------------------------------------------------------------
    def kaiser(self, num_taps: int, astop: float, fpass: float, fstop: float) -> np.ndarray:
        out = np.zeros(num_taps, dtype=np.float32)
        self.lib.refr2iq_kaiser(num_taps, astop, fpass, fstop, out.ctypes.data)
        return out

    def filter_hw(self, gain: float, d: int) -> np.ndarray:
        hf = self.fftn // 2
        out = np.zeros(2 * hf, dtype=np.float32)
        n = self.lib.refr2iq_filter(gain, d, out.ctypes.data)
        assert n == hf
        return out.view(np.complex64)

    def set_freq_offset(self, offset: float, d: int):
        tb = ctypes.c_int(0)
        r = self.lib.refr2iq_set_freq_offset(offset, d, ctypes.byref(tb))
        return tb.value, float(r)

    def convert(self, x: np.ndarray, rand: bool) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.int16)
        out = np.zeros(x.shape[0], dtype=np.float32)
        self.lib.refr2iq_convert(x.ctypes.data, out.ctypes.data, x.shape[0], int(rand))
        return out

    def run(self, blocks: np.ndarray, d: int, tunebin: int, *, gain: float = 1.0, rand: bool = False,
            sideband: bool = False, history: np.ndarray | None = None, variant: str = "dispatch") -> np.ndarray:
        blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
        assert blocks.shape[0] % 65536 == 0
        n_blocks = blocks.shape[0] // 65536
        out = np.zeros(2 * n_blocks * (32768 >> d), dtype=np.float32)
        hist_p = None
        if history is not None:
            history = np.ascontiguousarray(history, dtype=np.int16)
            assert history.shape[0] == self.fftn // 2
            hist_p = history.ctypes.data
        n = self.lib.refr2iq_run(blocks.ctypes.data, n_blocks, hist_p, gain, d, tunebin, int(rand), int(sideband),
                                 VARIANTS[variant], out.ctypes.data)
        assert n == n_blocks * (32768 >> d), n
        return out.view(np.complex64)

    def bench(self, blocks: np.ndarray, d: int, tunebin: int, *, warm_blocks: int, timed_blocks: int,
              gain: float = 1.0, rand: bool = False, sideband: bool = False, variant: str = "dispatch",
              instances: int = 1):
        """Returns (input samples per second aggregate, seconds)."""
        blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
        n_distinct = blocks.shape[0] // 65536
        sec = ctypes.c_double(0.0)
        sps = self.lib.refr2iq_bench(blocks.ctypes.data, n_distinct, warm_blocks, timed_blocks, gain, d, tunebin,
                                     int(rand), int(sideband), VARIANTS[variant], instances, ctypes.byref(sec))
        return float(sps), float(sec.value)


def mixer_available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "libref_mixer.so"))


def mixer_apply(iq: np.ndarray, fc: float, slot: int = 32768) -> np.ndarray:
    """The reference's fine-tune mixer over a complex64 stream, slot by slot with carried state, as
    RadioHandlerClass::OnDataPacket does (Core/RadioHandler.cpp:28-37)."""
    lib = ctypes.CDLL(os.path.join(REF_DIR, "libref_mixer.so"))
    lib.refmix_apply.argtypes = [ctypes.c_float, ctypes.c_void_p, ctypes.c_long, ctypes.c_long]
    lib.refmix_apply.restype = None
    out = np.array(iq, dtype=np.complex64, copy=True)
    lib.refmix_apply(fc, out.ctypes.data, out.shape[0], slot)
    return out


def get(fftn: int = 2048) -> RefLib:
    if fftn not in _libs:
        if not available(fftn):
            build()
        _libs[fftn] = RefLib(fftn)
    return _libs[fftn]
