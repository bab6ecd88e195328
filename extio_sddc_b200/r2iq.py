"""Host-side mirror of the reference's ``r2iqControlClass`` / ``fft_mt_r2iq`` operator
(Core/r2iq.h:16-48, Core/fft_mt_r2iq.h:21-118) over the CUDA backend's C ABI.

Method names, argument meaning and call order are the reference's: ``Init`` once, then per start
``setDecimate`` + ``TurnOn``; ``setFreqOffset`` / ``updateRand`` at any time; ``TurnOff``.  The
reference moves samples through two ring buffers owned by its caller; here the worker-loop body is
exposed directly as :meth:`process` (host numpy arrays) and :meth:`process_device` (device
pointers on a CUDA stream) -- the ring-buffer adapter itself is C++ (extio_sddc_b200/host).

All sample arithmetic runs in the sm_100a kernels behind ``libr2iq_cuda.so``; nothing here computes
on samples, and nothing falls back to the CPU.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi

NDECIDX = capi.NDECIDX
TRANSFER_SAMPLES = capi.TRANSFER_SAMPLES
HALF_FFT = capi.HALF_FFT


class CudaR2IQ:
    """``fft_mt_r2iq`` on a B200."""

    def __init__(self, device: int = -1, chunk_blocks: int = 0, fftn: int = 2048):
        self._lib = capi.load()
        self._h = ctypes.c_void_p(None)
        self._device = device
        self._chunk_blocks = chunk_blocks
        self.fftn = int(fftn)              # FFTN_R_ADC: 2048 (north star) or 8192 (reference as shipped)
        self.half_fft = self.fftn // 2

    # -- lifecycle ---------------------------------------------------------------------------
    def Init(self, gain: float, input=None, obuffers=None) -> None:  # noqa: N802 (reference naming)
        """fft_mt_r2iq::Init (Core/fft_mt_r2iq.cpp:147-227).  ``input`` / ``obuffers`` are the ring
        buffers of the reference signature; the Python mirror has none and ignores them."""
        if self._h:
            self.close()
        cfg = capi.Config(self._device, float(gain), int(self._chunk_blocks), self.fftn)
        h = ctypes.c_void_p(None)
        capi.check(self._lib.r2iq_cuda_create(ctypes.byref(cfg), ctypes.byref(h)))
        self._h = h

    def close(self) -> None:
        if self._h:
            self._lib.r2iq_cuda_destroy(self._h)
            self._h = ctypes.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _handle(self):
        if not self._h:
            raise RuntimeError("Init() has not been called")
        return self._h

    # -- r2iqControlClass ---------------------------------------------------------------------
    def setDecimate(self, dec: int) -> None:  # noqa: N802
        capi.check(self._lib.r2iq_cuda_set_decimate(self._handle(), int(dec)))

    def getRatio(self) -> int:  # noqa: N802
        return self._lib.r2iq_cuda_get_ratio(self._handle())

    def updateRand(self, v: bool) -> None:  # noqa: N802
        capi.check(self._lib.r2iq_cuda_set_rand(self._handle(), int(bool(v))))

    def getRand(self) -> bool:  # noqa: N802
        return bool(self._lib.r2iq_cuda_get_rand(self._handle()))

    def setSideband(self, lsb: bool) -> None:  # noqa: N802
        capi.check(self._lib.r2iq_cuda_set_sideband(self._handle(), int(bool(lsb))))

    def getSideband(self) -> bool:  # noqa: N802
        return bool(self._lib.r2iq_cuda_get_sideband(self._handle()))

    def TurnOn(self) -> None:  # noqa: N802
        capi.check(self._lib.r2iq_cuda_turn_on(self._handle()))

    def TurnOff(self) -> None:  # noqa: N802
        capi.check(self._lib.r2iq_cuda_turn_off(self._handle()))

    def IsOn(self) -> bool:  # noqa: N802
        return bool(self._lib.r2iq_cuda_is_on(self._handle()))

    def DataReady(self) -> None:  # noqa: N802
        """Core/r2iq.h:35 -- a no-op in the reference as well."""

    def setFreqOffset(self, offset: float) -> float:  # noqa: N802
        """fft_mt_r2iq::setFreqOffset (Core/fft_mt_r2iq.cpp:101-109); returns the fine-tune residual."""
        return float(self._lib.r2iq_cuda_set_freq_offset(self._handle(), float(offset)))

    def setFineTune(self, fc: float) -> float:  # noqa: N802
        """The residual mixer RadioHandlerClass applies after the operator (RadioHandler.cpp:33-37,292-297),
        run on the device instead: ``fc`` is what :meth:`setFreqOffset` returned; 0 switches it off."""
        return float(self._lib.r2iq_cuda_set_fine_tune(self._handle(), float(fc)))

    def setFineTuneMode(self, mode: int) -> None:  # noqa: N802
        """``capi.NCO_REFERENCE`` (bit-identical float32 recurrence, host-bound) or ``capi.NCO_DIRECT``
        (phase computed on the device, mixer free in the kernel's store epilogue, own tolerance)."""
        capi.check(self._lib.r2iq_cuda_set_fine_tune_mode(self._handle(), int(mode)))

    # -- extras the reference keeps private ---------------------------------------------------
    @property
    def tunebin(self) -> int:
        return self._lib.r2iq_cuda_get_tunebin(self._handle())

    @tunebin.setter
    def tunebin(self, tb: int) -> None:
        capi.check(self._lib.r2iq_cuda_set_tunebin(self._handle(), int(tb)))

    def filter_hw(self, d: int) -> np.ndarray:
        out = np.zeros(2 * self.half_fft, dtype=np.float32)
        capi.check(self._lib.r2iq_cuda_get_filter(self._handle(), int(d), out.ctypes.data, self.half_fft))
        return out.view(np.complex64)

    def set_filter_hw(self, d: int, hw: np.ndarray) -> None:
        hw = np.ascontiguousarray(hw, dtype=np.complex64)
        capi.check(self._lib.r2iq_cuda_set_filter(self._handle(), int(d), hw.ctypes.data, hw.shape[0]))

    def out_per_block(self) -> int:
        return int(self._lib.r2iq_cuda_out_per_block(self._handle()))

    def stats(self) -> capi.Stats:
        st = capi.Stats()
        capi.check(self._lib.r2iq_cuda_get_stats(self._handle(), ctypes.byref(st)))
        return st

    def kernel_name(self, d: int) -> str:
        buf = ctypes.create_string_buffer(128)
        capi.check(self._lib.r2iq_cuda_kernel_name(self._handle(), int(d), buf, 128))
        return buf.value.decode()

    def sync(self) -> None:
        capi.check(self._lib.r2iq_cuda_sync(self._handle()))

    # -- the worker-loop body -----------------------------------------------------------------
    def process(self, blocks: np.ndarray, history: np.ndarray | None = None, out: np.ndarray | None = None) -> np.ndarray:
        """Core/fft_mt_r2iq_impl.hpp:15-151 for ``len(blocks)/65536`` blocks in host memory.
        Returns complex64 of ``n_blocks * (32768 >> d)`` samples (``out`` may supply the buffer)."""
        blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
        if blocks.shape[0] % TRANSFER_SAMPLES:
            raise ValueError("input must be a whole number of 65536-sample blocks")
        n_blocks = blocks.shape[0] // TRANSFER_SAMPLES
        n_out = n_blocks * self.out_per_block()
        if out is None:
            out = np.empty(n_out, dtype=np.complex64)
        elif out.dtype != np.complex64 or out.size != n_out or not out.flags.c_contiguous:
            raise ValueError("out must be a contiguous complex64 array of n_blocks*(32768>>d) samples")
        hist_p = None
        if history is not None:
            history = np.ascontiguousarray(history, dtype=np.int16)
            if history.shape[0] != self.half_fft:
                raise ValueError(f"history must hold {self.half_fft} samples")
            hist_p = history.ctypes.data
        capi.check(self._lib.r2iq_cuda_process(self._handle(), blocks.ctypes.data, n_blocks, hist_p,
                                               out.ctypes.data if n_out else None))
        return out

    # -- submissions in flight (r2iq_cuda_process_async & co.) --------------------------------------
    def process_async(self, blocks: np.ndarray, out: np.ndarray | None = None, history: np.ndarray | None = None) -> int:
        """Enqueue one submission (<= chunk_blocks blocks) and return its ticket.  ``out`` None keeps the result on the
        device (fetch with :meth:`read`, then :meth:`release`).  The caller keeps ``blocks`` alive until
        :meth:`wait_input`, ``out`` until :meth:`wait`."""
        if blocks.dtype != np.int16 or not blocks.flags.c_contiguous or blocks.size % TRANSFER_SAMPLES:
            raise ValueError("blocks must be contiguous int16, a whole number of 65536-sample blocks")
        n_blocks = blocks.size // TRANSFER_SAMPLES
        if out is not None and (out.dtype != np.complex64 or out.size != n_blocks * self.out_per_block() or not out.flags.c_contiguous):
            raise ValueError("out must be a contiguous complex64 array of n_blocks*(32768>>d) samples")
        hist_p = None
        if history is not None:
            history = np.ascontiguousarray(history, dtype=np.int16)
            hist_p = history.ctypes.data
        ticket = ctypes.c_uint64(0)
        capi.check(self._lib.r2iq_cuda_process_async(self._handle(), blocks.ctypes.data, n_blocks, hist_p,
                                                     out.ctypes.data if out is not None else None, ctypes.byref(ticket)))
        return int(ticket.value)

    def wait_input(self, ticket: int) -> None:
        capi.check(self._lib.r2iq_cuda_wait_input(self._handle(), ctypes.c_uint64(ticket)))

    def wait(self, ticket: int) -> None:
        capi.check(self._lib.r2iq_cuda_wait(self._handle(), ctypes.c_uint64(ticket)))

    def read(self, ticket: int, first: int, out: np.ndarray) -> np.ndarray:
        capi.check(self._lib.r2iq_cuda_read(self._handle(), ctypes.c_uint64(ticket), int(first), int(out.size), out.ctypes.data))
        return out

    def release(self, ticket: int) -> None:
        capi.check(self._lib.r2iq_cuda_release(self._handle(), ctypes.c_uint64(ticket)))

    def process_device(self, d_in_ptr: int, n_blocks: int, d_out_ptr: int, stream_ptr: int = 0) -> None:
        """Device-resident variant: ``d_in_ptr`` -> [1024 history samples | n_blocks blocks] int16,
        ``d_out_ptr`` -> n_blocks*(32768>>d) complex64, launched on CUDA stream ``stream_ptr``."""
        capi.check(self._lib.r2iq_cuda_process_device(self._handle(), ctypes.c_void_p(d_in_ptr), int(n_blocks),
                                                      ctypes.c_void_p(d_out_ptr), ctypes.c_void_p(stream_ptr)))

    def process_device_static(self, d_in_ptr: int, n_blocks: int, d_out_ptr: int, stream_ptr: int = 0) -> None:
        """``process_device`` without any allocation (``r2iq_cuda_process_device_static``; no counterpart in the
        reference): work counter and tables live in the library's static device array.  FFTN 2048, at most 8192
        blocks per call; calls that overlap on the GPU race (include/r2iq_cuda.h)."""
        capi.check(self._lib.r2iq_cuda_process_device_static(self._handle(), ctypes.c_void_p(d_in_ptr), int(n_blocks),
                                                             ctypes.c_void_p(d_out_ptr), ctypes.c_void_p(stream_ptr)))

    def process_rates(self, blocks: np.ndarray, decimations, history: np.ndarray | None = None, outs=None) -> list:
        """One ADC stream at several output rates (``r2iq_cuda_process_rates``): the capture crosses PCIe once,
        every entry of ``decimations`` (indices 0..6) gets the stream ``process`` would give at that decimation.
        Returns a list of complex64 arrays (``outs`` may supply them, e.g. pinned buffers)."""
        blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
        if blocks.shape[0] % TRANSFER_SAMPLES:
            raise ValueError("input must be a whole number of 65536-sample blocks")
        n_blocks = blocks.shape[0] // TRANSFER_SAMPLES
        dec = np.ascontiguousarray(decimations, dtype=np.int32)
        sizes = [n_blocks * ((TRANSFER_SAMPLES // 2) >> int(d)) for d in dec]
        if outs is None:
            outs = [np.empty(n, dtype=np.complex64) for n in sizes]
        for o, n in zip(outs, sizes):
            if o.dtype != np.complex64 or o.size != n or not o.flags.c_contiguous:
                raise ValueError("outs[r] must be a contiguous complex64 array of n_blocks*(32768>>d) samples")
        ptrs = (ctypes.c_void_p * len(outs))(*[o.ctypes.data for o in outs])
        hist_p = None
        if history is not None:
            history = np.ascontiguousarray(history, dtype=np.int16)
            if history.shape[0] != self.half_fft:
                raise ValueError(f"history must hold {self.half_fft} samples")
            hist_p = history.ctypes.data
        capi.check(self._lib.r2iq_cuda_process_rates(self._handle(), blocks.ctypes.data, n_blocks, hist_p,
                                                     dec.ctypes.data, dec.shape[0], ptrs))
        return list(outs)

    def process_channels(self, blocks: np.ndarray, tunebins, history: np.ndarray | None = None) -> np.ndarray:
        """Channelizer: every tune bin in ``tunebins`` from one pass over ``blocks`` (any latched d, either FFTN).
        Returns complex64 [n_channels, n_blocks*(32768>>d)]; row c equals a separate operator tuned to
        ``tunebins[c]``."""
        blocks = np.ascontiguousarray(blocks, dtype=np.int16).reshape(-1)
        if blocks.shape[0] % TRANSFER_SAMPLES:
            raise ValueError("input must be a whole number of 65536-sample blocks")
        n_blocks = blocks.shape[0] // TRANSFER_SAMPLES
        tb = np.ascontiguousarray(tunebins, dtype=np.int32)
        out = np.empty((tb.shape[0], n_blocks * self.out_per_block()), dtype=np.complex64)
        hist_p = None
        if history is not None:
            history = np.ascontiguousarray(history, dtype=np.int16)
            hist_p = history.ctypes.data
        capi.check(self._lib.r2iq_cuda_process_channels(self._handle(), blocks.ctypes.data, n_blocks, hist_p,
                                                        tb.ctypes.data, tb.shape[0], out.ctypes.data))
        return out

    def process_channels_device(self, d_in_ptr: int, n_blocks: int, tunebins, d_out_ptr: int, stream_ptr: int = 0) -> None:
        tb = np.ascontiguousarray(tunebins, dtype=np.int32)
        capi.check(self._lib.r2iq_cuda_process_channels_device(self._handle(), ctypes.c_void_p(d_in_ptr), int(n_blocks),
                                                               tb.ctypes.data, tb.shape[0], ctypes.c_void_p(d_out_ptr),
                                                               ctypes.c_void_p(stream_ptr)))

    def convert_float(self, x: np.ndarray, rand: bool) -> np.ndarray:
        """convert_float<rand> (Core/fft_mt_r2iq.h:36-51) through the device code path."""
        x = np.ascontiguousarray(x, dtype=np.int16).reshape(-1)
        out = np.empty(x.shape[0], dtype=np.float32)
        capi.check(self._lib.r2iq_cuda_convert(self._handle(), x.ctypes.data, x.shape[0], int(bool(rand)), out.ctypes.data))
        return out


def design_filter(gain: float, d: int) -> np.ndarray:
    """Host-side filter design of Init for one decimation (device-free): halfFft complex64 bins."""
    out = np.zeros(2 * HALF_FFT, dtype=np.float32)
    capi.check(capi.load().r2iq_cuda_design_filter(float(gain), int(d), out.ctypes.data, HALF_FFT))
    return out.view(np.complex64)


def kaiser_taps(num_taps: int, astop_db: float, f_pass: float, f_stop: float) -> np.ndarray:
    out = np.zeros(int(num_taps), dtype=np.float32)
    capi.check(capi.load().r2iq_cuda_kaiser(int(num_taps), float(astop_db), float(f_pass), float(f_stop), out.ctypes.data))
    return out


def freq_to_tunebin(offset: float, d: int):
    """(tunebin, residual) of setFreqOffset without touching an operator (device-free)."""
    tb = ctypes.c_int(0)
    r = capi.load().r2iq_cuda_freq_to_tunebin(float(offset), int(d), ctypes.byref(tb))
    return tb.value, float(r)


class PinnedBuffer:
    """Page-locked host buffer (r2iq_cuda_host_alloc) exposed as a numpy array, so the H2D/D2H
    copies inside process() run asynchronously and overlap with the kernels."""

    def __init__(self, n: int, dtype, device: int = -1):
        # device: the GPU whose copy engines will use the buffer; its pages go to that GPU's NUMA node
        self._lib = capi.load()
        dt = np.dtype(dtype)
        self.nbytes = int(n) * dt.itemsize
        self._ptr = self._lib.r2iq_cuda_host_alloc_on(self.nbytes, int(device))
        if not self._ptr:
            raise capi.R2iqError(capi.E_NOMEM, capi.last_error())
        buf = (ctypes.c_char * self.nbytes).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=dt, count=int(n))

    def free(self) -> None:
        if self._ptr:
            self.array = None
            self._lib.r2iq_cuda_host_free(ctypes.c_void_p(self._ptr))
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
