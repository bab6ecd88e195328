"""Python mirror of the reference's FFT backend abstraction (Core/fft_backend.h:22-50) over the batched CUDA transforms
of include/r2iq_cuda.h (`r2iq_cuda_fft_*`).  Same member names and argument meaning as `FFTBackend`; one plan runs
`batch` transforms per launch (the reference's backends run one).  FFTW conventions: unnormalised, forward =
exp(-2 pi i nk/n), r2c of n reals -> n/2 + 1 bins.  CUDA only: planning fails without a device."""
import ctypes

import numpy as np

from . import capi
from .r2iq import PinnedBuffer

FORWARD, BACKWARD = "forward", "backward"


class CudaFFTPlan:
    def __init__(self, handle, kind: int, n: int, batch: int):
        self.handle, self.kind, self.n, self.batch = handle, kind, n, batch

    @property
    def in_shape(self):
        return (self.batch, self.n)

    @property
    def out_shape(self):
        return (self.batch, self.n // 2 + 1 if self.kind == capi.FFT_R2C else self.n)


class CudaFFTBackend:
    """FFTBackend (Core/fft_backend.h:22) on one CUDA device."""

    def __init__(self, device: int = -1, batch: int = 1):
        self._lib = capi.load()
        self.device, self.batch = int(device), int(batch)
        self._buffers = {}

    def name(self) -> str:                                        # fft_backend.h:27
        return "CUDA sm_100a register FFT (batch %d)" % self.batch

    def _plan(self, kind: int, n: int) -> CudaFFTPlan:
        h = ctypes.c_void_p()
        capi.check(self._lib.r2iq_cuda_fft_plan(self.device, kind, int(n), self.batch, ctypes.byref(h)))
        return CudaFFTPlan(h, kind, int(n), self.batch)

    def plan_r2c(self, n: int) -> CudaFFTPlan:                    # fft_backend.h:30
        return self._plan(capi.FFT_R2C, n)

    def plan_c2c(self, n: int, direction: str) -> CudaFFTPlan:    # fft_backend.h:31
        if direction not in (FORWARD, BACKWARD):
            raise ValueError("direction must be 'forward' or 'backward'")
        return self._plan(capi.FFT_C2C_FORWARD if direction == FORWARD else capi.FFT_C2C_BACKWARD, n)

    def execute_r2c(self, plan: CudaFFTPlan, real_in: np.ndarray, complex_out: np.ndarray) -> None:   # fft_backend.h:34
        self._execute(plan, real_in, np.float32, complex_out)

    def execute_c2c(self, plan: CudaFFTPlan, complex_in: np.ndarray, complex_out: np.ndarray) -> None:   # fft_backend.h:35
        self._execute(plan, complex_in, np.complex64, complex_out)

    def _execute(self, plan, src, src_dtype, dst) -> None:
        if src.dtype != src_dtype or dst.dtype != np.complex64:
            raise TypeError("float32 / complex64 arrays expected")
        if src.size != plan.batch * plan.n or dst.size != int(np.prod(plan.out_shape)):
            raise ValueError("array sizes do not match the plan (batch %d, n %d)" % (plan.batch, plan.n))
        if not (src.flags.c_contiguous and dst.flags.c_contiguous):
            raise ValueError("contiguous arrays expected")
        capi.check(self._lib.r2iq_cuda_fft_execute(plan.handle, src.ctypes.data, dst.ctypes.data))

    def execute_device(self, plan: CudaFFTPlan, d_in: int, d_out: int, batch: int, stream: int = 0) -> None:
        """The same kernel on device pointers, asynchronous on `stream` (any batch)."""
        capi.check(self._lib.r2iq_cuda_fft_execute_device(plan.handle, ctypes.c_void_p(d_in), ctypes.c_void_p(d_out),
                                                          int(batch), ctypes.c_void_p(stream)))

    def destroy_plan(self, plan: CudaFFTPlan) -> None:            # fft_backend.h:38
        if plan.handle:
            self._lib.r2iq_cuda_fft_destroy(plan.handle)
            plan.handle = None

    def alloc(self, n: int, dtype) -> np.ndarray:                 # fft_backend.h:45: page-locked, device-addressable
        buf = PinnedBuffer(n, dtype, self.device)
        self._buffers[buf.array.ctypes.data] = buf
        return buf.array

    def free(self, array: np.ndarray) -> None:                    # fft_backend.h:46
        buf = self._buffers.pop(array.ctypes.data, None)
        if buf is not None:
            buf.free()
