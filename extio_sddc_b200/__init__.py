"""extio_sddc_b200 -- B200-native (sm_100a) backend for ExtIO_sddc's real->IQ down-converter.

Only what the ``fft_mt_r2iq`` hot path needs lives here:
  csrc/   hand-written CUDA kernels + the C-ABI shared library (include/r2iq_cuda.h)
  host/   C++ ``cuda_r2iq : r2iqControlClass`` adapter over the reference's ring-buffer contract
  capi.py / r2iq.py   ctypes binding and the Python mirror of the operator interface
"""
from .r2iq import CudaR2IQ, PinnedBuffer  # noqa: F401

__all__ = ["CudaR2IQ", "PinnedBuffer"]
