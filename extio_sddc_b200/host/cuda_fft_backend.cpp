// cuda_fft_backend.cpp -- see cuda_fft_backend.h.  No arithmetic here: every call forwards to the C ABI.
#include "cuda_fft_backend.h"

#include <cstdio>

const char* CudaFFTBackend::name() const {
    snprintf(name_, sizeof(name_), "CUDA sm_100a register FFT (batch %ld)", batch_);
    return name_;
}

FFTPlanHandle CudaFFTBackend::plan_r2c(int n, float*, fft_complex*) {
    r2iq_cuda_fft_t p = nullptr;
    if (r2iq_cuda_fft_plan(device_, R2IQ_FFT_R2C, n, batch_, &p) != R2IQ_OK) return nullptr;
    return static_cast<FFTPlanHandle>(p);
}

FFTPlanHandle CudaFFTBackend::plan_c2c(int n, fft_complex*, fft_complex*, FFTDirection dir) {
    r2iq_cuda_fft_t p = nullptr;
    const int kind = dir == FFTDirection::Forward ? R2IQ_FFT_C2C_FORWARD : R2IQ_FFT_C2C_BACKWARD;
    if (r2iq_cuda_fft_plan(device_, kind, n, batch_, &p) != R2IQ_OK) return nullptr;
    return static_cast<FFTPlanHandle>(p);
}

void CudaFFTBackend::execute_r2c(FFTPlanHandle plan, float* in, fft_complex* out) {
    if (r2iq_cuda_fft_execute(static_cast<r2iq_cuda_fft_t>(plan), in, out) != R2IQ_OK) failures_++;
}

void CudaFFTBackend::execute_c2c(FFTPlanHandle plan, fft_complex* in, fft_complex* out) {
    if (r2iq_cuda_fft_execute(static_cast<r2iq_cuda_fft_t>(plan), in, out) != R2IQ_OK) failures_++;
}

int CudaFFTBackend::execute_device(FFTPlanHandle plan, const void* d_in, void* d_out, long batch, void* stream) {
    return r2iq_cuda_fft_execute_device(static_cast<r2iq_cuda_fft_t>(plan), d_in, d_out, batch, stream);
}

void CudaFFTBackend::destroy_plan(FFTPlanHandle plan) { r2iq_cuda_fft_destroy(static_cast<r2iq_cuda_fft_t>(plan)); }

void* CudaFFTBackend::alloc(size_t bytes) { return r2iq_cuda_host_alloc_on(bytes, device_); }

void CudaFFTBackend::free(void* ptr) { r2iq_cuda_host_free(ptr); }

#ifdef R2IQ_FFT_BACKEND_FACTORY
// the compiled-in backend of a reference build that selects CUDA (Core/fft_backend.h:50)
FFTBackend* getFFTBackend() {
    static CudaFFTBackend instance;
    return &instance;
}
#endif
