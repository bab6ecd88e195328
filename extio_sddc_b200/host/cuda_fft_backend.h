// cuda_fft_backend.h -- `CudaFFTBackend : FFTBackend`, the reference's FFT backend abstraction
// (Core/fft_backend.h:22-50) over the batched CUDA transforms of include/r2iq_cuda.h (r2iq_cuda_fft_*).
//
// Inside the reference tree compile with -DR2IQ_USE_REFERENCE_HEADERS (the class then derives from the reference's
// own FFTBackend, and -DR2IQ_FFT_BACKEND_FACTORY makes this file's getFFTBackend() the compiled-in backend, in the
// place of Core/fft_backend_fftw.cpp / fft_backend_mkl.cpp: Core/CMakeLists.txt:19-37).  Stand-alone, the interface
// is restated below with the same members in the same order, so the vtable layout is the reference's.
//
// Contract differences, all additive:
//   * one plan executes `batch` transforms per execute_* call (set_batch() before planning; 1 = the reference's shape);
//   * alloc() returns page-locked host memory; buffers from alloc() of up to 256 KiB are transformed in place by the
//     kernel (no staging copies), anything else is staged through device memory;
//   * there is no CPU path: plan_* returns nullptr and last_error() says why when no sm_100 device is present.
#pragma once
#include <cstddef>

#include "r2iq_cuda.h"

#ifdef R2IQ_USE_REFERENCE_HEADERS
#include "fft_backend.h"
#else
#include <complex>
using fft_complex = std::complex<float>;
typedef void* FFTPlanHandle;
enum class FFTDirection { Forward, Backward };
class FFTBackend {
public:
    virtual ~FFTBackend() = default;
    virtual const char* name() const = 0;
    virtual FFTPlanHandle plan_r2c(int n, float* in, fft_complex* out) = 0;
    virtual FFTPlanHandle plan_c2c(int n, fft_complex* in, fft_complex* out, FFTDirection dir) = 0;
    virtual void execute_r2c(FFTPlanHandle plan, float* in, fft_complex* out) = 0;
    virtual void execute_c2c(FFTPlanHandle plan, fft_complex* in, fft_complex* out) = 0;
    virtual void destroy_plan(FFTPlanHandle plan) = 0;
    virtual void import_wisdom(const char*) {}
    virtual void export_wisdom(const char*) {}
    virtual void* alloc(size_t bytes) = 0;
    virtual void free(void* ptr) = 0;
};
FFTBackend* getFFTBackend();
#endif

class CudaFFTBackend : public FFTBackend {
public:
    explicit CudaFFTBackend(int device = -1, long batch = 1) : device_(device), batch_(batch < 1 ? 1 : batch) {}
    ~CudaFFTBackend() override = default;

    const char* name() const override;
    FFTPlanHandle plan_r2c(int n, float* in, fft_complex* out) override;
    FFTPlanHandle plan_c2c(int n, fft_complex* in, fft_complex* out, FFTDirection dir) override;
    void execute_r2c(FFTPlanHandle plan, float* in, fft_complex* out) override;
    void execute_c2c(FFTPlanHandle plan, fft_complex* in, fft_complex* out) override;
    void destroy_plan(FFTPlanHandle plan) override;
    void* alloc(size_t bytes) override;
    void free(void* ptr) override;

    // transforms per execute_* call of the plans created from now on; buffers hold `batch` transforms back to back
    void set_batch(long batch) { batch_ = batch < 1 ? 1 : batch; }
    long batch() const { return batch_; }
    // device-resident variant: the same kernel on device pointers, asynchronous on `stream`
    int execute_device(FFTPlanHandle plan, const void* d_in, void* d_out, long batch, void* stream);
    const char* last_error() const { return r2iq_cuda_last_error(); }
    int failures() const { return failures_; }   // execute_* calls that reported an error (the interface returns void)

private:
    int device_;
    long batch_;
    int failures_ = 0;
    mutable char name_[96] = "";
};
