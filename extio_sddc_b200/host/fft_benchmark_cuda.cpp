// fft_benchmark_cuda.cpp -- the reference's FFT backend benchmark (Core/fft_benchmark.cpp:39-128: r2c / c2c forward /
// c2c backward at 4096 ... 128, 10 warm-up + 1000 timed calls, table of microseconds per call) for CudaFFTBackend.
//
// Two tables:
//   1. the reference's call shape: ONE transform per execute_* call on buffers from backend->alloc() (host memory),
//      blocking -- a launch-latency measurement on a GPU, printed because it is what the reference's program measures;
//   2. the batched shape this backend is built for: `batch` transforms per call on the same kind of buffers
//      (host -> device -> host inside every call), reported per transform.
// Device-resident throughput (no PCIe in the timed region) is measured with CUDA events by tools/fft_benchmark.py.
//   usage: fft_benchmark_cuda [batch=4096] [iters=1000]
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "cuda_fft_backend.h"

namespace {
const int kSizes[] = {4096, 2048, 1024, 512, 256, 128};
constexpr int kWarmup = 10;

struct Row {
    int n;
    double r2c_us, fwd_us, bwd_us;
};

template <typename F>
double time_us(int iters, F&& call) {
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < iters; i++) call();
    const auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double, std::micro>(t1 - t0).count() / iters;
}

bool run_size(CudaFFTBackend& be, int n, int iters, Row* row) {
    const long b = be.batch();
    auto* x = static_cast<float*>(be.alloc(sizeof(float) * n * b));
    auto* X = static_cast<fft_complex*>(be.alloc(sizeof(fft_complex) * (n / 2 + 1) * b));
    auto* u = static_cast<fft_complex*>(be.alloc(sizeof(fft_complex) * n * b));
    auto* v = static_cast<fft_complex*>(be.alloc(sizeof(fft_complex) * n * b));
    if (!x || !X || !u || !v) return false;
    for (long i = 0; i < (long)n * b; i++) {
        x[i] = (float)rand() / RAND_MAX - 0.5f;
        u[i] = fft_complex((float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f);
    }
    FFTPlanHandle pr = be.plan_r2c(n, x, X);
    FFTPlanHandle pf = be.plan_c2c(n, u, v, FFTDirection::Forward);
    FFTPlanHandle pb = be.plan_c2c(n, v, u, FFTDirection::Backward);
    if (!pr || !pf || !pb) return false;
    for (int i = 0; i < kWarmup; i++) {
        be.execute_r2c(pr, x, X);
        be.execute_c2c(pf, u, v);
        be.execute_c2c(pb, v, u);
    }
    row->n = n;
    row->r2c_us = time_us(iters, [&] { be.execute_r2c(pr, x, X); });
    row->fwd_us = time_us(iters, [&] { be.execute_c2c(pf, u, v); });
    row->bwd_us = time_us(iters, [&] { be.execute_c2c(pb, v, u); });
    be.destroy_plan(pr);
    be.destroy_plan(pf);
    be.destroy_plan(pb);
    be.free(x);
    be.free(X);
    be.free(u);
    be.free(v);
    return be.failures() == 0;
}

bool table(CudaFFTBackend& be, int iters, bool per_transform) {
    const double div = per_transform ? (double)be.batch() : 1.0;
    printf("=======================================================================\n");
    printf("                    FFT BACKEND BENCHMARK\n");
    printf("=======================================================================\n");
    printf("Backend: %s%s\n", be.name(), per_transform ? ", microseconds PER TRANSFORM, host buffers" : ", one transform per call, host buffers");
    printf("Iterations: %d (warmup: %d)\n", iters, kWarmup);
    printf("-----------------------------------------------------------------------\n");
    printf("  Size  |   R2C (us)   |  C2C Fwd (us) |  C2C Bwd (us) |  Total (us)\n");
    printf("-----------------------------------------------------------------------\n");
    double total = 0;
    for (int n : kSizes) {
        Row r{};
        if (!run_size(be, n, iters, &r)) {
            fprintf(stderr, "fft_benchmark_cuda: size %d failed: %s\n", n, be.last_error());
            return false;
        }
        const double row_total = (r.r2c_us + r.fwd_us + r.bwd_us) / div;
        total += row_total;
        printf("  %4d  |  %9.4f   |   %9.4f   |   %9.4f   |  %9.4f\n", r.n, r.r2c_us / div, r.fwd_us / div, r.bwd_us / div, row_total);
    }
    printf("-----------------------------------------------------------------------\n");
    printf("Total time per iteration: %.4f us\n", total);
    printf("=======================================================================\n\n");
    return true;
}
}  // namespace

int main(int argc, char** argv) {
    const long batch = argc > 1 ? atol(argv[1]) : 4096;
    const int iters = argc > 2 ? atoi(argv[2]) : 1000;
    CudaFFTBackend single(-1, 1);
    if (!table(single, iters, false)) return 1;
    CudaFFTBackend batched(-1, batch);
    if (!table(batched, iters >= 50 ? iters / 50 : 1, true)) return 1;
    return 0;
}
