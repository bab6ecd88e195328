// test_fft_backend.cpp -- CudaFFTBackend driven only through the reference's FFTBackend interface
// (Core/fft_backend.h:22-50), the way Core/fft_benchmark.cpp:39-97 drives a backend: alloc, plan, execute, destroy, free.
// Checked against a float64 DFT evaluated here (O(n^2), exact to 1e-15): rel-L2 <= 1e-6 for every size the benchmark
// sweeps, all three kinds, one transform per call and a batch of 5; pageable buffers (staged copies) as well.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "cuda_fft_backend.h"

namespace {
int g_failed = 0;
void check(bool ok, const char* what, int n, long batch, double err) {
    printf("%s %s n=%d batch=%ld rel_l2=%.3g\n", ok ? "PASS" : "FAIL", what, n, batch, err);
    if (!ok) g_failed++;
}

using cd = std::complex<double>;
std::vector<cd> dft(const std::vector<cd>& x, int sign) {
    const int n = (int)x.size();
    std::vector<cd> w(n), X(n);
    for (int j = 0; j < n; j++) w[j] = std::polar(1.0, sign * 2.0 * M_PI * j / n);
    for (int k = 0; k < n; k++) {
        cd acc = 0;
        for (int j = 0; j < n; j++) acc += x[j] * w[(long)j * k % n];
        X[k] = acc;
    }
    return X;
}
double rel_l2(const fft_complex* got, const std::vector<cd>& want, int count) {
    double num = 0, den = 0;
    for (int i = 0; i < count; i++) {
        num += std::norm(cd(got[i].real(), got[i].imag()) - want[i]);
        den += std::norm(want[i]);
    }
    return std::sqrt(num / den);
}

void run(FFTBackend* be, CudaFFTBackend* cuda, int n, long batch, bool pageable) {
    cuda->set_batch(batch);
    auto get = [&](size_t bytes) { return pageable ? malloc(bytes) : be->alloc(bytes); };
    auto put = [&](void* p) { pageable ? free(p) : be->free(p); };
    float* x = (float*)get(sizeof(float) * n * batch);
    fft_complex* X = (fft_complex*)get(sizeof(fft_complex) * (n / 2 + 1) * batch);
    fft_complex* u = (fft_complex*)get(sizeof(fft_complex) * n * batch);
    fft_complex* v = (fft_complex*)get(sizeof(fft_complex) * n * batch);
    for (long i = 0; i < (long)n * batch; i++) {
        x[i] = (float)rand() / RAND_MAX - 0.5f;
        u[i] = fft_complex((float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f);
    }
    FFTPlanHandle pr = be->plan_r2c(n, x, X);
    FFTPlanHandle pf = be->plan_c2c(n, u, v, FFTDirection::Forward);
    if (!pr || !pf) {
        printf("FAIL plan n=%d: %s\n", n, cuda->last_error());
        g_failed++;
        return;
    }
    be->execute_r2c(pr, x, X);
    be->execute_c2c(pf, u, v);
    double e_r = 0, e_f = 0, e_b = 0;
    const long probe[2] = {0, batch - 1};
    for (long b : probe) {
        std::vector<cd> xr(n), xu(n);
        for (int i = 0; i < n; i++) {
            xr[i] = x[b * n + i];
            xu[i] = cd(u[b * n + i].real(), u[b * n + i].imag());
        }
        e_r = std::fmax(e_r, rel_l2(X + b * (n / 2 + 1), dft(xr, -1), n / 2 + 1));
        e_f = std::fmax(e_f, rel_l2(v + b * n, dft(xu, -1), n));
    }
    FFTPlanHandle pb = be->plan_c2c(n, v, u, FFTDirection::Backward);
    std::vector<cd> vin(n);
    for (int i = 0; i < n; i++) vin[i] = cd(v[(batch - 1) * n + i].real(), v[(batch - 1) * n + i].imag());
    be->execute_c2c(pb, v, u);
    e_b = rel_l2(u + (batch - 1) * n, dft(vin, +1), n);
    const char* mem = pageable ? " (pageable)" : "";
    char what[64];
    snprintf(what, sizeof(what), "r2c%s", mem);
    check(e_r <= 1e-6, what, n, batch, e_r);
    snprintf(what, sizeof(what), "c2c forward%s", mem);
    check(e_f <= 1e-6, what, n, batch, e_f);
    snprintf(what, sizeof(what), "c2c backward%s", mem);
    check(e_b <= 1e-6, what, n, batch, e_b);
    be->destroy_plan(pr);
    be->destroy_plan(pf);
    be->destroy_plan(pb);
    put(x);
    put(X);
    put(u);
    put(v);
}
}  // namespace

int main() {
    CudaFFTBackend cuda;
    FFTBackend* be = &cuda;
    printf("backend: %s\n", be->name());
    const int sizes[] = {4096, 2048, 1024, 512, 256, 128};   // Core/fft_benchmark.cpp:12
    for (int n : sizes) {
        run(be, &cuda, n, 1, false);
        run(be, &cuda, n, 5, false);
    }
    run(be, &cuda, 1024, 3, true);
    run(be, &cuda, 4096, 300, false);       // beyond the in-place limit: staged through device memory
    // error behaviour: unsupported length -> null plan and a message, nothing thrown
    FFTPlanHandle bad = be->plan_c2c(1000, nullptr, nullptr, FFTDirection::Forward);
    check(bad == nullptr && cuda.last_error()[0] != 0, "unsupported length refused", 1000, 1, 0.0);
    check(cuda.failures() == 0, "no execute failures", 0, 0, 0.0);
    if (g_failed) {
        printf("%d FAILED\n", g_failed);
        return 1;
    }
    printf("ALL PASS\n");
    return 0;
}
