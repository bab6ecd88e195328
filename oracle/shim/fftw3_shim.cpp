// oracle/shim/fftw3_shim.cpp -- TEST INFRASTRUCTURE, not product code.
//
// Implements the handful of fftwf_* entry points declared in oracle/shim/fftw3.h so that the
// reference's own fft_mt_r2iq sources link without libfftw3f (absent from this image).
// Transform semantics follow FFTW's documented definitions (unnormalised, sign = exponent sign):
//   c2c : out[k] = sum_n in[n] * exp(sign * 2*pi*i*k*n/N)
//   r2c : out[k] = sum_n in[n] * exp(-2*pi*i*k*n/N), k = 0 .. N/2
// Two backends:
//   DOUBLE (default) -- iterative radix-2 FFT evaluated in float64 with a float64 twiddle table,
//                       result rounded once to float32.  Portable and deterministic; this is what
//                       the parity fixtures are generated with.
//   MKL              -- Intel MKL DFTI single precision resolved with dlopen/dlsym from torch's
//                       libtorch_cpu.so.  Used for the timed CPU baseline, because it is a
//                       production-grade float32 FFT of the kind the reference links against.
#include "fftw3.h"

#include <dlfcn.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <complex>
#include <mutex>
#include <vector>

namespace {

typedef std::complex<double> cd;

// ---------------------------------------------------------------------------------------------
// portable float64 FFT (power-of-two sizes only; the reference only plans powers of two)
// ---------------------------------------------------------------------------------------------
struct DoubleFFT {
    int n = 0;
    int log2n = 0;
    std::vector<cd> tw;        // exp(-2*pi*i*k/n), k < n/2
    std::vector<uint32_t> rev;

    explicit DoubleFFT(int n_) : n(n_) {
        while ((1 << log2n) < n) log2n++;
        if ((1 << log2n) != n) {
            fprintf(stderr, "fftw3 shim: size %d is not a power of two\n", n);
            abort();
        }
        tw.resize(n / 2 > 0 ? n / 2 : 1);
        for (int k = 0; k < n / 2; k++) {
            long double a = -2.0L * 3.141592653589793238462643383279502884L * k / n;
            tw[k] = cd((double)cosl(a), (double)sinl(a));
        }
        rev.resize(n);
        for (int i = 0; i < n; i++) {
            uint32_t r = 0;
            for (int b = 0; b < log2n; b++)
                if (i & (1 << b)) r |= 1u << (log2n - 1 - b);
            rev[i] = r;
        }
    }

    // in-place on `a` (already bit-reverse permuted); sign = -1 forward, +1 backward
    void run(std::vector<cd>& a, int sign) const {
        for (int len = 2; len <= n; len <<= 1) {
            int half = len >> 1, step = n / len;
            for (int i = 0; i < n; i += len) {
                for (int j = 0; j < half; j++) {
                    cd w = tw[j * step];
                    if (sign > 0) w = std::conj(w);
                    cd u = a[i + j], v = a[i + j + half] * w;
                    a[i + j] = u + v;
                    a[i + j + half] = u - v;
                }
            }
        }
    }
};

// ---------------------------------------------------------------------------------------------
// MKL DFTI, hand-declared (no MKL headers in the image).  Enumerator values are those of
// mkl_dfti.h; they are verified at run time by a self-check against the float64 FFT before the
// backend is accepted (fftwshim_set_backend).
// ---------------------------------------------------------------------------------------------
typedef void* DFTI_HANDLE;
enum {
    DFTI_FORWARD_SCALE = 4,
    DFTI_BACKWARD_SCALE = 5,
    DFTI_CONJUGATE_EVEN_STORAGE = 10,
    DFTI_PLACEMENT = 11,
    DFTI_COMPLEX = 32,
    DFTI_REAL = 33,
    DFTI_SINGLE = 35,
    DFTI_COMPLEX_COMPLEX = 39,
    DFTI_INPLACE = 43,
    DFTI_NOT_INPLACE = 44,
};
struct MklApi {
    void* lib = nullptr;
    long (*create_s_1d)(DFTI_HANDLE*, int domain, long n) = nullptr;
    long (*set_value)(DFTI_HANDLE, int param, ...) = nullptr;
    long (*commit)(DFTI_HANDLE) = nullptr;
    long (*fwd)(DFTI_HANDLE, void*, ...) = nullptr;
    long (*bwd)(DFTI_HANDLE, void*, ...) = nullptr;
    long (*free_desc)(DFTI_HANDLE*) = nullptr;
    bool ok() const { return create_s_1d && set_value && commit && fwd && bwd && free_desc; }
};
MklApi g_mkl;
int g_backend = FFTWSHIM_BACKEND_DOUBLE;
std::mutex g_mu;

bool mkl_load(const char* path) {
    if (g_mkl.ok()) return true;
    void* lib = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
        fprintf(stderr, "fftw3 shim: dlopen(%s) failed: %s\n", path ? path : "(null)", dlerror());
        return false;
    }
    g_mkl.lib = lib;
    g_mkl.create_s_1d = (long (*)(DFTI_HANDLE*, int, long))dlsym(lib, "DftiCreateDescriptor_s_1d");
    g_mkl.set_value = (long (*)(DFTI_HANDLE, int, ...))dlsym(lib, "DftiSetValue");
    g_mkl.commit = (long (*)(DFTI_HANDLE))dlsym(lib, "DftiCommitDescriptor");
    g_mkl.fwd = (long (*)(DFTI_HANDLE, void*, ...))dlsym(lib, "DftiComputeForward");
    g_mkl.bwd = (long (*)(DFTI_HANDLE, void*, ...))dlsym(lib, "DftiComputeBackward");
    g_mkl.free_desc = (long (*)(DFTI_HANDLE*))dlsym(lib, "DftiFreeDescriptor");
    return g_mkl.ok();
}

}  // namespace

struct fftwshim_plan_s {
    int n = 0;
    int sign = -1;
    bool r2c = false;
    int backend = FFTWSHIM_BACKEND_DOUBLE;
    DoubleFFT* dfft = nullptr;        // DOUBLE backend
    DFTI_HANDLE mkl_inplace = nullptr;   // MKL backend, in == out
    DFTI_HANDLE mkl_outplace = nullptr;  // MKL backend, in != out
};

// A program that never calls fftwshim_set_backend (the reference's own fft_benchmark main) selects MKL through the
// environment: FFTW3_SHIM_MKL=<path of a library exporting the DFTI entry points> is honoured at the first plan.
static void backend_from_env() {
    static std::once_flag once;
    std::call_once(once, [] {
        const char* path = getenv("FFTW3_SHIM_MKL");
        if (path && *path && fftwshim_set_backend(FFTWSHIM_BACKEND_MKL, path) != 0)
            fprintf(stderr, "fftw3 shim: FFTW3_SHIM_MKL=%s not usable, staying on the float64 FFT\n", path);
    });
}

static fftwf_plan make_plan(int n, int sign, bool r2c) {
    fftwshim_plan_s* p = new fftwshim_plan_s();
    p->n = n;
    p->sign = sign;
    p->r2c = r2c;
    p->backend = g_backend;
    if (p->backend == FFTWSHIM_BACKEND_MKL) {
        std::lock_guard<std::mutex> lk(g_mu);
        long st = 0;
        if (r2c) {
            st |= g_mkl.create_s_1d(&p->mkl_outplace, DFTI_REAL, n);
            st |= g_mkl.set_value(p->mkl_outplace, DFTI_PLACEMENT, DFTI_NOT_INPLACE);
            st |= g_mkl.set_value(p->mkl_outplace, DFTI_CONJUGATE_EVEN_STORAGE, DFTI_COMPLEX_COMPLEX);
            st |= g_mkl.commit(p->mkl_outplace);
        } else {
            st |= g_mkl.create_s_1d(&p->mkl_outplace, DFTI_COMPLEX, n);
            st |= g_mkl.set_value(p->mkl_outplace, DFTI_PLACEMENT, DFTI_NOT_INPLACE);
            st |= g_mkl.commit(p->mkl_outplace);
            st |= g_mkl.create_s_1d(&p->mkl_inplace, DFTI_COMPLEX, n);
            st |= g_mkl.set_value(p->mkl_inplace, DFTI_PLACEMENT, DFTI_INPLACE);
            st |= g_mkl.commit(p->mkl_inplace);
        }
        if (st != 0) {
            fprintf(stderr, "fftw3 shim: MKL descriptor setup failed (status %ld), using float64 FFT\n", st);
            p->backend = FFTWSHIM_BACKEND_DOUBLE;
        }
    }
    if (p->backend == FFTWSHIM_BACKEND_DOUBLE) p->dfft = new DoubleFFT(n);
    return p;
}

extern "C" {

void* fftwf_malloc(size_t n) {
    void* p = nullptr;
    if (posix_memalign(&p, 64, n ? n : 64) != 0) return nullptr;
    return p;
}
void fftwf_free(void* p) { free(p); }

fftwf_plan fftwf_plan_dft_1d(int n, fftwf_complex*, fftwf_complex*, int sign, unsigned) {
    backend_from_env();
    return make_plan(n, sign, false);
}
fftwf_plan fftwf_plan_dft_r2c_1d(int n, float*, fftwf_complex*, unsigned) {
    backend_from_env();
    return make_plan(n, -1, true);
}

void fftwf_execute_dft(const fftwf_plan p, fftwf_complex* in, fftwf_complex* out) {
    if (p->backend == FFTWSHIM_BACKEND_MKL) {
        if (in == out) {
            if (p->sign < 0) g_mkl.fwd(p->mkl_inplace, in);
            else g_mkl.bwd(p->mkl_inplace, in);
        } else {
            if (p->sign < 0) g_mkl.fwd(p->mkl_outplace, in, out);
            else g_mkl.bwd(p->mkl_outplace, in, out);
        }
        return;
    }
    const DoubleFFT& f = *p->dfft;
    std::vector<cd> a(f.n);
    for (int i = 0; i < f.n; i++) a[f.rev[i]] = cd(in[i][0], in[i][1]);
    f.run(a, p->sign);
    for (int i = 0; i < f.n; i++) {
        out[i][0] = (float)a[i].real();
        out[i][1] = (float)a[i].imag();
    }
}

void fftwf_execute_dft_r2c(const fftwf_plan p, float* in, fftwf_complex* out) {
    if (p->backend == FFTWSHIM_BACKEND_MKL) {
        g_mkl.fwd(p->mkl_outplace, in, out);
        return;
    }
    const DoubleFFT& f = *p->dfft;
    std::vector<cd> a(f.n);
    for (int i = 0; i < f.n; i++) a[f.rev[i]] = cd(in[i], 0.0);
    f.run(a, -1);
    for (int i = 0; i <= f.n / 2; i++) {
        out[i][0] = (float)a[i].real();
        out[i][1] = (float)a[i].imag();
    }
}

void fftwf_destroy_plan(fftwf_plan p) {
    if (!p) return;
    if (p->mkl_inplace) g_mkl.free_desc(&p->mkl_inplace);
    if (p->mkl_outplace) g_mkl.free_desc(&p->mkl_outplace);
    delete p->dfft;
    delete p;
}

int fftwf_import_wisdom_from_filename(const char*) { return 0; }
int fftwf_export_wisdom_to_filename(const char*) { return 0; }

int fftwshim_get_backend(void) { return g_backend; }

int fftwshim_set_backend(int backend, const char* libpath) {
    if (backend == FFTWSHIM_BACKEND_DOUBLE) {
        g_backend = backend;
        return 0;
    }
    if (backend != FFTWSHIM_BACKEND_MKL) return -1;
    if (!mkl_load(libpath)) return -2;
    // Self-check the hand-declared DFTI enumerators: r2c(256) and backward c2c(64) must agree
    // with the float64 FFT to float32 accuracy before MKL is accepted.
    g_backend = FFTWSHIM_BACKEND_MKL;
    bool good = true;
    {
        const int n = 256;
        std::vector<float> x(n);
        for (int i = 0; i < n; i++) x[i] = (float)((i * 37 % 101) - 50) + 0.25f * (i % 7);
        std::vector<float> o1(2 * (n / 2 + 1)), o2(2 * (n / 2 + 1));
        fftwf_plan pm = make_plan(n, -1, true);
        good = good && pm->backend == FFTWSHIM_BACKEND_MKL;
        fftwf_execute_dft_r2c(pm, x.data(), (fftwf_complex*)o1.data());
        g_backend = FFTWSHIM_BACKEND_DOUBLE;
        fftwf_plan pd = make_plan(n, -1, true);
        fftwf_execute_dft_r2c(pd, x.data(), (fftwf_complex*)o2.data());
        g_backend = FFTWSHIM_BACKEND_MKL;
        double num = 0, den = 0;
        for (size_t i = 0; i < o1.size(); i++) {
            num += (double)(o1[i] - o2[i]) * (o1[i] - o2[i]);
            den += (double)o2[i] * o2[i];
        }
        good = good && den > 0 && sqrt(num / den) < 1e-5;
        fftwf_destroy_plan(pm);
        fftwf_destroy_plan(pd);
    }
    {
        const int n = 64;
        std::vector<float> x(2 * n), o1(2 * n), o2(2 * n), o3(2 * n);
        for (int i = 0; i < 2 * n; i++) x[i] = (float)((i * 29 % 53) - 26) * 0.5f;
        fftwf_plan pm = make_plan(n, +1, false);
        good = good && pm->backend == FFTWSHIM_BACKEND_MKL;
        fftwf_execute_dft(pm, (fftwf_complex*)x.data(), (fftwf_complex*)o1.data());
        o3 = x;
        fftwf_execute_dft(pm, (fftwf_complex*)o3.data(), (fftwf_complex*)o3.data());
        g_backend = FFTWSHIM_BACKEND_DOUBLE;
        fftwf_plan pd = make_plan(n, +1, false);
        fftwf_execute_dft(pd, (fftwf_complex*)x.data(), (fftwf_complex*)o2.data());
        g_backend = FFTWSHIM_BACKEND_MKL;
        double num = 0, num3 = 0, den = 0;
        for (size_t i = 0; i < o1.size(); i++) {
            num += (double)(o1[i] - o2[i]) * (o1[i] - o2[i]);
            num3 += (double)(o3[i] - o2[i]) * (o3[i] - o2[i]);
            den += (double)o2[i] * o2[i];
        }
        good = good && den > 0 && sqrt(num / den) < 1e-5 && sqrt(num3 / den) < 1e-5;
        fftwf_destroy_plan(pm);
        fftwf_destroy_plan(pd);
    }
    if (!good) {
        fprintf(stderr, "fftw3 shim: MKL backend failed its self-check, staying on float64 FFT\n");
        g_backend = FFTWSHIM_BACKEND_DOUBLE;
        return -3;
    }
    return 0;
}

}  // extern "C"
