/* oracle/shim/fftw3.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Minimal stand-in for the FFTW3 single-precision API so that the reference's own hot-path
 * sources (Core/fft_mt_r2iq*.cpp, compiled in place from /root/reference by oracle/Makefile)
 * build on a box without libfftw3f.  Only the entry points those sources call are declared:
 *   Core/fft_mt_r2iq.cpp:76,154      wisdom import/export   (no-ops here)
 *   Core/fft_mt_r2iq.cpp:173-180     fftwf_malloc, fftwf_plan_dft_1d
 *   Core/fft_mt_r2iq.cpp:205         fftwf_execute_dft
 *   Core/fft_mt_r2iq.cpp:221-225     fftwf_plan_dft_r2c_1d, fftwf_plan_dft_1d
 *   Core/fft_mt_r2iq_impl.hpp:88,98  fftwf_execute_dft_r2c, fftwf_execute_dft
 * The transforms themselves are provided by oracle/shim/fftw3_shim.cpp: either a portable
 * double-precision FFT (default; deterministic everywhere) or Intel MKL DFTI single precision
 * resolved at run time from torch's libtorch_cpu.so (MKL's FFTW3 wrapper is a backend the
 * reference itself supports, Core/fft_backend_mkl.cpp:1-8).  FFTW proper is not installed in
 * this image, so "the reference FFTW path" means the reference's loop over this shim.
 */
#ifndef ORACLE_SHIM_FFTW3_H
#define ORACLE_SHIM_FFTW3_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef float fftwf_complex[2];
typedef struct fftwshim_plan_s* fftwf_plan;

#define FFTW_FORWARD (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE (0U)
#define FFTW_ESTIMATE (1U << 6)

void* fftwf_malloc(size_t n);
void fftwf_free(void* p);
fftwf_plan fftwf_plan_dft_1d(int n, fftwf_complex* in, fftwf_complex* out, int sign, unsigned flags);
fftwf_plan fftwf_plan_dft_r2c_1d(int n, float* in, fftwf_complex* out, unsigned flags);
void fftwf_execute_dft(const fftwf_plan p, fftwf_complex* in, fftwf_complex* out);
void fftwf_execute_dft_r2c(const fftwf_plan p, float* in, fftwf_complex* out);
void fftwf_destroy_plan(fftwf_plan p);
int fftwf_import_wisdom_from_filename(const char* filename);
int fftwf_export_wisdom_to_filename(const char* filename);

/* shim controls (not part of FFTW) */
enum { FFTWSHIM_BACKEND_DOUBLE = 0, FFTWSHIM_BACKEND_MKL = 1 };
/* Select the backend used by plans created afterwards.  For MKL, `libpath` is the shared object
 * exporting the Dfti* symbols (torch/lib/libtorch_cpu.so).  Returns 0 on success, <0 if the
 * backend is unavailable or failed its self-check (the double backend stays selected). */
int fftwshim_set_backend(int backend, const char* libpath);
int fftwshim_get_backend(void);

#ifdef __cplusplus
}
#endif
#endif
