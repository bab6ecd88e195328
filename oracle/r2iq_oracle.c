/* oracle/r2iq_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's real->IQ down-converter (the fft_mt_r2iq path of
 * ShadowLive/ExtIO_sddc), scalar and single-threaded.  It is the portable "port" baseline and a
 * second opinion beside the NumPy restatement (oracle/r2iq_oracle.py); only tests/, smoke() and
 * bench.py's cpu_baseline leg may load it.  Parity status: PINNED -- tests/test_oracle_c.py checks
 * it against the golden vectors produced by the unmodified reference (tests/golden/).
 *
 * Reference lines restated (paths relative to /root/reference):
 *   Core/fir.cpp:7-25,48-105            Izero, KaiserWindow            -> izero(), kaiser()
 *   Core/fft_mt_r2iq.cpp:181-205        filter placement + spectrum    -> oracle_design_filter()
 *   Core/fft_mt_r2iq.h:36-51            convert_float<rand>            -> convert()
 *   Core/fft_mt_r2iq_impl.hpp:15-151    worker loop                    -> oracle_r2iq()
 * FFTs (third-party FFTW in the reference) are evaluated here by a float64 radix-2 transform
 * following FFTW's documented conventions (unnormalised, sign of the exponent = direction).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TRANSFER_SAMPLES 65536
#define ORACLE_PI 3.14159265358979323846

typedef struct { double re, im; } cplx;

/* in-place iterative radix-2 FFT, n a power of two, sign = -1 forward / +1 backward */
static void fft_inplace(cplx* a, int n, int sign) {
    for (int i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) { cplx t = a[i]; a[i] = a[j]; a[j] = t; }
    }
    for (int len = 2; len <= n; len <<= 1) {
        for (int i = 0; i < n; i += len) {
            for (int k = 0; k < len / 2; k++) {
                double ang = sign * 2.0 * ORACLE_PI * (double)k / (double)len;
                double wr = cos(ang), wi = sin(ang);
                cplx u = a[i + k], v = a[i + k + len / 2];
                double tr = v.re * wr - v.im * wi, ti = v.re * wi + v.im * wr;
                a[i + k].re = u.re + tr; a[i + k].im = u.im + ti;
                a[i + k + len / 2].re = u.re - tr; a[i + k + len / 2].im = u.im - ti;
            }
        }
    }
}

/* Core/fir.cpp:7-25 */
static float izero(float x) {
    float x2 = x / 2.0f, sum = 1.0f, ds = 1.0f, di = 1.0f, tmp;
    do {
        tmp = x2 / di;
        tmp *= tmp;
        ds *= tmp;
        sum += ds;
        di += 1.0f;
    } while (ds >= 1e-9f * sum);
    return sum;
}

/* Core/fir.cpp:48-105, num_taps > 0 form, float32 arithmetic throughout */
static void kaiser(int ntaps, float astop, float fpass, float fstop, float* coef) {
    const float kpi = 3.141592653f, k2pi = 2 * kpi;
    float fcut = (fstop + fpass) / 2.0f, beta;
    if (astop < 20.96f) beta = 0.0f;
    else if (astop >= 50.0f) beta = .1102f * (astop - 8.71f);
    else beta = .5842f * powf((astop - 20.96f), 0.4f) + .07886f * (astop - 20.96f);
    float fcenter = .5f * (float)(ntaps - 1), izb = izero(beta);
    for (int n = 0; n < ntaps; n++) {
        float x = (float)n - fcenter, c;
        if ((float)n == fcenter) c = 2.0f * fcut;
        else c = (float)sinf(k2pi * x * fcut) / (kpi * x);
        x = ((float)n - ((float)ntaps - 1.0f) / 2.0f) / (((float)ntaps - 1.0f) / 2.0f);
        coef[n] = 1.0f * c * izero(beta * sqrtf(1 - (x * x))) / izb;
    }
}

/* Core/fft_mt_r2iq.cpp:181-205: hw = halfFft complex floats (interleaved) */
int oracle_design_filter(float gain, int d, int fftn, float* hw) {
    int half = fftn / 2, ntaps = half / 4 + 1;
    if (d < 0 || d > 6) return -1;
    float* pht = (float*)malloc(sizeof(float) * ntaps);
    cplx* ht = (cplx*)calloc(half, sizeof(cplx));
    float bw = 64.0f / (float)(1 << d);
    kaiser(ntaps, 120.0f, 0.85f * bw / 128.0f, 1.1f * bw / 128.0f, pht);
    float gainadj = gain * 2048.0f / (float)fftn;
    for (int t = 0; t < ntaps; t++) ht[half - 1 - t].re = (double)(gainadj * pht[t]);
    fft_inplace(ht, half, -1);
    for (int k = 0; k < half; k++) { hw[2 * k] = (float)ht[k].re; hw[2 * k + 1] = (float)ht[k].im; }
    free(pht); free(ht);
    return half;
}

/* Core/fft_mt_r2iq.h:36-51 */
static void convert(const int16_t* in, double* out, int n, int rnd) {
    for (int m = 0; m < n; m++) {
        int16_t v = in[m];
        if (rnd && (v & 1)) v = (int16_t)(v ^ (-2));
        out[m] = (double)(float)v;
    }
}

void oracle_convert(const int16_t* in, float* out, long n, int rnd) {
    for (long m = 0; m < n; m++) {
        int16_t v = in[m];
        if (rnd && (v & 1)) v = (int16_t)(v ^ (-2));
        out[m] = (float)v;
    }
}

/* Core/fft_mt_r2iq_impl.hpp:15-151.  blocks: n_blocks*65536 int16; history: halfFft samples or
 * NULL (zeros); hw: filterHw[d] (halfFft complex floats) or NULL to design it from `gain`;
 * iq_out: n_blocks*(32768>>d) complex floats.  Returns complex samples written, <0 on error. */
long oracle_r2iq(const int16_t* blocks, long n_blocks, const int16_t* history, const float* hw, float gain, int d,
                 int tunebin, int rnd, int sideband, int fftn, float* iq_out) {
    if (d < 0 || d > 6 || n_blocks < 0) return -1;
    const int half = fftn / 2, hop = 3 * half / 2, nfr = TRANSFER_SAMPLES / hop + 1, mfft = half >> d;
    const long opb = (TRANSFER_SAMPLES / 2) >> d;
    float* H = (float*)malloc(sizeof(float) * 2 * half);
    if (hw) memcpy(H, hw, sizeof(float) * 2 * half);
    else oracle_design_filter(gain, d, fftn, H);
    double* t = (double*)malloc(sizeof(double) * (half + TRANSFER_SAMPLES));
    cplx* X = (cplx*)malloc(sizeof(cplx) * fftn);
    cplx* Y = (cplx*)malloc(sizeof(cplx) * mfft);
    int16_t* prev = (int16_t*)calloc(half, sizeof(int16_t));
    if (history) memcpy(prev, history, sizeof(int16_t) * half);
    /* impl.hpp:76-80 */
    const int count = (mfft / 2 < half - tunebin) ? mfft / 2 : half - tunebin;
    const int start = (mfft / 2 - tunebin > 0) ? mfft / 2 - tunebin : 0;
    for (long b = 0; b < n_blocks; b++) {
        const int16_t* blk = blocks + b * (long)TRANSFER_SAMPLES;
        convert(prev, t, half, rnd);                     /* impl.hpp:42|52 history          */
        convert(blk, t + half, TRANSFER_SAMPLES, rnd);   /* impl.hpp:48|53 the block        */
        memcpy(prev, blk + TRANSFER_SAMPLES - half, sizeof(int16_t) * half);
        float* out = iq_out + 2 * b * opb;
        for (int k = 0; k < nfr; k++) {
            for (int n = 0; n < fftn; n++) { X[n].re = t[(long)k * hop + n]; X[n].im = 0.0; }
            fft_inplace(X, fftn, -1);                    /* impl.hpp:88 r2c, bins 0..half   */
            memset(Y, 0, sizeof(cplx) * mfft);           /* impl.hpp:91-92,95-96 zero fill  */
            for (int m = 0; m < count; m++) {            /* impl.hpp:90                     */
                cplx x = X[tunebin + m]; double hr = H[2 * m], hi = H[2 * m + 1];
                Y[m].re = x.re * hr - x.im * hi; Y[m].im = x.im * hr + x.re * hi;
            }
            for (int m = start; m < mfft / 2; m++) {     /* impl.hpp:94                     */
                cplx x = X[tunebin - mfft / 2 + m];
                double hr = H[2 * (half - mfft / 2 + m)], hi = H[2 * (half - mfft / 2 + m) + 1];
                Y[mfft / 2 + m].re = x.re * hr - x.im * hi; Y[mfft / 2 + m].im = x.im * hr + x.re * hi;
            }
            fft_inplace(Y, mfft, +1);                    /* impl.hpp:98 backward c2c        */
            const double sg = sideband ? -1.0 : 1.0;     /* impl.hpp:124-137 copy<flip>     */
            if (k == 0) {
                for (int i = 0; i < mfft / 2; i++) {
                    out[2 * i] = (float)Y[mfft / 4 + i].re; out[2 * i + 1] = (float)(sg * Y[mfft / 4 + i].im);
                }
            } else {
                float* o = out + 2 * (mfft / 2 + (long)(3 * mfft / 4) * (k - 1));
                for (int i = 0; i < 3 * mfft / 4; i++) { o[2 * i] = (float)Y[i].re; o[2 * i + 1] = (float)(sg * Y[i].im); }
            }
        }
    }
    free(H); free(t); free(X); free(Y); free(prev);
    return n_blocks * opb;
}
