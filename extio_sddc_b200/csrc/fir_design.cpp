// fir_design.cpp -- host-side design of the decimation filters (once per Init).
//
// Same filter family as the reference: a Kaiser-windowed ideal low-pass, evaluated in single
// precision because the roundings of that evaluation are part of what "the reference's filter"
// is (Core/fir.cpp:48-105 KaiserWindow, Core/fir.cpp:7-25 Izero; call site and parameters
// Core/fft_mt_r2iq.cpp:181-205).  Frequency response: direct DFT in double precision, rounded once.
#include "fir_design.h"

#include <cmath>
#include <vector>

namespace r2iq {

namespace {

// Zeroth-order modified Bessel function of the first kind by its power series
// I0(x) = sum_k ((x/2)^k / k!)^2, summed in float until the term drops below 1e-9 of the sum.
float bessel_i0(float x) {
    const float half_x = x / 2.0f;
    float total = 1.0f;
    float term = 1.0f;
    float k = 1.0f;
    for (;;) {
        float q = half_x / k;
        q *= q;
        term *= q;
        total += term;
        k += 1.0f;
        if (!(term >= 1e-9f * total)) break;
    }
    return total;
}

}  // namespace

std::vector<float> kaiser_lowpass(int num_taps, float stop_db, float f_pass, float f_stop) {
    const float pi = 3.141592653f;
    const float two_pi = 2 * pi;
    const float f_cut = (f_stop + f_pass) / 2.0f;   // -6 dB point, relative to the sample rate

    float beta;
    if (stop_db < 20.96f) beta = 0.0f;
    else if (stop_db >= 50.0f) beta = .1102f * (stop_db - 8.71f);
    else beta = .5842f * powf(stop_db - 20.96f, 0.4f) + .07886f * (stop_db - 20.96f);

    std::vector<float> taps(num_taps);
    const float centre = .5f * (float)(num_taps - 1);
    const float half_span = ((float)num_taps - 1.0f) / 2.0f;
    const float i0_beta = bessel_i0(beta);
    for (int n = 0; n < num_taps; n++) {
        const float t = (float)n - centre;
        float ideal;
        if ((float)n == centre) ideal = 2.0f * f_cut;
        else ideal = (float)sinf(two_pi * t * f_cut) / (pi * t);
        const float r = ((float)n - half_span) / half_span;
        taps[n] = 1.0f * ideal * bessel_i0(beta * sqrtf(1 - (r * r))) / i0_beta;
    }
    return taps;
}

std::vector<std::complex<float>> design_filter_response(float gain, int d, int fftn) {
    const int half = fftn / 2;
    const int ntaps = half / 4 + 1;
    const float bw = 64.0f / (float)(1 << d);
    const std::vector<float> taps = kaiser_lowpass(ntaps, 120.0f, 0.85f * bw / 128.0f, 1.1f * bw / 128.0f);
    const float gainadj = gain * 2048.0f / (float)fftn;

    // taps sit time-reversed at the END of the halfFft window: an anticausal circular filter,
    // which is why the LAST quarter of every inverse transform is the part that gets discarded
    std::vector<float> ht(half, 0.0f);
    for (int t = 0; t < ntaps; t++) ht[half - 1 - t] = gainadj * taps[t];

    std::vector<std::complex<float>> hw(half);
    const double w0 = -2.0 * 3.14159265358979323846 / (double)half;
    for (int k = 0; k < half; k++) {
        double re = 0.0, im = 0.0;
        for (int t = half - ntaps; t < half; t++) {
            // reduce k*t mod half before scaling so the angle stays small and accurate
            const double ang = w0 * (double)(((long long)k * t) % half);
            re += (double)ht[t] * cos(ang);
            im += (double)ht[t] * sin(ang);
        }
        hw[k] = std::complex<float>((float)re, (float)im);
    }
    return hw;
}

}  // namespace r2iq
