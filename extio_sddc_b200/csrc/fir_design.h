// fir_design.h -- host-side filter design for the r2iq CUDA backend (see fir_design.cpp).
#pragma once
#include <complex>
#include <vector>

namespace r2iq {

// Kaiser-windowed sinc low-pass with `num_taps` taps, stop-band attenuation `stop_db`, pass/stop
// edges relative to the sample rate.  Single-precision evaluation (Core/fir.cpp:48-105).
std::vector<float> kaiser_lowpass(int num_taps, float stop_db, float f_pass, float f_stop);

// filterHw[d] of fft_mt_r2iq::Init (Core/fft_mt_r2iq.cpp:181-205): halfFft = fftn/2 complex bins.
std::vector<std::complex<float>> design_filter_response(float gain, int d, int fftn);

}  // namespace r2iq
