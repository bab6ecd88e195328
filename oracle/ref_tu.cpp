// oracle/ref_tu.cpp -- TEST INFRASTRUCTURE, not product code.
//
// Wrapper translation unit that compiles ONE of the reference's own hot-path sources in place
// (from $(REF)/Core, passed with -I; nothing is copied into this repo) with FFTN_R_ADC overridden.
// The reference hard-codes `#define FFTN_R_ADC (8192)` (Core/config.h:49) while the north-star
// configuration is FFTN 2048 ("tested at 2048, 8192, ..." ibid.); every size on the hot path is
// derived from that macro (Core/fft_mt_r2iq.h:18-19), so including config.h first, re-defining
// the macro and relying on config.h's include guard changes only that one constant.
//
// Build with  -DORACLE_FFTN=<2048|8192> -DREF_TU=<n>  (see oracle/Makefile):
//   1  Core/fft_mt_r2iq.cpp          ctor/Init/setFreqOffset/TurnOn/TurnOff/dispatcher
//   2  Core/fft_mt_r2iq_def.cpp      worker loop, baseline ISA
//   3  Core/fft_mt_r2iq_avx.cpp      worker loop, -mavx
//   4  Core/fft_mt_r2iq_avx2.cpp     worker loop, AVX2 intrinsics, -mavx2
//   5  Core/fft_mt_r2iq_avx512.cpp   worker loop, -mavx512f
#include "config.h"

#ifndef ORACLE_FFTN
#error "define ORACLE_FFTN"
#endif
#undef FFTN_R_ADC
#define FFTN_R_ADC (ORACLE_FFTN)

#if REF_TU == 1
#include "fft_mt_r2iq.cpp"
#elif REF_TU == 2
#include "fft_mt_r2iq_def.cpp"
#elif REF_TU == 3
#include "fft_mt_r2iq_avx.cpp"
#elif REF_TU == 4
#include "fft_mt_r2iq_avx2.cpp"
#elif REF_TU == 5
#include "fft_mt_r2iq_avx512.cpp"
#else
#error "REF_TU must be 1..5"
#endif
