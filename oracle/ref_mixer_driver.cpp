// oracle/ref_mixer_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// C-ABI access to the reference's fine-tune mixer exactly as RadioHandlerClass uses it
// (Core/RadioHandler.cpp:33-37,292-297): shift_limited_unroll_C_sse_init(fc, 0) once, then
// shift_limited_unroll_C_sse_inp_c on consecutive output slots with the state carried.
// Core/pffft/pf_mixer.cpp is compiled in place by oracle/Makefile into oracle/_ref/libref_mixer.so.
#include "pf_mixer.h"

extern "C" void refmix_apply(float fc, float* iq_interleaved, long n_complex, long slot_complex) {
    shift_limited_unroll_C_sse_data_t st = shift_limited_unroll_C_sse_init(fc, 0.0F);
    complexf* p = reinterpret_cast<complexf*>(iq_interleaved);
    if (slot_complex <= 0) slot_complex = n_complex;
    for (long done = 0; done < n_complex; done += slot_complex) {
        long n = n_complex - done < slot_complex ? n_complex - done : slot_complex;
        shift_limited_unroll_C_sse_inp_c(p + done, (int)n, &st);
    }
}
