// FFTN 8192 kernel for decimation index 5, direct-phase mixer in the store epilogue: a module of its own (see r2iq_wide_pair.cu)
#define R2IQ_WIDE_TU_D 5
#define R2IQ_WIDE_TU_MIX
#include "../r2iq_wide_pair.cu"
