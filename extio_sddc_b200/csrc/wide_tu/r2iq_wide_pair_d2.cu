// FFTN 8192 kernel for decimation index 2: a module of its own (see r2iq_wide_pair.cu)
#define R2IQ_WIDE_TU_D 2

#include "../r2iq_wide_pair.cu"
