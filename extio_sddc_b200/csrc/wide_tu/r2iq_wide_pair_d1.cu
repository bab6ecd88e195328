// FFTN 8192 kernel for decimation index 1: a module of its own (see r2iq_wide_pair.cu)
#define R2IQ_WIDE_TU_D 1

#include "../r2iq_wide_pair.cu"
