// r2iq_launch.cu -- kernel selection per decimation (launch_fused), and the two small stand-alone kernels: the residual
// fine-tune mixer as a separate pass and the bit-exact int16 -> float stage exposed for parity tests.
#include "r2iq_kernels.cuh"

#include <cstdio>
#include <cstdlib>

#include "dual_frame.cuh"

namespace r2iq {

static int g_kernel_gen = 0;
int kernel_generation() { return g_kernel_gen; }
#ifdef R2IQ_AB_VARIANTS
void set_kernel_generation(int gen) { g_kernel_gen = gen; }
int kernel_variant(int dflt) {
    // unset: every kernel family runs its fastest measured combination (`dflt`)
    const char* v = getenv("R2IQ_VARIANT");
    return v ? atoi(v) : dflt;
}
#else
void set_kernel_generation(int) {}                 // the shipped library has no alternatives to select
int kernel_variant(int dflt) { return dflt; }      // ... and reads no environment on the launch path
#endif

cudaError_t setup_fused_kernels() {
    cudaError_t e;
    if ((e = setup_pair_kernels()) != cudaSuccess) return e;
#ifdef R2IQ_AB_VARIANTS
    if ((e = setup_dual_kernels()) != cudaSuccess) return e;
    if ((e = setup_dual_d0_kernels()) != cudaSuccess) return e;
    if ((e = setup_dual_d1_kernels()) != cudaSuccess) return e;
    if ((e = setup_dual_dn_kernels()) != cudaSuccess) return e;
    if ((e = setup_gen1_kernels()) != cudaSuccess) return e;
#endif
    return cudaSuccess;
}

// Name (as ncu and cuobjdump print it) of the kernel launch_fused picks for decimation index d -- reported through the
// C ABI so that measurements name the kernel that actually ran.
const char* fused_kernel_name(int d, bool direct_mixer) {
    static char names[kNDec][2][64];
    if (d < 0 || d >= kNDec) return "";
    char* n = names[d][direct_mixer ? 1 : 0];
    if (!n[0]) {
        if (g_kernel_gen == 0) snprintf(n, 64, "r2iq_pair_kernel<%d, %d, %d>", d, d == 0 ? 8 : 12, direct_mixer ? 2 : 0);
        else snprintf(n, 64, "kernel generation %d (A/B build), d=%d", g_kernel_gen, d);
    }
    return n;
}

cudaError_t launch_fused(int d, const FusedParams& p, int sm_count, cudaStream_t stream) {
    // The shipped library has ONE kernel per decimation: r2iq_pair.cu -- 8 free-running warps per SM for d = 0, 12 for
    // d = 1..6 (B200, 8192 blocks: 529 / 663 / 855 / 989 / 1111 / 1284 / 1297 Gsamples/s at d = 0..6 against 454 / 601 /
    // 759 / 865 / 929 / 963 / 833 of the round-1 kernels).
    if (g_kernel_gen == 0) return launch_pair(d, d == 0 ? 8 : 12, p, sm_count, stream);
#ifdef R2IQ_AB_VARIANTS
    // Comparison builds (-DR2IQ_AB_VARIANTS, env R2IQ_KERNEL_GEN): 1 = first generation (r2iq_gen1.cu); 2 / 3 = first
    // dual-frame family without / with TMA (r2iq_dual.cu); 4 = generic packed-backward kernel also for d = 1;
    // 5 = the round-1 default selection (r2iq_dual_d0 / _d1 / _dn.cu).  All meet the parity bar.
    if (g_kernel_gen == 5 || g_kernel_gen == 4) {
        if (d == 0) return launch_dual_d0(p, sm_count, stream, true);
        if (d == 1 && g_kernel_gen == 5) return launch_dual_d1(p, sm_count, stream, true);
        if (d <= 5) return launch_dual_dn(d, p, sm_count, stream);
    } else if (g_kernel_gen != 1 && d <= 5) {
        return launch_dual(d, p, sm_count, stream, g_kernel_gen == 3);
    }
    return launch_gen1(d, p, sm_count, stream);
#else
    return cudaErrorInvalidValue;
#endif
}

// ---------------------------------------------------------------------------------------------
// Residual fine-tune mixer: the complex NCO RadioHandlerClass applies to every output
// slot right after the operator (Core/RadioHandler.cpp:33-37 -> shift_limited_unroll_C_sse_inp_c,
// Core/pffft/pf_mixer.cpp:791-856).  The reference advances four start phasors once per 128 samples
// (multiply by a table entry, renormalise) and derives the phasor of sample group g inside the block as
// table[g-1] * start; `starts` holds the start phasors of every 128-sample block of this launch (the
// sequential recurrence is run on the host, r2iq_capi.cu), so every sample is independent here.  Products
// are formed with separately rounded multiplies and add/sub exactly like the reference's SSE code (its
// function clones carry no FMA target, Core/pffft/fmv.h:9): the result is bit-identical to the CPU mixer.
// ---------------------------------------------------------------------------------------------
__global__ void nco_mix_kernel(float2* __restrict__ iq, long long n, const float2* __restrict__ starts,
                               const float2* __restrict__ tab) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const float2 s = __ldg(&starts[(i >> 7) * 4 + (i & 3)]);
        const int g = (int)((i & 127) >> 2);
        float2 ph = s;
        if (g) {
            const float2 t = __ldg(&tab[g - 1]);
            ph.x = __fsub_rn(__fmul_rn(t.x, s.x), __fmul_rn(t.y, s.y));
            ph.y = __fadd_rn(__fmul_rn(t.y, s.x), __fmul_rn(t.x, s.y));
        }
        const float2 v = iq[i];
        iq[i] = make_float2(__fsub_rn(__fmul_rn(v.x, ph.x), __fmul_rn(v.y, ph.y)),
                            __fadd_rn(__fmul_rn(v.y, ph.x), __fmul_rn(v.x, ph.y)));
    }
}

// Direct-phase mixer as a stand-alone in-place pass: one accurate phasor per thread and trip, then 4 samples by
// multiplying with exp(2 pi j fc k), k = 1..3 (rounding of the order of the float32 ulp).
__global__ void nco_mix_direct_kernel(float2* __restrict__ iq, long long n, double phase0, double fc) {
    float2 step[3];
#pragma unroll
    for (int k = 0; k < 3; k++) step[k] = nco_direct_phasor(0.0, fc, k + 1);
    long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 4;
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    for (; i < n; i += stride) {
        const float2 ph0 = nco_direct_phasor(phase0, fc, i);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (i + k >= n) break;
            float2 ph = ph0;
            if (k) ph = make_float2(ph0.x * step[k - 1].x - ph0.y * step[k - 1].y, ph0.x * step[k - 1].y + ph0.y * step[k - 1].x);
            const float2 v = iq[i + k];
            iq[i + k] = make_float2(v.x * ph.x - v.y * ph.y, v.x * ph.y + v.y * ph.x);
        }
    }
}

cudaError_t launch_nco_mix_direct(float* iq, long long n_complex, double phase0, double fc, cudaStream_t stream) {
    if (n_complex <= 0) return cudaSuccess;
    long long blocks = (n_complex / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    nco_mix_direct_kernel<<<(unsigned)blocks, 256, 0, stream>>>((float2*)iq, n_complex, phase0, fc);
    return cudaGetLastError();
}

cudaError_t launch_nco_mix(float* iq, long long n_complex, const float* starts, const float* tab, cudaStream_t stream) {
    if (n_complex <= 0) return cudaSuccess;
    long long blocks = (n_complex + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    nco_mix_kernel<<<(unsigned)blocks, 256, 0, stream>>>((float2*)iq, n_complex, (const float2*)starts, (const float2*)tab);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// the bit-exact stage on its own (parity hook): convert_float<rand> of Core/fft_mt_r2iq.h:36-51
// ---------------------------------------------------------------------------------------------
__global__ void convert_kernel(const uint32_t* __restrict__ in, float2* __restrict__ out, long long n_words,
                               uint32_t rand_mul) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n_words; i += stride) {
        // the fused kernels' own conversion (dual_frame.cuh: unpack_words handles a word of frame A and one of frame B)
        c2 v;
        const uint32_t w = __ldg(in + i);
        dualf::unpack_words(w, w, rand_mul, v);
        float re, im, unused;
        upk(v.re, re, unused);
        upk(v.im, im, unused);
        out[i] = make_float2(re, im);
    }
}

cudaError_t launch_convert(const int16_t* in, float* out, long long n, int rand, cudaStream_t stream) {
    if (n & 1) return cudaErrorInvalidValue;
    const long long n_words = n / 2;
    if (n_words == 0) return cudaSuccess;
    long long blocks = (n_words + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    convert_kernel<<<(unsigned)blocks, 256, 0, stream>>>((const uint32_t*)in, (float2*)out, n_words,
                                                         rand ? 0xFFFEu : 0u);
    return cudaGetLastError();
}

}  // namespace r2iq
