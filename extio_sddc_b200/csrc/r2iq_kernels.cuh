// r2iq_kernels.cuh -- launch parameters shared by the kernels and the C-ABI layer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace r2iq {

// Sizes of the FFTN = 2048 path (north-star configuration).  Reference: Core/fft_mt_r2iq.h:18-19,
// Core/config.h:80-81.
constexpr int kFftN = 2048;                       // FFTN_R_ADC
constexpr int kHalfFft = kFftN / 2;               // halfFft: complex bins of the forward FFT
constexpr int kTransferSamples = 65536;           // int16 samples per input block
constexpr int kHop = 3 * kHalfFft / 2;            // frame hop inside a block (1536 samples)
constexpr int kFramesPerBlock = kTransferSamples / kHop + 1;   // fftPerBuf = 43
constexpr int kNDec = 7;                          // NDECIDX

struct FusedParams {
    const uint32_t* in_words;   // [halfFft/2 halo words | n_blocks * 32768 words]; one word = 2 int16
    float2* out;                // n_blocks * (32768 >> d) complex
    long long total_frames;     // n_blocks * 43
    const float2* tw_fwd;       // [32][32]  W_1024^(k2*lane), forward inter-pass twiddles
    const float2* gk;           // [1024]    0.5*H[(k-tunebin) mod 1024] inside the band, else 0
    const float2* gck;          // [1024]    (-i*W_2048^k) * gk[k]
    const float2* gp;           // [1024]    gk + gck   (dual-frame kernel: Y = gp*a + gq*conj(b))
    const float2* gq;           // [1024]    gk - gck
                                // shipped kernels (r2iq_pair.cu), d >= 1: `gp` carries ONE float4 table (P, Q) in band order
                                // ([mfft]: position m holds bin tunebin + s(m)), `gq` is null
    const float2* tw_inv;       // d=0: [32][32] conj(W_1024)^((lane-tunebin)*k2) -- shipped kernels: in READER order, row n =
                                // twiddle of the element read from row n of the transposition; d>=1: [32][L] conj(W_M)^(n1*k2)
    const float2* tw_aux;       // d=1..4: [R][32] exp(+2 pi i lane k2 / mfft), R = 32>>d, then the pass-2 weights [32][32]
                                // exp(+2 pi i (lane % (32/R)) n1 / 32) -- shipped kernels: times exp(+2 pi i n1 (lane / G) / mfft),
                                // the folded pass-1 twiddle; d = 5, 6 (shipped kernels): [32 columns][32] float4 coefficients
                                // exp(-2 pi i n1 k / 1024) of the column's band bin and mirror bin (direct evaluation)
    int tunebin;
    int band_lo, band_hi;       // forward-FFT bins [lo,hi) that feed the inverse transform
    uint32_t row_mask;          // bit k1 set: spectrum row [32*k1, 32*k1+32) holds a band bin or the mirror of one
    uint32_t rand_mul;          // 0xFFFE when the ADC randomizer is on, else 0
    float imag_sign;            // -1 when sideband inversion (copy<true>) is requested, else +1
    unsigned long long* work_counter;   // device counter handing out frame pairs (dynamic work distribution)
    // Fine-tune NCO folded into the store epilogue (variant bit 5 of the default dual-frame kernels; null = off):
    // start phasors of every 128-sample block of this launch's output and the 33-entry group table, as nco_mix_kernel
    const float2* nco_starts;
    const float2* nco_tab;
    // Direct-phase form of the same mixer (R2IQ_NCO_DIRECT; null = off): output sample i of this launch is multiplied
    // by exp(2 pi j (nco_phase0 + nco_fc * i)); nco_rot[k] = exp(2 pi j nco_fc * 32 k), k < 32.  No host recurrence.
    const float2* nco_rot;
    double nco_phase0;          // cycles, in [0, 1)
    double nco_fc;              // cycles per output sample
};
// bit 5 of a kernel's kVar: multiply every output sample by the fine-tune phasor before it is stored
constexpr int kVarMix = 32;

#ifdef __CUDACC__
// Residual fine-tune mixer on one output sample (see nco_mix_kernel in r2iq_kernels.cu for the derivation): i is the
// sample's index in this launch's output stream.  Separately rounded products and sums, the order of the reference's
// SSE code (Core/pffft/pf_mixer.cpp:791-856), so a kernel that mixes in its store epilogue delivers the same bits as
// the stand-alone pass -- without reading and writing the IQ a second time.
__device__ __forceinline__ float2 nco_mix_sample(float2 v, long long i, const float2* __restrict__ starts,
                                                 const float2* __restrict__ tab) {
    const float2 s = __ldg(&starts[(i >> 7) * 4 + (i & 3)]);
    const int g = (int)((i & 127) >> 2);
    float2 ph = s;
    if (g) {
        const float2 t = __ldg(&tab[g - 1]);
        ph.x = __fsub_rn(__fmul_rn(t.x, s.x), __fmul_rn(t.y, s.y));
        ph.y = __fadd_rn(__fmul_rn(t.y, s.x), __fmul_rn(t.x, s.y));
    }
    return make_float2(__fsub_rn(__fmul_rn(v.x, ph.x), __fmul_rn(v.y, ph.y)),
                       __fadd_rn(__fmul_rn(v.y, ph.x), __fmul_rn(v.x, ph.y)));
}
// Direct-phase mixer: the phasor of output sample i, argument reduced in float64 (fc * i reaches 2^27 cycles within
// one launch), sine and cosine in float32.
__device__ __forceinline__ float2 nco_direct_phasor(double phase0, double fc, long long i) {
    double ph = fma(fc, (double)i, phase0);
    ph -= floor(ph);
    float s, c;
    sincospif((float)(2.0 * ph), &s, &c);
    return make_float2(c, s);
}
#endif
// Measurement variants of the persistent kernels (env R2IQ_VARIANT, read at every launch): bit 0 = dynamic work
// distribution (work_dispenser.cuh), bit 1 = no block barriers between phases (warps run free), bit 2 = TMA
// input lands in the transposition tile itself (12 instead of 8 warps per SM; d = 0 and d >= 2 kernels), bit 3 =
// inter-pass twiddles from a packed power recurrence instead of table loads (d = 0 kernel), bit 4 = mirror bins of the
// real-FFT split by warp shuffle instead of a shared-memory round trip (d = 0 kernel; measured: no gain).  Bit 5
// (kVarMix) is not a measurement switch: it selects the instantiation with the fine-tune mixer in the store epilogue
// and is set by the launcher when FusedParams::nco_starts is non-null.  Without the
// variable every family runs its fastest measured combination: 1 for d = 0 and d = 1, 7 for d = 2..5, bit 0 for
// the one-frame, FFTN 8192 and channelizer kernels.
int kernel_variant(int dflt = 1);

// Channelizer (r2iq_channels.cu): n_channels tuned channels from one input stream, d >= 1.
struct ChannelParams {
    const uint32_t* in_words;   // as FusedParams
    float2* out;                // channel c writes out[c*out_stride + ...], n_blocks*(32768>>d) complex each
    long long out_stride;       // complex samples between channels
    long long total_frames;
    const float2* tw_fwd;
    const float2* ctab;         // [1024]  -i * W_2048^k  (odd-part twiddle of the real-FFT split)
    const float2* gh;           // [mfft]  0.5 * H_d at inverse-FFT bin m (lower half m, upper half m - mfft)
    const float2* tw_inv;       // [32][L] as FusedParams for d >= 1
    const int* tunebins;        // [n_channels] device array, multiples of 4 in [0, halfFft)
    int n_channels;
    uint32_t rand_mul;
    float imag_sign;
    unsigned long long* work_counter;   // as FusedParams
};
cudaError_t setup_channel_kernels();
cudaError_t launch_channels(int d, const ChannelParams& p, int sm_count, cudaStream_t stream);

// FFTN = 8192 family (r2iq_wide_pair.cu; r2iq_wide.cu in comparison builds): halfFft 4096, 11 frames per block,
// hop 6144 samples; a team of four warps per frame pair.
struct WideParams {
    const uint32_t* in_words;   // [2048 halo words | n_blocks * 32768 words]
    float2* out;                // n_blocks * (32768 >> d) complex
    long long total_frames;     // n_blocks * 11
    const float2* tw1024;       // [32][32] W_1024^(k2*lane) (shared with the 2048 family)
    const float2* twq;          // [3][1024] W_4096^(q*n), q = 1..3
    const float2* ctab;         // [2049]    -i * W_8192^p
    const float2* g;            // [4096]    0.5 * H_d[(k - tunebin) mod 4096] inside the band, else 0
    const float2* tw_sub;       // d >= 1: [R][32] exp(+2 pi i lane k2 / S), then [32][32] exp(+2 pi i (lane % (32/R)) n1 / 32),
                                // R = S/32, S = the per-plane size mfft/4 (d = 1..5) or mfft = 64 (d = 6)
    const float4* pq;           // [4][1024] plane-major: (P, Q) of bin 4 m + r at [r][m]; Y[k] = P[k] Z[k] + Q[k] conj(Z[4096-k]),
                                // P = g (1 + c), Q = g (1 - c), zero outside the band (r2iq_wide_pair.cu)
    const float2* tw_inv;       // d = 0: [32][32] exp(+2 pi i ((l - tunebin/4) mod 1024) k2 / 1024), backward inter-pass twiddles
    uint32_t row_mask[4];       // per plane: rows of 32 slots that the band cut reads (band bins and their mirrors)
    int tunebin;
    uint32_t rand_mul;
    float imag_sign;
    unsigned long long* work_counter;   // as FusedParams
};
// Direct-phase fine-tune mixer in the store epilogue of the FFTN 8192 kernels (nco_rot null = off), as the fields of the
// same names in FusedParams; nco_rot has 128 entries here.  A kernel parameter of its own on purpose: with these three
// fields inside WideParams the mixer-less d = 1 kernel compiled to 40 more instructions and 88 bytes of spills
// (418 -> 396 Gsamples/s) -- code generation follows the size of the parameter block.
struct WideNco {
    const float2* nco_rot;
    double nco_phase0;
    double nco_fc;
};
cudaError_t setup_wide_kernels();
cudaError_t launch_wide(int d, const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream);

// FFTN = 8192 channelizer at d = 4, 5 (r2iq_wide_channels.cu): one forward transform and split per frame pair, every channel
// served from it.
struct WideChannelParams {
    const uint32_t* in_words;   // [2048 halo words | n_blocks * 32768 words]
    float2* out;                // channel c writes out[c*out_stride + ...], n_blocks*(32768>>d) complex each
    long long out_stride;
    long long total_frames;     // n_blocks * 11
    const float2* tw1024;       // [32][32] W_1024^(k2*lane)
    const float2* twq;          // [1024]   W_4096^n
    const float2* ctab;         // [4096]   -i * W_8192^k
    const float2* gh;           // [mfft]   0.5 * H_d at inverse-FFT bin m (lower half m, upper half m - mfft)
    const int* tunebins;        // [n_channels] device array, multiples of 4 in [0, 4096)
    int n_channels;
    uint32_t rand_mul;
    float imag_sign;
    unsigned long long* work_counter;
};
cudaError_t setup_wide_channel_kernels();
cudaError_t launch_wide_channels(int d, const WideChannelParams& p, int sm_count, cudaStream_t stream);   // d = 4, 5
const char* wide_kernel_name(int d, bool direct_mixer = false);   // as profilers print it
cudaError_t setup_wide_gen1_kernels();               // r2iq_wide.cu, comparison builds only
cudaError_t launch_wide_gen1(int d, const WideParams& p, int sm_count, cudaStream_t stream);
int kernel_generation();

const char* fused_kernel_name(int d, bool direct_mixer);
// Launch the fused kernel for decimation index d on `stream`.  Returns cudaGetLastError().
cudaError_t launch_fused(int d, const FusedParams& p, int sm_count, cudaStream_t stream);
// Kernel generation selector for A/B measurements (env R2IQ_KERNEL_GEN): 0 = the fastest measured kernel
// per decimation (default, see launch_fused); 1..4 = older or alternative families kept for comparison.
// All meet the parity bar.
void set_kernel_generation(int gen);
// first-generation kernels (r2iq_gen1.cu; comparison builds only, like the four families below)
cudaError_t setup_gen1_kernels();
cudaError_t launch_gen1(int d, const FusedParams& p, int sm_count, cudaStream_t stream);
// second-generation kernels (r2iq_dual.cu): two frames per warp on packed fp32x2 lanes + TMA input
cudaError_t setup_dual_kernels();
cudaError_t launch_dual(int d, const FusedParams& p, int sm_count, cudaStream_t stream, bool tma);
// dual-frame kernels with packed backward transform for d = 1..5 (r2iq_dual_dn.cu)
cudaError_t setup_dual_dn_kernels();
cudaError_t launch_dual_dn(int d, const FusedParams& p, int sm_count, cudaStream_t stream);
// specialised d = 1 dual-frame kernel (r2iq_dual_d1.cu)
cudaError_t setup_dual_d1_kernels();
cudaError_t launch_dual_d1(const FusedParams& p, int sm_count, cudaStream_t stream, bool tma);
// specialised d = 0 dual-frame kernel (r2iq_dual_d0.cu)
cudaError_t setup_dual_d0_kernels();
cudaError_t launch_dual_d0(const FusedParams& p, int sm_count, cudaStream_t stream, bool tma);
// third-generation dual-frame kernels (r2iq_pair.cu): one CTA of `warps` (8 or 12) free-running warps per SM, tables in
// shared memory, one TMA copy per frame pair; d = 0..5 (12 warps: d >= 1 only).  Mixes when p.nco_starts is set.
cudaError_t setup_pair_kernels();
cudaError_t launch_pair(int d, int warps, const FusedParams& p, int sm_count, cudaStream_t stream);
// One-time per-device setup (opt-in shared memory sizes).
cudaError_t setup_fused_kernels();
// Fine-tune NCO applied in place to n_complex output samples (see nco_mix_kernel).
cudaError_t launch_nco_mix(float* iq, long long n_complex, const float* starts, const float* tab, cudaStream_t stream);
// The direct-phase form as a stand-alone pass (kernel families without a mixing epilogue: FFTN 8192).
cudaError_t launch_nco_mix_direct(float* iq, long long n_complex, double phase0, double fc, cudaStream_t stream);
// int16 -> float with optional de-randomisation (the bit-exact stage, exposed for parity tests).
cudaError_t launch_convert(const int16_t* in, float* out, long long n, int rand, cudaStream_t stream);

}  // namespace r2iq
