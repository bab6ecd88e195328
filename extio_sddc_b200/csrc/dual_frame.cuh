// dual_frame.cuh -- device helpers shared by the dual-frame (two frames per warp on packed fp32x2 lanes) kernels:
// mbarrier / TMA bulk-copy wrappers, the two-plane shared-memory tile accessors, the de-randomise + int16 -> float
// unpack of two frames at once, frame addressing and output placement.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dft_regs2.cuh"
#include "r2iq_kernels.cuh"

namespace r2iq {
namespace dualf {

constexpr int kTile = 32 * 33;                // elements per plane: 32x32 tile, pitch 33 (conflict-free for 8 B accesses)
constexpr int kFrameBytes = kFftN * 2;        // 4096 B of int16 per frame

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t mbar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(mbar), "r"(parity)
            : "memory");
    } while (!done);
}
// TMA 1-D bulk copy global -> shared, completion signalled on the mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(mbar)
                 : "memory");
}

// first word of frame f = 43*block + k inside [halo | blocks] (hop 1536 samples = 768 words)
// (frame indices fit 32 bits: the launchers refuse captures of 2^31 frames or more -- 6.5 TB of samples -- so the
// division by 43 is one multiply-high instead of a 64-bit sequence)
template <class Params>
__device__ __forceinline__ const uint32_t* frame_src(const Params& p, long long f) {
    const uint32_t b = (uint32_t)f / (uint32_t)kFramesPerBlock;
    const uint32_t k = (uint32_t)f - b * (uint32_t)kFramesPerBlock;
    return p.in_words + (size_t)b * (kTransferSamples / 2) + k * (kHop / 2);
}

__device__ __forceinline__ int brev5(int v) { return (int)(__brev((unsigned)v) >> 27); }
__device__ __forceinline__ int brev_n(int v, int bits) { return bits == 0 ? 0 : (int)(__brev((unsigned)v) >> (32 - bits)); }
__device__ __forceinline__ float flip(float x, uint32_t sgn) { return __uint_as_float(__float_as_uint(x) ^ sgn); }

// Shared-memory exchanges use two 64-bit planes (re pairs, im pairs) rather than one 128-bit element: a v2.b64
// access needs re and im in four consecutive registers (ptxas paid ~400 MOVs per iteration to arrange that); plane
// accesses have no such constraint and stay conflict-free at pitch 33.  Stored as float2 halves: a plain 64-bit
// store of the packed register makes ptxas copy the pair into a scratch pair first.
__device__ __forceinline__ void sts2(u64* S, int idx, c2 v) {
    float a, b, c, d;
    upk(v.re, a, b);
    upk(v.im, c, d);
    reinterpret_cast<float2*>(S)[idx] = make_float2(a, b);
    reinterpret_cast<float2*>(S)[kTile + idx] = make_float2(c, d);
}
__device__ __forceinline__ c2 lds2(const u64* S, int idx) {
    const float2 r = reinterpret_cast<const float2*>(S)[idx];
    const float2 i = reinterpret_cast<const float2*>(S)[kTile + idx];
    c2 v;
    v.re = pk(r.x, r.y);
    v.im = pk(i.x, i.y);
    return v;
}

// One word of frame A and one of frame B (two int16 each = one packed complex input point per frame):
// de-randomise both halves (Core/fft_mt_r2iq.h:41-44: x ^= 0xFFFE when the LSB is set), then int16 -> float.
// The conversion is the integer-pipe "magic number" form: 0x4B000000 | (x ^ 0x8000) is the float 2^23 + (x + 32768)
// exactly, so one exact packed subtraction of 2^23 + 32768 yields float(x) -- the same bits as I2F (every int16 is
// representable), without the half-rate conversion unit.
__device__ __forceinline__ void unpack_words(uint32_t wa, uint32_t wb, uint32_t rand_mul, c2& v) {
    wa ^= ((wa & 0x00010001u) * rand_mul) ^ 0x80008000u;
    wb ^= ((wb & 0x00010001u) * rand_mul) ^ 0x80008000u;
    const uint32_t are = __byte_perm(wa, 0x4B000000u, 0x7610), aim = __byte_perm(wa, 0x4B000000u, 0x7632);
    const uint32_t bre = __byte_perm(wb, 0x4B000000u, 0x7610), bim = __byte_perm(wb, 0x4B000000u, 0x7632);
    const u64 bias = bc(8421376.0f);   // 2^23 + 2^15
    v.re = sub2(pk(__uint_as_float(are), __uint_as_float(bre)), bias);
    v.im = sub2(pk(__uint_as_float(aim), __uint_as_float(bim)), bias);
}

// unpack_words without the final subtraction: v = float(x) + kUnpackBias in every component.  The first butterfly stage
// of the transform that follows removes the bias for less (dit2_biased): b' = a - b needs no correction at all, and
// a' = (a - 2 bias) + b is two exact additions -- 3 instead of 4 packed instructions per component and butterfly,
// 32 fewer per frame pair, same bits (every intermediate is an integer below 2^24).
constexpr float kUnpackBias = 8421376.0f;   // 2^23 + 2^15
__device__ __forceinline__ void unpack_words_biased(uint32_t wa, uint32_t wb, uint32_t rand_mul, c2& v) {
    wa ^= ((wa & 0x00010001u) * rand_mul) ^ 0x80008000u;
    wb ^= ((wb & 0x00010001u) * rand_mul) ^ 0x80008000u;
    const uint32_t are = __byte_perm(wa, 0x4B000000u, 0x7610), aim = __byte_perm(wa, 0x4B000000u, 0x7632);
    const uint32_t bre = __byte_perm(wb, 0x4B000000u, 0x7610), bim = __byte_perm(wb, 0x4B000000u, 0x7632);
    v.re = pk(__uint_as_float(are), __uint_as_float(bre));
    v.im = pk(__uint_as_float(aim), __uint_as_float(bim));
}
// dit2<N, INV> on values that all carry the additive bias of unpack_words_biased
template <int N, bool INV>
__device__ __forceinline__ void dit2_biased(c2 (&v)[N]) {
    static_assert(N >= 4, "first stage written out, the rest is dit2");
    const u64 m2b = bc(-2.0f * kUnpackBias);
#pragma unroll
    for (int i = 0; i < N / 2; i++) {
        c2& a = v[2 * i];
        c2& b = v[2 * i + 1];
        const c2 t = b;
        b.re = sub2(a.re, t.re);
        b.im = sub2(a.im, t.im);
        a.re = add2(add2(a.re, m2b), t.re);
        a.im = add2(add2(a.im, m2b), t.im);
    }
    dit2<N, INV, 4>(v);
}

struct FrameOut2 {
    long long base;   // out[base + idx] receives backward-transform sample idx of this frame
    int lo;           // first kept idx (the kept range ends at 3*mfft/4)
    bool live;        // false for the phantom second frame of an odd total
};
// impl.hpp:117-138: frame 0 of a block keeps y[mfft/4, 3mfft/4), the others y[0, 3mfft/4) at mfft/2 + (k-1)*3mfft/4
template <int D>
__device__ __forceinline__ FrameOut2 frame_out2(long long f, long long total) {
    constexpr int M = kHalfFft >> D;
    const uint32_t b = (uint32_t)f / (uint32_t)kFramesPerBlock;
    const int k = (int)((uint32_t)f - b * (uint32_t)kFramesPerBlock);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = (long long)((unsigned long long)b * (unsigned)((kTransferSamples / 2) >> D)) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

// Direct-phase mixer on the packed lanes: the phasors of the two frames' samples at output positions ia / ib ...
__device__ __forceinline__ c2 direct_phasors(double phase0, double fc, long long ia, long long ib) {
    const float2 pa = nco_direct_phasor(phase0, fc, ia);
    const float2 pb = nco_direct_phasor(phase0, fc, ib);
    c2 ph;
    ph.re = pk(pa.x, pb.x);
    ph.im = pk(pa.y, pb.y);
    return ph;
}
// ... advanced by `rot` (32 k samples further on: the lane's k-th stored row) and applied to t after the sideband
// conjugation (the mixer follows the operator, Core/RadioHandler.cpp:33-37): 9 packed instructions per row
__device__ __forceinline__ c2 direct_mix(c2 t, c2 ph0, float2 rot, float imag_sign) {
    const c2 ph = cmul2s(ph0, rot);
    const u64 ti = mul2(bc(imag_sign), t.im);
    c2 o;
    o.re = fma2(t.re, ph.re, neg2(mul2(ti, ph.im)));
    o.im = fma2(t.re, ph.im, mul2(ti, ph.re));
    return o;
}

// ---- 32-point inverse DFT across the LANES of a warp (d = 4, 5: mfft = 64, 32) -------------------------------------
// Decimation in frequency, five butterfly stages at lane distances 16, 8, 4, 2, 1: lane m holds X[m] on entry and
// x[brev5(lane)] = sum_m X[m] exp(+2 pi i m brev5(lane) / 32) on exit.  Per stage 4 SHFL (both frames' re and im) and 6
// packed instructions: t = partner + sg * own  (sg = -1 in the upper lane of a pair: a - b there, a + b in the lower),
// then t * w with w = 1 in the lower lane and exp(+2 pi i (lane mod dist) / (2 dist)) in the upper (last stage: w = 1).
// ~50 instructions instead of the ~210 of the 32-term multiply-accumulate through shared memory it replaces.
__device__ __forceinline__ c2 shfl_xor2(c2 v, int dist) {
    float a, b, c, d;
    upk(v.re, a, b);
    upk(v.im, c, d);
    a = __shfl_xor_sync(0xffffffffu, a, dist);
    b = __shfl_xor_sync(0xffffffffu, b, dist);
    c = __shfl_xor_sync(0xffffffffu, c, dist);
    d = __shfl_xor_sync(0xffffffffu, d, dist);
    return c2{pk(a, b), pk(c, d)};
}
// NS butterfly stages: a 2^NS-point transform inside every aligned group of 2^NS lanes (NS = 4: the 16-point transform
// of d = 6, run twice side by side; w[0..NS-2] are the lane's stage twiddles, distance 2^(NS-1) first)
template <int NS>
__device__ __forceinline__ c2 warp_idft(c2 x, const float2 (&w)[4], int lane) {
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const int dist = (1 << (NS - 1)) >> s;
        const c2 pr = shfl_xor2(x, dist);
        const u64 sg = bc((lane & dist) ? -1.0f : 1.0f);
        c2 t;
        t.re = fma2(sg, x.re, pr.re);
        t.im = fma2(sg, x.im, pr.im);
        x = s < NS - 1 ? cmul2s(t, w[s]) : t;
    }
    return x;
}
__device__ __forceinline__ c2 warp_idft32(c2 x, const float2 (&w)[4], int lane) { return warp_idft<5>(x, w, lane); }
// this lane's stage twiddles (entries 0..3) and, for mfft = 32 R > 32, exp(+2 pi i lane e / mfft), e = 1..R-1 (entries 4..)
__device__ __forceinline__ float2 warp_idft_twiddle(int lane, int e, int mfft) {
    float sn, cs;
    if (e >= 4) {
        sincospif((float)(2 * lane * (e - 3)) / (float)mfft, &sn, &cs);
        return make_float2(cs, sn);
    }
    const int dist = (mfft == 16 ? 8 : 16) >> e;     // mfft = 16: four stages inside each half warp
    if (dist == 0 || !(lane & dist)) return make_float2(1.0f, 0.0f);
    sincospif((float)(lane & (dist - 1)) / (float)dist, &sn, &cs);
    return make_float2(cs, sn);
}

}  // namespace dualf
}  // namespace r2iq
