// r2iq_common.cuh -- pieces shared by the one-frame (r2iq_kernels.cu) and dual-frame (r2iq_dual.cu)
// fused kernels: geometry per decimation, output placement, and the batched backward transform
// for d >= 1.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dft_regs.cuh"
#include "r2iq_kernels.cuh"

namespace r2iq {

constexpr int kScratch = 32 * 33;   // float2 per warp: 32x32 transposition tile, row pitch 33
constexpr int kStage = 64 * 17;     // float2 per warp: staged bands of 2^d frames (max over d)

template <int D>
struct Geo {
    static constexpr int M = kHalfFft >> D;                  // mfft: inverse FFT size
    static constexpr int ROUNDS = 1 << D;                    // frames a warp runs per unit of work
    static constexpr int L = (D <= 5) ? (32 >> D) : 1;       // lanes per frame in the inverse transform
    static constexpr int GS = (D == 6) ? 17 : (M + L);       // staging pitch per frame (bank-conflict free)
    static constexpr int OUT_PER_BLOCK = (kTransferSamples / 2) >> D;
    static constexpr int WARP_SMEM = kScratch + (D > 0 ? kStage : 0);   // float2 per warp
};

__device__ __forceinline__ int brev(int v, int bits) { return (int)(__brev((unsigned)v) >> (32 - bits)); }

// De-randomise (Core/fft_mt_r2iq.h:41-44: `val = input ^ (-2)` when the LSB is set) both int16
// halves of a word at once, then convert each to float (exact).
__device__ __forceinline__ float2 unpack_pair(uint32_t w, uint32_t rand_mul) {
    w ^= (w & 0x00010001u) * rand_mul;
    return make_float2((float)(short)(w & 0xFFFFu), (float)((int)w >> 16));
}

struct FrameOut {
    long long base;   // out[base + idx] receives inverse-FFT sample idx
    int lo, hi;       // kept range of idx
};

// Output placement of frame f = block*43 + k (impl.hpp:117-138): k == 0 keeps y[mfft/4, 3mfft/4),
// k >= 1 keeps y[0, 3mfft/4) at offset mfft/2 + (k-1)*3mfft/4 inside the block's output.
template <int D>
__device__ __forceinline__ FrameOut frame_out(long long f) {
    constexpr int M = Geo<D>::M;
    long long b = f / kFramesPerBlock;
    int k = (int)(f - b * kFramesPerBlock);
    FrameOut o;
    o.lo = k ? 0 : M / 4;
    o.hi = 3 * M / 4;
    o.base = b * (long long)Geo<D>::OUT_PER_BLOCK + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

__device__ __forceinline__ void store_iq(float2* out, const FrameOut& fo, int idx, float2 v, float imag_sign) {
    if (idx >= fo.lo && idx < fo.hi) out[fo.base + idx] = make_float2(v.x, v.y * imag_sign);
}


// Where the backward transform of staged band g goes: output pointer, kept range and liveness.
struct Placement {
    float2* out;
    FrameOut fo;
    bool live;
};

// Coalesced output for the small transforms (d >= 5), one band per lane: lane g deposits the kept samples
// of its band (value(idx), idx in [lo, hi)) and the band's destination in shared memory; the warp then
// writes all 32 bands as runs of consecutive lanes.  `vals` needs 32*(KEEP+1) float2, `S` 32*16 bytes.
template <int D, typename ValueFn>
__device__ __forceinline__ void staged_store(int lane, float2* S, float2* vals, const Placement& pl, float imag_sign,
                                             ValueFn value) {
    constexpr int M = Geo<D>::M;
    constexpr int KEEP = 3 * M / 4;          // 24 (d = 5) or 12 (d = 6); frame 0 of a block keeps M/2 of them
    constexpr int PITCH = KEEP + 1;          // odd: conflict-free deposits
    float2** dst = reinterpret_cast<float2**>(S);
    int* keep = reinterpret_cast<int*>(S + 32);
    __syncwarp();
    dst[lane] = pl.out + pl.fo.base + pl.fo.lo;
    keep[lane] = pl.live ? (pl.fo.hi - pl.fo.lo) : 0;
#pragma unroll
    for (int r = 0; r < KEEP; r++) {
        // frame 0 of a block keeps [M/4, 3M/4): its r-th kept sample is idx M/4 + r; others keep [0, 3M/4)
        const float2 a = value(r);
        const float2 b = value(r + M / 4);
        const float2 v = pl.fo.lo ? b : a;
        vals[lane * PITCH + r] = make_float2(v.x, v.y * imag_sign);
    }
    __syncwarp();
#pragma unroll
    for (int it = 0; it < KEEP; it++) {
        const int item = it * 32 + lane;
        const int band = item / KEEP;
        const int r = item - band * KEEP;
        if (r < keep[band]) dst[band][r] = vals[band * PITCH + r];
    }
    __syncwarp();
}

// Backward transforms of 2^D staged bands (band g at stage + g*GS, bins in inverse-FFT order), 32>>D
// lanes per band and 32 points per lane; overlap-save discard, optional conjugate and placement as
// impl.hpp:117-138.  `place(g)` says where band g goes: consecutive frames of one channel in the
// fused kernels, one frame of consecutive channels in the channelizer.
// S is a float2[kScratch] transposition tile private to the warp.
// kStaged selects the coalescing output path for d >= 5 (measured: it wins when bands are far apart in
// memory -- the channelizer, 3.95 -> 1.71 ms for 64 channels -- and loses ~13 % when they are consecutive
// frames of one channel, whose 8-byte stores already land in adjacent sectors).
// NB (a power of two <= 2^D) is the number of bands actually staged: with NB < 2^D the lane groups
// beyond NB recompute band (g mod NB) and store nothing.
template <int D, bool kStaged = false, int NB = (1 << D), typename PlaceFn>
__device__ __forceinline__ void inverse_bands(int lane, float2* S, const float2* stage, const float2* tw_inv,
                                              float imag_sign, PlaceFn place) {
    using G = Geo<D>;
    static_assert(D >= 1, "d = 0 keeps its spectrum in registers");
    __syncwarp();
    if constexpr (D <= 5) {
        constexpr int L = G::L;
        constexpr int NI = 32 / L;
        constexpr int LB = ilog2(L);
        const int g = lane / L;
        const int n = lane % L;
        float2 t[32];
#pragma unroll
        for (int n2 = 0; n2 < 32; n2++) t[brev(n2, 5)] = stage[(g & (NB - 1)) * G::GS + n + L * n2];
        __syncwarp();
        dit<32, true>(t);                        // pass 1: DFT over n2 within the band
        Placement pl = place(g);
        pl.live = pl.live && g < NB;
        if constexpr (L > 1) {
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) t[k2] = cmul(t[k2], __ldg(&tw_inv[k2 * L + n]));
            float2* Sg = S + g * (33 * L);
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) Sg[n * 33 + k2] = t[k2];
            __syncwarp();
#pragma unroll
            for (int i = 0; i < NI; i++) {
                float2 x[L];
#pragma unroll
                for (int n1 = 0; n1 < L; n1++) x[brev(n1, LB)] = Sg[n1 * 33 + n + L * i];
                dit<L, true>(x);                 // pass 2: DFT over n1
                if (pl.live) {
#pragma unroll
                    for (int k1 = 0; k1 < L; k1++) store_iq(pl.out, pl.fo, n + L * i + 32 * k1, x[k1], imag_sign);
                }
            }
            __syncwarp();
        } else {
            // d = 5: lane g holds all 32 samples of band g.  Writing them straight out would issue 8-byte
            // stores 192 bytes apart (one 32-byte sector transaction per lane); instead the kept samples go
            // through shared memory and leave as runs of consecutive lanes (full sectors).
            if constexpr (kStaged) {
                staged_store<D>(lane, S, const_cast<float2*>(stage), pl, imag_sign, [&](int idx) { return t[idx]; });
            } else if (pl.live) {
#pragma unroll
                for (int k2 = 0; k2 < 32; k2++) store_iq(pl.out, pl.fo, k2, t[k2], imag_sign);
            }
        }
    } else {
        // d = 6: mfft = 16, 64 bands per unit, two whole 16-point transforms per lane
        float2 x0[16], x1[16];
#pragma unroll
        for (int n = 0; n < 16; n++) {
            x0[brev(n, 4)] = stage[(lane & (NB - 1)) * G::GS + n];
            x1[brev(n, 4)] = stage[((lane + 32) & (NB - 1)) * G::GS + n];
        }
        __syncwarp();
        dit<16, true>(x0);
        dit<16, true>(x1);
        if constexpr (kStaged) {
            staged_store<D>(lane, S, const_cast<float2*>(stage), place(lane), imag_sign, [&](int idx) { return x0[idx]; });
            staged_store<D>(lane, S, const_cast<float2*>(stage), place(lane + 32), imag_sign, [&](int idx) { return x1[idx]; });
        } else {
            const Placement p0 = place(lane), p1 = place(lane + 32);
#pragma unroll
            for (int n = 0; n < 16; n++) {
                if (p0.live && lane < NB) store_iq(p0.out, p0.fo, n, x0[n], imag_sign);
                if (p1.live && lane + 32 < NB) store_iq(p1.out, p1.fo, n, x1[n], imag_sign);
            }
        }
    }
}

// The fused kernels' use: band g is frame f0 + g of the one tuned channel.
template <int D>
__device__ __forceinline__ void inverse_batch(const FusedParams& p, long long f0, int lane, float2* S,
                                              const float2* stage) {
    inverse_bands<D>(lane, S, stage, p.tw_inv, p.imag_sign, [&](int g) {
        Placement pl;
        pl.out = p.out;
        pl.fo = frame_out<D>(f0 + g);
        pl.live = f0 + g < p.total_frames;
        return pl;
    });
}

}  // namespace r2iq
