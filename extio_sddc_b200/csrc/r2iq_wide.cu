// r2iq_wide.cu -- COMPARISON BUILDS ONLY (R2IQ_AB_VARIANTS; the shipped FFTN 8192 kernels are r2iq_wide_pair.cu).
// FFTN = 8192 (the value the reference tree ships with, Core/config.h:49), second
// generation: FOUR WARPS SHARE ONE FRAME.  halfFft = 4096, hop 6144 samples, 11 frames per 65536-sample block,
// inverse sizes mfft = 4096 >> d; worker-loop semantics of Core/fft_mt_r2iq_impl.hpp:15-151.
//
// A first wide kernel (one frame per warp, 41 KB of shared memory each; in the history of this file's
// earlier revisions) fitted five warps on an SM and sat on load latency (ncu: 1.27 warps per scheduler,
// FMA pipe 20 %, 98-239 Gsamples/s).  Here a CTA of four warps owns one frame
// and its 4096-bin half spectrum (four skewed planes of 1060 float2 = 33.9 KB), four CTAs per SM:
//   stage A   radix-4 decimation-in-frequency stage straight from global memory, 128 threads wide
//             (coalesced, every sample converted once), twiddle W_4096^(q n), result to plane q;
//   stage B   warp q runs the 1024-point register FFT on plane q, transposing through the plane itself
//             (all 1024 values are in registers by then); output bin 4k'+q lands at plane q, slot k';
//   split     real-FFT split + filter in place on mirror pairs (k, 4096-k), 128 threads wide;
//   backward  decimation in TIME, mfft = 4 x (mfft/4): the tune bin is a multiple of 4
//             (Core/fft_mt_r2iq.cpp:101-109), so the bins Y[4m'+r] of the selected band are exactly plane r,
//             rotated by tunebin/4: warp r runs the (mfft/4)-point inverse transform in place on plane r
//             (1024-point register FFT for d = 0, the R = mfft/128 register scheme of r2iq_dual_dn.cu for
//             d = 1..5), applies W_mfft^(-r n'), and a last radix-4 stage, 128 threads wide, combines the four
//             planes and writes y[n' + (mfft/4) a] for a = 0..2 (a = 3 is always in the discarded quarter).
//   d = 6     mfft = 64 is too small to split: warp 0 transforms the band alone.
#include "r2iq_common.cuh"
#include "work_dispenser.cuh"

namespace r2iq {

namespace {

constexpr int kQHalf = 4096;                       // halfFft
constexpr int kQHopWords = 3 * kQHalf / 2 / 2;     // 6144 samples = 3072 words
constexpr int kQFrames = kTransferSamples / (3 * kQHalf / 2) + 1;   // fftPerBuf = 11
constexpr int kQPlane = 1060;                      // float2 per plane: >= 32*33 (transposition tile), == 4 mod 32 (skew)
constexpr int kQThreads = 128;
constexpr int kQCtasPerSm = 4;
constexpr int kQSmemBytes = 4 * kQPlane * (int)sizeof(float2);

__device__ __forceinline__ int qaddr(int k) { return (k & 3) * kQPlane + (k >> 2); }

template <int BITS>
__device__ __forceinline__ constexpr int brev_c(int v) {
    int r = 0;
    for (int i = 0; i < BITS; i++) r |= ((v >> i) & 1) << (BITS - 1 - i);
    return r;
}

// 1024-point transform of x[lane + 32 j] (v[brev(j)] on entry) -> X[lane + 32 k1] (v[k1] on return): two
// radix-32 register passes around one transposition through T (32 rows, pitch 33).
template <bool INV>
__device__ __forceinline__ void fft1024q(float2 (&v)[32], float2* T, const float2* tw1024, int lane) {
    dit<32, INV>(v);
    // inter-pass twiddles W_1024^(lane k2), k2 = 8a + b: ten table loads (b = 1..7, 8a = 8, 16, 24) and 21
    // products instead of 31 loads -- the L1 data pipe is the busiest unit of this kernel, the FMA pipe is not
    float2 wb[8], wa[4];
#pragma unroll
    for (int b = 1; b < 8; b++) wb[b] = __ldg(&tw1024[b * 32 + lane]);
#pragma unroll
    for (int a = 1; a < 4; a++) wa[a] = __ldg(&tw1024[8 * a * 32 + lane]);
#pragma unroll
    for (int k2 = 1; k2 < 32; k2++) {
        const int a = k2 >> 3, b = k2 & 7;
        const float2 w = a == 0 ? wb[b] : (b == 0 ? wa[a] : cmul(wa[a], wb[b]));
        v[k2] = INV ? cmul_conj(v[k2], w) : cmul(v[k2], w);   // conjugate for the inverse
    }
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) T[lane * 33 + k2] = v[k2];
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) v[brev_c<5>(n1)] = T[n1 * 33 + lane];
    __syncwarp();
    dit<32, INV>(v);
}

template <int D>
__device__ __forceinline__ FrameOut quad_frame_out(long long f) {
    constexpr int M = kQHalf >> D;
    const long long b = f / kQFrames;
    const int k = (int)(f - b * kQFrames);
    FrameOut o;
    o.lo = k ? 0 : M / 4;
    o.hi = 3 * M / 4;
    o.base = b * (long long)((kTransferSamples / 2) >> D) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

}  // namespace

// kDyn: frames from the device-wide counter instead of a fixed stride (work_dispenser.cuh; one frame per CTA)
template <int D, bool kDyn = false>
__global__ void __launch_bounds__(kQThreads, kQCtasPerSm) r2iq_wide_kernel(const WideParams p) {
    constexpr int M = kQHalf >> D;
    extern __shared__ float2 Z[];
    const int t = threadIdx.x;
    const int lane = t & 31;
    const int warp = t >> 5;
    float2* Zw = Z + warp * kQPlane;                   // the plane this warp transforms
    const int tb = p.tunebin;
    const int tb4 = tb >> 2;

    __shared__ long long s_base[3];
    PairDispenser<kDyn, false, 1> work;
    work.init(p.work_counter, s_base, lane, warp);
    for (; work.more(p.total_frames); work.advance()) {
        work.top(p.work_counter, s_base);
        const long long f = work.cur(0);
        // ---------------- stage A: radix-4 DIF over the four quarters of the frame ---------------------------
        const long long b = f / kQFrames;
        const int kf = (int)(f - b * kQFrames);
        const uint32_t* src = p.in_words + b * (kTransferSamples / 2) + kf * kQHopWords + t;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int n = t + 128 * j;
            const float2 z0 = unpack_pair(__ldg(src + 128 * j), p.rand_mul);
            const float2 z1 = unpack_pair(__ldg(src + 128 * j + 1024), p.rand_mul);
            const float2 z2 = unpack_pair(__ldg(src + 128 * j + 2048), p.rand_mul);
            const float2 z3 = unpack_pair(__ldg(src + 128 * j + 3072), p.rand_mul);
            const float2 a0 = make_float2(z0.x + z2.x, z0.y + z2.y), a1 = make_float2(z0.x - z2.x, z0.y - z2.y);
            const float2 a2 = make_float2(z1.x + z3.x, z1.y + z3.y), a3 = make_float2(z1.x - z3.x, z1.y - z3.y);
            const float2 u0 = make_float2(a0.x + a2.x, a0.y + a2.y);
            const float2 u2 = make_float2(a0.x - a2.x, a0.y - a2.y);
            const float2 u1 = make_float2(a1.x + a3.y, a1.y - a3.x);      // a1 - i a3
            const float2 u3 = make_float2(a1.x - a3.y, a1.y + a3.x);      // a1 + i a3
            Z[0 * kQPlane + n] = u0;
            const float2 w1 = __ldg(&p.twq[0 * 1024 + n]);                // W_4096^n
            const float2 w2 = __ldg(&p.twq[1 * 1024 + n]);                // W_4096^(2 n)
            Z[1 * kQPlane + n] = cmul(u1, w1);
            Z[2 * kQPlane + n] = cmul(u2, w2);
            Z[3 * kQPlane + n] = cmul(u3, cmul(w1, w2));                  // W_4096^(3 n)
        }
        __syncthreads();

        // ---------------- stage B: warp q transforms plane q in place ---------------------------------------
        {
            float2 v[32];
#pragma unroll
            for (int j = 0; j < 32; j++) v[brev_c<5>(j)] = Zw[lane + 32 * j];
            __syncwarp();
            fft1024q<false>(v, Zw, p.tw1024, lane);
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) Zw[lane + 32 * k1] = v[k1];   // bin 4 (lane + 32 k1) + warp
        }
        __syncthreads();

        // ---------------- real-FFT split + filter, in place on mirror pairs ------------------------------
#pragma unroll 4
        for (int i = 0; i <= 16; i++) {
            const int pp = i < 16 ? t + 128 * i : 2048;
            if (i < 16 || t == 0) {
                const int pm = (kQHalf - pp) & (kQHalf - 1);
                if constexpr (D > 0) {      // only bins of the selected band are read by the backward transform
                    const bool inp = ((pp - tb + M / 2) & (kQHalf - 1)) < M, inm = ((pm - tb + M / 2) & (kQHalf - 1)) < M;
                    if (!(inp || inm)) continue;
                }
                const float2 a = Z[qaddr(pp)];
                const float2 bq = Z[qaddr(pm)];
                const float2 c = __ldg(&p.ctab[pp]);                 // -i * W_8192^p
                const float2 gp = __ldg(&p.g[pp]);
                const float2 gm = __ldg(&p.g[pm]);
                const float ux = a.x + bq.x, uy = a.y - bq.y;         // u = a + conj b
                const float wx = a.x - bq.x, wy = a.y + bq.y;         // w = a - conj b
                const float tx = c.x * wx - c.y * wy, ty = c.x * wy + c.y * wx;
                const float2 xp = make_float2(ux + tx, uy + ty);      // 2 X[p]
                const float2 xm = make_float2(ux - tx, -(uy - ty));   // 2 X[4096-p] = conj(u - t)
                Z[qaddr(pp)] = cmul(xp, gp);
                if (pp != 0 && pm != pp) Z[qaddr(pm)] = cmul(xm, gm);
            }
        }
        __syncthreads();

        // ---------------- backward transform, discard, placement ------------------------------------------
        const FrameOut fo = quad_frame_out<D>(f);
        if constexpr (D <= 5) {
            constexpr int MS = M / 4;                                 // points per plane: 1024 >> D
            if constexpr (D == 0) {
                float2 v[32];
#pragma unroll
                for (int j = 0; j < 32; j++) v[brev_c<5>(j)] = Zw[(tb4 + lane + 32 * j) & 1023];
                __syncwarp();
                fft1024q<true>(v, Zw, p.tw1024, lane);
                if (warp) {
#pragma unroll
                    for (int k1 = 0; k1 < 32; k1++)                   // W_4096^(-warp n')
                        v[k1] = cmul_conj(v[k1], __ldg(&p.twq[(warp - 1) * 1024 + lane + 32 * k1]));
                }
#pragma unroll
                for (int k1 = 0; k1 < 32; k1++) Zw[lane + 32 * k1] = v[k1];
            } else {
                constexpr int R = MS / 32;                            // band bins of this plane per lane
                constexpr int G = 32 / R;
                constexpr int RB = ilog2(R);
                float2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int m = lane + 32 * j;
                    y[brev_c<RB>(j)] = Zw[(tb4 + (m < MS / 2 ? m : m - MS)) & 1023];
                }
                __syncwarp();
                dit<R, true>(y);                                      // pass 1 over n2 (m' = lane + 32 n2)
#pragma unroll
                for (int k2 = 1; k2 < R; k2++) y[k2] = cmul(y[k2], __ldg(&p.tw_sub[k2 * 32 + lane]));   // exp(+2 pi i lane k2 / MS)
#pragma unroll
                for (int k2 = 0; k2 < R; k2++) Zw[lane * (R + 1) + k2] = y[k2];
                __syncwarp();
                const int k2p = lane / G, h = lane % G;
                float2 u[R];
#pragma unroll
                for (int n1 = 0; n1 < 32; n1++) {                     // u_j = sum_r x[j + R r] exp(+2 pi i h (j + R r) / 32)
                    const float2 x = Zw[n1 * (R + 1) + k2p];
                    const float2 w = __ldg(&p.tw_sub[(R + n1) * 32 + lane]);
                    float2& acc = u[brev_c<RB>(n1 % R)];
                    if (n1 < R) acc = cmul(x, w);
                    else acc = make_float2(fmaf(x.x, w.x, fmaf(-x.y, w.y, acc.x)), fmaf(x.x, w.y, fmaf(x.y, w.x, acc.y)));
                }
                __syncwarp();
                dit<R, true>(u);                                      // u[q] = S_warp[k2p + R h + 32 q]
#pragma unroll
                for (int q = 0; q < R; q++) {
                    const int n = k2p + R * h + 32 * q;
                    float2 s = u[q];
                    if (warp) s = cmul_conj(s, __ldg(&p.twq[(warp - 1) * 1024 + (n << D)]));            // W_mfft^(-warp n')
                    Zw[n] = s;
                }
            }
            __syncthreads();
            // last radix-4 stage across the planes: y[n' + MS a] = sum_r i^(r a) S_r[n'], a = 0..2 kept
#pragma unroll
            for (int i = 0; i < (MS + kQThreads - 1) / kQThreads; i++) {
                const int n = t + kQThreads * i;
                if (MS >= kQThreads || n < MS) {
                    const float2 s0 = Z[n], s1 = Z[kQPlane + n], s2 = Z[2 * kQPlane + n], s3 = Z[3 * kQPlane + n];
                    const float2 a0 = make_float2(s0.x + s2.x, s0.y + s2.y), a1 = make_float2(s0.x - s2.x, s0.y - s2.y);
                    const float2 a2 = make_float2(s1.x + s3.x, s1.y + s3.y), a3 = make_float2(s1.x - s3.x, s1.y - s3.y);
                    store_iq(p.out, fo, n, make_float2(a0.x + a2.x, a0.y + a2.y), p.imag_sign);
                    store_iq(p.out, fo, n + MS, make_float2(a1.x - a3.y, a1.y + a3.x), p.imag_sign);      // a1 + i a3
                    store_iq(p.out, fo, n + 2 * MS, make_float2(a0.x - a2.x, a0.y - a2.y), p.imag_sign);
                }
            }
        } else {
            // d = 6: mfft = 64, warp 0 alone (R = 2 register scheme over the whole band)
            if (warp == 0) {
                constexpr int R = M / 32, G = 32 / R, RB = ilog2(R);
                auto ysrc = [&](int m) { return qaddr((tb + (m < M / 2 ? m : m - M)) & (kQHalf - 1)); };
                float2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) y[brev_c<RB>(j)] = Z[ysrc(lane + 32 * j)];
                __syncwarp();
                dit<R, true>(y);
#pragma unroll
                for (int k2 = 1; k2 < R; k2++) y[k2] = cmul(y[k2], __ldg(&p.tw_sub[k2 * 32 + lane]));
#pragma unroll
                for (int k2 = 0; k2 < R; k2++) Z[lane * (R + 1) + k2] = y[k2];
                __syncwarp();
                const int k2p = lane / G, h = lane % G;
                float2 u[R];
#pragma unroll
                for (int n1 = 0; n1 < 32; n1++) {
                    const float2 x = Z[n1 * (R + 1) + k2p];
                    const float2 w = __ldg(&p.tw_sub[(R + n1) * 32 + lane]);
                    float2& acc = u[brev_c<RB>(n1 % R)];
                    if (n1 < R) acc = cmul(x, w);
                    else acc = make_float2(fmaf(x.x, w.x, fmaf(-x.y, w.y, acc.x)), fmaf(x.x, w.y, fmaf(x.y, w.x, acc.y)));
                }
                dit<R, true>(u);
#pragma unroll
                for (int q = 0; q < R; q++) store_iq(p.out, fo, k2p + R * h + 32 * q, u[q], p.imag_sign);
            }
        }
        __syncthreads();                                              // the planes are rewritten by the next frame
    }
}

template <int D, bool kDyn>
static cudaError_t setup_qv() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_wide_kernel<D, kDyn>, cudaFuncAttributeMaxDynamicSharedMemorySize, kQSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_wide_kernel<D, kDyn>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}
template <int D>
static cudaError_t setup_q() {
#ifdef R2IQ_AB_VARIANTS
    cudaError_t e = setup_qv<D, false>();      // fixed-stride form, comparison builds only
    if (e != cudaSuccess) return e;
#endif
    return setup_qv<D, true>();
}

cudaError_t setup_wide_gen1_kernels() {
    cudaError_t e;
    if ((e = setup_q<0>()) != cudaSuccess) return e;
    if ((e = setup_q<1>()) != cudaSuccess) return e;
    if ((e = setup_q<2>()) != cudaSuccess) return e;
    if ((e = setup_q<3>()) != cudaSuccess) return e;
    if ((e = setup_q<4>()) != cudaSuccess) return e;
    if ((e = setup_q<5>()) != cudaSuccess) return e;
    if ((e = setup_q<6>()) != cudaSuccess) return e;
    return cudaSuccess;
}

template <int D, bool kDyn>
static cudaError_t launch_qv(const WideParams& p, int sm_count, cudaStream_t stream) {
    long long grid = (long long)sm_count * kQCtasPerSm;
    if (grid > p.total_frames) grid = p.total_frames;
    if (grid < 1) grid = 1;
    if constexpr (kDyn) {
        cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    r2iq_wide_kernel<D, kDyn><<<(unsigned)grid, kQThreads, kQSmemBytes, stream>>>(p);
    return cudaGetLastError();
}
template <int D>
static cudaError_t launch_q(const WideParams& p, int sm_count, cudaStream_t stream) {
#ifdef R2IQ_AB_VARIANTS
    if (!(kernel_variant() & 1)) return launch_qv<D, false>(p, sm_count, stream);
#endif
    return launch_qv<D, true>(p, sm_count, stream);
}

cudaError_t launch_wide_gen1(int d, const WideParams& p, int sm_count, cudaStream_t stream) {
    switch (d) {
        case 0: return launch_q<0>(p, sm_count, stream);
        case 1: return launch_q<1>(p, sm_count, stream);
        case 2: return launch_q<2>(p, sm_count, stream);
        case 3: return launch_q<3>(p, sm_count, stream);
        case 4: return launch_q<4>(p, sm_count, stream);
        case 5: return launch_q<5>(p, sm_count, stream);
        case 6: return launch_q<6>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
