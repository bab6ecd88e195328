// r2iq_wide.cu -- the fused r2iq kernel family for FFTN = 8192, the value the reference tree ships with
// (`#define FFTN_R_ADC (8192)`, Core/config.h:49; every size of the path derives from it,
// Core/fft_mt_r2iq.h:18-19): halfFft = 4096, frame hop 6144 samples, 11 frames per 65536-sample block,
// 1025 filter taps, inverse sizes mfft = 4096 >> d.  Same worker-loop semantics as the FFTN = 2048 family
// (Core/fft_mt_r2iq_impl.hpp:15-151); first-generation style (one frame per warp, scalar FP32).
//
// A warp keeps the 4096-bin half spectrum of its frame in shared memory (32 KB) and reuses the
// 1024-point two-pass register FFT of the 2048 family as a building block:
//   forward   4096 = 4 x 1024, decimation in frequency: one radix-4 stage on the four contiguous quarters
//             of the packed frame (coalesced loads, each sample converted once), twiddle W_4096^(q n),
//             then four 1024-point FFTs whose outputs are the bins 4k'+q;
//   split     real-FFT split + filter IN PLACE on mirror pairs (k, 4096-k): slot k then holds X[k]*G[k], with
//             G = 0.5*H_d[(k - tunebin)] inside the band and 0 outside (edge zero fill, impl.hpp:90-96);
//   backward  the tune rotation is addressing (bin m of the band lives in slot tunebin + s(m));
//             mfft = 4096 / 2048 (d = 0, 1): one in-place radix-4 / radix-2 DIF stage, then 1024-point
//             transforms whose outputs interleave in time; mfft <= 1024 (d >= 2): the R = mfft/32 register
//             scheme of r2iq_dual_dn.cu in scalar form.
// Spectrum storage is skewed -- bin k at (k & 3) * 1028 + (k >> 2) -- so that both the per-q planes written
// by the sub-FFTs and the consecutive-bin accesses of split and band cut are bank-conflict free.
#include "r2iq_common.cuh"

namespace r2iq {

namespace {

constexpr int kWHalf = 4096;                       // halfFft
constexpr int kWHopWords = 3 * kWHalf / 2 / 2;     // 6144 samples = 3072 words
constexpr int kWFrames = kTransferSamples / (3 * kWHalf / 2) + 1;   // fftPerBuf = 11
constexpr int kWPlane = 1028;                      // float2 per q-plane (1024 + skew)
constexpr int kWZ = 4 * kWPlane;                   // spectrum slots per warp
constexpr int kWWarpFloat2 = kWZ + kScratch;       // + the 32x33 transposition tile
constexpr int kWWarps = 5;                         // 5 x 41.3 KB = 207 KB: one CTA per SM
constexpr int kWThreads = kWWarps * 32;

__device__ __forceinline__ int zaddr(int k) { return (k & 3) * kWPlane + (k >> 2); }

// 1024-point transform of x[lane + 32 j] (v[brev(j)] on entry) -> X[lane + 32 k1] (v[k1] on return),
// two radix-32 register passes around one transposition through T (32x33 tile).
template <bool INV>
__device__ __forceinline__ void fft1024(float2 (&v)[32], float2* T, const float2* tw1024, int lane) {
    dit<32, INV>(v);
#pragma unroll
    for (int k2 = 1; k2 < 32; k2++) {
        const float2 w = __ldg(&tw1024[k2 * 32 + lane]);      // W_1024^(lane k2); conjugate for the inverse
        v[k2] = INV ? cmul_conj(v[k2], w) : cmul(v[k2], w);
    }
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) T[lane * 33 + k2] = v[k2];
    __syncwarp();
    float2 t[32];
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) t[brev(n1, 5)] = T[n1 * 33 + lane];
    __syncwarp();
    dit<32, INV>(t);
#pragma unroll
    for (int k1 = 0; k1 < 32; k1++) v[k1] = t[k1];
}

template <int D>
struct WGeo {
    static constexpr int M = kWHalf >> D;
    static constexpr int OUT_PER_BLOCK = (kTransferSamples / 2) >> D;
};

template <int D>
__device__ __forceinline__ FrameOut wide_frame_out(long long f) {
    constexpr int M = WGeo<D>::M;
    const long long b = f / kWFrames;
    const int k = (int)(f - b * kWFrames);
    FrameOut o;
    o.lo = k ? 0 : M / 4;
    o.hi = 3 * M / 4;
    o.base = b * (long long)WGeo<D>::OUT_PER_BLOCK + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

}  // namespace

template <int D>
__global__ void __launch_bounds__(kWThreads, 1) r2iq_wide_kernel(const WideParams p) {
    constexpr int M = WGeo<D>::M;
    extern __shared__ float2 smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float2* Z = smem + warp * kWWarpFloat2;
    float2* T = Z + kWZ;
    const int tb = p.tunebin;
    auto ysrc = [&](int m) { return zaddr((tb + (m < M / 2 ? m : m - M)) & (kWHalf - 1)); };

    for (long long f = (long long)blockIdx.x * kWWarps + warp; f < p.total_frames;
         f += (long long)gridDim.x * kWWarps) {
        // ---------------- forward: radix-4 DIF stage over the four quarters, each word converted once -------
        const long long b = f / kWFrames;
        const int kf = (int)(f - b * kWFrames);
        const uint32_t* src = p.in_words + b * (kTransferSamples / 2) + kf * kWHopWords + lane;
        float2 v[32];
#pragma unroll 8
        for (int j = 0; j < 32; j++) {
            const int n = lane + 32 * j;
            const float2 z0 = unpack_pair(__ldg(src + 32 * j), p.rand_mul);
            const float2 z1 = unpack_pair(__ldg(src + 32 * j + 1024), p.rand_mul);
            const float2 z2 = unpack_pair(__ldg(src + 32 * j + 2048), p.rand_mul);
            const float2 z3 = unpack_pair(__ldg(src + 32 * j + 3072), p.rand_mul);
            const float2 a0 = make_float2(z0.x + z2.x, z0.y + z2.y), a1 = make_float2(z0.x - z2.x, z0.y - z2.y);
            const float2 a2 = make_float2(z1.x + z3.x, z1.y + z3.y), a3 = make_float2(z1.x - z3.x, z1.y - z3.y);
            const float2 u0 = make_float2(a0.x + a2.x, a0.y + a2.y);
            const float2 u2 = make_float2(a0.x - a2.x, a0.y - a2.y);
            const float2 u1 = make_float2(a1.x + a3.y, a1.y - a3.x);      // a1 - i a3
            const float2 u3 = make_float2(a1.x - a3.y, a1.y + a3.x);      // a1 + i a3
            Z[0 * kWPlane + n] = u0;
            Z[1 * kWPlane + n] = cmul(u1, __ldg(&p.twq[0 * 1024 + n]));   // W_4096^(1 n)
            Z[2 * kWPlane + n] = cmul(u2, __ldg(&p.twq[1 * 1024 + n]));   // W_4096^(2 n)
            Z[3 * kWPlane + n] = cmul(u3, __ldg(&p.twq[2 * 1024 + n]));   // W_4096^(3 n)
        }
        __syncwarp();
#pragma unroll 1
        for (int q = 0; q < 4; q++) {
#pragma unroll
            for (int j = 0; j < 32; j++) v[brev(j, 5)] = Z[q * kWPlane + lane + 32 * j];
            __syncwarp();
            fft1024<false>(v, T, p.tw1024, lane);
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) Z[q * kWPlane + lane + 32 * k1] = v[k1];   // bin 4(lane+32k1)+q
        }
        __syncwarp();

        // ---------------- real-FFT split + filter, in place on mirror pairs ------------------------------
#pragma unroll 5
        for (int i = 0; i <= 64; i++) {
            const int pp = i < 64 ? lane + 32 * i : 2048;
            if (i < 64 || lane == 0) {
                const int pm = (kWHalf - pp) & (kWHalf - 1);
                const float2 a = Z[zaddr(pp)];
                const float2 bq = Z[zaddr(pm)];
                const float2 c = __ldg(&p.ctab[pp]);                 // -i * W_8192^p
                const float2 gp = __ldg(&p.g[pp]);
                const float2 gm = __ldg(&p.g[pm]);
                const float ux = a.x + bq.x, uy = a.y - bq.y;         // u = a + conj b
                const float wx = a.x - bq.x, wy = a.y + bq.y;         // w = a - conj b
                const float tx = c.x * wx - c.y * wy, ty = c.x * wy + c.y * wx;
                const float2 xp = make_float2(ux + tx, uy + ty);      // 2 X[p]
                const float2 xm = make_float2(ux - tx, -(uy - ty));   // 2 X[4096-p] = conj(u - t)
                Z[zaddr(pp)] = cmul(xp, gp);
                if (pp != 0 && pm != pp) Z[zaddr(pm)] = cmul(xm, gm);
            }
        }
        __syncwarp();

        // ---------------- backward transform, discard, placement ------------------------------------------
        const FrameOut fo = wide_frame_out<D>(f);
        if constexpr (D <= 1) {
            constexpr int A = M / 1024;                              // 4 or 2 interleaved 1024-point transforms
#pragma unroll 4
            for (int i = 0; i < 32; i++) {
                const int n = lane + 32 * i;
                if constexpr (A == 4) {
                    const int s0 = ysrc(n), s1 = ysrc(n + 1024), s2 = ysrc(n + 2048), s3 = ysrc(n + 3072);
                    const float2 y0 = Z[s0], y1 = Z[s1], y2 = Z[s2], y3 = Z[s3];
                    const float2 a0 = make_float2(y0.x + y2.x, y0.y + y2.y), a1 = make_float2(y0.x - y2.x, y0.y - y2.y);
                    const float2 a2 = make_float2(y1.x + y3.x, y1.y + y3.y), a3 = make_float2(y1.x - y3.x, y1.y - y3.y);
                    const float2 u1 = make_float2(a1.x - a3.y, a1.y + a3.x);      // a1 + i a3
                    const float2 u3 = make_float2(a1.x + a3.y, a1.y - a3.x);      // a1 - i a3
                    Z[s0] = make_float2(a0.x + a2.x, a0.y + a2.y);
                    Z[s1] = cmul_conj(u1, __ldg(&p.twq[0 * 1024 + n]));
                    Z[s2] = cmul_conj(make_float2(a0.x - a2.x, a0.y - a2.y), __ldg(&p.twq[1 * 1024 + n]));
                    Z[s3] = cmul_conj(u3, __ldg(&p.twq[2 * 1024 + n]));
                } else {
                    const int s0 = ysrc(n), s1 = ysrc(n + 1024);
                    const float2 y0 = Z[s0], y1 = Z[s1];
                    Z[s0] = make_float2(y0.x + y1.x, y0.y + y1.y);
                    Z[s1] = cmul_conj(make_float2(y0.x - y1.x, y0.y - y1.y), __ldg(&p.twq[1 * 1024 + n]));   // W_2048^n
                }
            }
            __syncwarp();
#pragma unroll 1
            for (int q = 0; q < A; q++) {
#pragma unroll
                for (int j = 0; j < 32; j++) v[brev(j, 5)] = Z[ysrc(lane + 32 * j + 1024 * q)];
                fft1024<true>(v, T, p.tw1024, lane);
#pragma unroll
                for (int k1 = 0; k1 < 32; k1++) store_iq(p.out, fo, A * (lane + 32 * k1) + q, v[k1], p.imag_sign);
            }
        } else {
            constexpr int R = M / 32;                                // band bins per lane
            constexpr int G = 32 / R;
            constexpr int RB = ilog2(R);
            float2 y[R];
#pragma unroll
            for (int j = 0; j < R; j++) y[brev(j, RB)] = Z[ysrc(lane + 32 * j)];
            dit<R, true>(y);                                         // pass 1 over n2 (n = lane + 32 n2)
#pragma unroll
            for (int k2 = 1; k2 < R; k2++) y[k2] = cmul(y[k2], __ldg(&p.tw_r[k2 * 32 + lane]));   // exp(+2 pi i lane k2 / M)
#pragma unroll
            for (int k2 = 0; k2 < R; k2++) T[lane * (R + 1) + k2] = y[k2];
            __syncwarp();
            const int k2p = lane / G, h = lane % G;
            float2 u[R];
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) {                        // u_j = sum_r x[j + R r] exp(+2 pi i h (j + R r) / 32)
                const float2 x = T[n1 * (R + 1) + k2p];
                const float2 w = __ldg(&p.tw_r[(R + n1) * 32 + lane]);
                float2& acc = u[brev(n1 % R, RB)];
                if (n1 < R) acc = cmul(x, w);
                else acc = make_float2(fmaf(x.x, w.x, fmaf(-x.y, w.y, acc.x)), fmaf(x.x, w.y, fmaf(x.y, w.x, acc.y)));
            }
            __syncwarp();
            dit<R, true>(u);                                         // u[q] = y_out[k2p + R h + 32 q]
#pragma unroll
            for (int q = 0; q < R; q++) store_iq(p.out, fo, k2p + R * h + 32 * q, u[q], p.imag_sign);
        }
        __syncwarp();
    }
}

template <int D>
static cudaError_t setup_w() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_wide_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kWWarps * kWWarpFloat2 * sizeof(float2)));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_wide_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

cudaError_t setup_wide_kernels() {
    cudaError_t e;
    if ((e = setup_wide4_kernels()) != cudaSuccess) return e;
    if ((e = setup_w<0>()) != cudaSuccess) return e;
    if ((e = setup_w<1>()) != cudaSuccess) return e;
    if ((e = setup_w<2>()) != cudaSuccess) return e;
    if ((e = setup_w<3>()) != cudaSuccess) return e;
    if ((e = setup_w<4>()) != cudaSuccess) return e;
    if ((e = setup_w<5>()) != cudaSuccess) return e;
    if ((e = setup_w<6>()) != cudaSuccess) return e;
    return cudaSuccess;
}

template <int D>
static cudaError_t launch_w(const WideParams& p, int sm_count, cudaStream_t stream) {
    const size_t smem = kWWarps * kWWarpFloat2 * sizeof(float2);
    long long want = (p.total_frames + kWWarps - 1) / kWWarps;
    long long grid = sm_count;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    r2iq_wide_kernel<D><<<(unsigned)grid, kWThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_wide(int d, const WideParams& p, int sm_count, cudaStream_t stream) {
    if (kernel_generation() != 1) return launch_wide4(d, p, sm_count, stream);
    switch (d) {
        case 0: return launch_w<0>(p, sm_count, stream);
        case 1: return launch_w<1>(p, sm_count, stream);
        case 2: return launch_w<2>(p, sm_count, stream);
        case 3: return launch_w<3>(p, sm_count, stream);
        case 4: return launch_w<4>(p, sm_count, stream);
        case 5: return launch_w<5>(p, sm_count, stream);
        case 6: return launch_w<6>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
