// r2iq_wide_pair.cu -- FFTN = 8192 (the value the reference tree ships with, Core/config.h:49) on the packed
// two-frame lanes of the FFTN 2048 kernels.  halfFft = 4096, hop 6144 samples, 11 frames per 65536-sample block,
// inverse sizes mfft = 4096 >> d; worker-loop semantics of Core/fft_mt_r2iq_impl.hpp:15-151.
//
// The decomposition is that of the first wide kernel (r2iq_wide.cu, comparison builds): 4096 = 4 x 1024, a radix-4
// stage across four shared-memory planes and one 1024-point register FFT per plane and warp, so a TEAM of four warps
// owns a transform.  What changed:
//   * a team transforms a PAIR of consecutive frames at once, every value an (A, B) pair in one 64-bit register,
//     every butterfly one FFMA2 / FADD2 / FMUL2 (dft_regs2.cuh): half the issue slots per sample;
//   * ONE persistent CTA per SM of T teams (T = 3 at 168 registers; T = 2 where the d = 0 backward transform needs
//     more), teams synchronise among their own 128 threads with named barriers and take frame pairs from the
//     device-wide counter -- a team waiting at a barrier leaves the SM to the other teams, and no CTA is relaunched;
//   * the forward twiddles (W_1024 and W_4096^n, W_4096^2n; W^3n is their product) live in shared memory, loaded once
//     per CTA, instead of being fetched through a 28 KB L1 by every frame.
// Phases per pair (B = team barrier):
//   A  128 threads: de-randomise + int16 -> float (integer pipe, dual_frame.cuh) and radix-4 decimation in frequency
//      over the four quarters of both frames, twiddle W_4096^(q n), result to plane q                              B
//   B  warp q: 1024-point forward transform of plane q in place (two radix-32 register passes)                      B
//   C  128 threads: real-FFT split + filter on mirror pairs (k, 4096-k), in place, only the bins of the band         B
//   D  warp r: backward transform of plane r (the tune bin is a multiple of 4, Core/fft_mt_r2iq.cpp:101-109, so the
//      band's bins 4m'+r are plane r rotated by tunebin/4): 1024 points for d = 0, the R = mfft/128 register scheme
//      for d = 1..5, then W_mfft^(-r n')                                                                              B
//   E  128 threads: last radix-4 stage across the planes, overlap-save discard, placement of both frames            B
//   d = 6 (mfft = 64 is too small to split): warp 0 of the team transforms the band alone.
#include "dual_frame.cuh"
#include "r2iq_common.cuh"
#include "r2iq_kernels.cuh"

#include <cstdio>

namespace r2iq {

using namespace dualf;

namespace {

constexpr int kWHalf = 4096;                        // halfFft
constexpr int kWHopWords = 3 * kWHalf / 2 / 2;      // 6144 samples = 3072 words
constexpr int kWFrames = kTransferSamples / (3 * kWHalf / 2) + 1;   // fftPerBuf = 11
constexpr int kWPlane = 1060;                       // 64-bit slots per plane: >= 32*33 (transposition tile), == 4 mod 16 (skew)
constexpr int kTeamThreads = 128;
constexpr int kTeamBytes = 2 * 4 * kWPlane * 8;     // four planes of (re_A, re_B) then four of (im_A, im_B)
constexpr int kWTabFloat2 = 3 * 1024;               // tw1024 | W_4096^n | W_4096^(2n)

template <int T>
__host__ __device__ constexpr int wide_smem_bytes() {
    return kWTabFloat2 * 8 + T * kTeamBytes;
}

struct Planes {
    float2* re;   // [4][kWPlane] (re_A, re_B)
    float2* im;   // [4][kWPlane] (im_A, im_B)
};
__device__ __forceinline__ c2 pld(const Planes& P, int idx) {
    const float2 r = P.re[idx], i = P.im[idx];
    c2 v;
    v.re = pk(r.x, r.y);
    v.im = pk(i.x, i.y);
    return v;
}
__device__ __forceinline__ void pst(const Planes& P, int idx, c2 v) {
    float a, b, c, d;
    upk(v.re, a, b);
    upk(v.im, c, d);
    P.re[idx] = make_float2(a, b);
    P.im[idx] = make_float2(c, d);
}
__device__ __forceinline__ int waddr(int k) { return (k & 3) * kWPlane + (k >> 2); }
__device__ __forceinline__ c2 cadd(c2 a, c2 b) { return c2{add2(a.re, b.re), add2(a.im, b.im)}; }
__device__ __forceinline__ c2 csub(c2 a, c2 b) { return c2{sub2(a.re, b.re), sub2(a.im, b.im)}; }
__device__ __forceinline__ c2 cmul_conj2s(c2 v, float2 w) { return cmul2s(v, make_float2(w.x, -w.y)); }
__device__ __forceinline__ void team_barrier(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(kTeamThreads) : "memory");
}

// 1024-point transform of x[lane + 32 j] (v[brev5(j)] on entry) -> X[lane + 32 k1] (v[k1] on return): two radix-32
// register passes around one transposition through this warp's own plane (pitch 33); twiddles from shared memory
template <bool INV>
__device__ __forceinline__ void fft1024p(c2 (&v)[32], const Planes& P, int base, const float2* tw1024, int lane) {
    dit2<32, INV>(v);
#pragma unroll
    for (int k2 = 1; k2 < 32; k2++) {
        const float2 w = tw1024[32 * k2 + lane];
        v[k2] = INV ? cmul_conj2s(v[k2], w) : cmul2s(v[k2], w);
    }
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) pst(P, base + lane * 33 + k2, v[k2]);
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = pld(P, base + n1 * 33 + lane);
    __syncwarp();
    dit2<32, INV>(v);
}

template <int D>
__device__ __forceinline__ FrameOut2 wide_frame_out(long long f, long long total) {
    constexpr int M = kWHalf >> D;
    const long long b = f / kWFrames;
    const int k = (int)(f - b * kWFrames);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = b * (long long)((kTransferSamples / 2) >> D) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

// both frames' sample idx of the backward transform; idx < 3 mfft / 4 by construction of the callers
__device__ __forceinline__ void store_pair(float2* out, const FrameOut2& oa, const FrameOut2& ob, int idx, c2 v, uint32_t sgn) {
    float ar, br, ai, bi;
    upk(v.re, ar, br);
    upk(v.im, ai, bi);
    if (idx >= oa.lo) out[oa.base + idx] = make_float2(ar, flip(ai, sgn));
    if (ob.live && idx >= ob.lo) out[ob.base + idx] = make_float2(br, flip(bi, sgn));
}

}  // namespace

template <int D, int T>
__global__ void __launch_bounds__(T* kTeamThreads, 1) r2iq_wide_pair_kernel(const WideParams p) {
    constexpr int M = kWHalf >> D;
    extern __shared__ __align__(128) unsigned char wsmem[];
    __shared__ long long s_pair[T];
    const int team = threadIdx.x / kTeamThreads;
    const int tt = threadIdx.x % kTeamThreads;
    const int lane = tt & 31;
    const int warp = tt >> 5;                          // plane this warp transforms
    float2* tab = reinterpret_cast<float2*>(wsmem);
    const float2* tw1024 = tab;
    const float2* twq1 = tab + 1024;                   // W_4096^n
    const float2* twq2 = tab + 2048;                   // W_4096^(2n)
    Planes P;
    P.re = reinterpret_cast<float2*>(wsmem + kWTabFloat2 * 8 + (size_t)team * kTeamBytes);
    P.im = P.re + 4 * kWPlane;
    const int wbase = warp * kWPlane;
    const int tb = p.tunebin;
    const int tb4 = tb >> 2;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;

    for (int i = threadIdx.x; i < 1024; i += T * kTeamThreads) {
        tab[i] = __ldg(p.tw1024 + i);
        tab[1024 + i] = __ldg(p.twq + i);
        tab[2048 + i] = __ldg(p.twq + 1024 + i);
    }
    const long long n_pairs = (p.total_frames + 1) / 2;
    if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);
    __syncthreads();

    while (true) {
        const long long q = s_pair[team];
        if (q >= n_pairs) break;
        // ---------------- A: radix-4 DIF over the four quarters of both frames -------------------------------
        {
            const long long fa = 2 * q;
            const long long b = fa / kWFrames;
            const int kf = (int)(fa - b * kWFrames);
            const uint32_t* srcA = p.in_words + b * (kTransferSamples / 2) + kf * kWHopWords + tt;
            // frame B starts one hop later (2048 words later across a block edge); the phantom B of an odd total reads A
            const int offB = fa + 1 >= p.total_frames ? 0 : (kf == kWFrames - 1 ? kWHalf / 2 : kWHopWords);
            const uint32_t* srcB = srcA + offB;
#pragma unroll 4
            for (int j = 0; j < 8; j++) {
                const int n = tt + 128 * j;
                c2 z0, z1, z2, z3;
                unpack_words(__ldg(srcA + 128 * j), __ldg(srcB + 128 * j), p.rand_mul, z0);
                unpack_words(__ldg(srcA + 128 * j + 1024), __ldg(srcB + 128 * j + 1024), p.rand_mul, z1);
                unpack_words(__ldg(srcA + 128 * j + 2048), __ldg(srcB + 128 * j + 2048), p.rand_mul, z2);
                unpack_words(__ldg(srcA + 128 * j + 3072), __ldg(srcB + 128 * j + 3072), p.rand_mul, z3);
                const c2 a0 = cadd(z0, z2), a1 = csub(z0, z2), a2 = cadd(z1, z3), a3 = csub(z1, z3);
                const c2 u0 = cadd(a0, a2), u2 = csub(a0, a2);
                const c2 u1 = c2{add2(a1.re, a3.im), sub2(a1.im, a3.re)};      // a1 - i a3
                const c2 u3 = c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)};      // a1 + i a3
                const float2 w1 = twq1[n], w2 = twq2[n];
                pst(P, 0 * kWPlane + n, u0);
                pst(P, 1 * kWPlane + n, cmul2s(u1, w1));
                pst(P, 2 * kWPlane + n, cmul2s(u2, w2));
                pst(P, 3 * kWPlane + n, cmul2s(u3, cmul(w1, w2)));               // W_4096^(3 n)
            }
        }
        team_barrier(team);
        if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);   // every thread has read q; next read is after barrier E

        // ---------------- B: warp q transforms plane q in place -------------------------------------------------
        {
            c2 v[32];
#pragma unroll
            for (int j = 0; j < 32; j++) v[brev5(j)] = pld(P, wbase + lane + 32 * j);
            __syncwarp();
            fft1024p<false>(v, P, wbase, tw1024, lane);
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) pst(P, wbase + lane + 32 * k1, v[k1]);   // bin 4 (lane + 32 k1) + warp
        }
        team_barrier(team);

        // ---------------- C: real-FFT split + filter, in place on mirror pairs ------------------------------
#pragma unroll 2
        for (int i = 0; i <= 16; i++) {
            const int pp = i < 16 ? tt + 128 * i : 2048;
            if (i < 16 || tt == 0) {
                const int pm = (kWHalf - pp) & (kWHalf - 1);
                if constexpr (D > 0) {      // only bins of the selected band are read by the backward transform
                    const bool inp = ((pp - tb + M / 2) & (kWHalf - 1)) < M, inm = ((pm - tb + M / 2) & (kWHalf - 1)) < M;
                    if (!(inp || inm)) continue;
                }
                const c2 a = pld(P, waddr(pp));
                const c2 bq = pld(P, waddr(pm));
                const float2 c = __ldg(&p.ctab[pp]);                 // -i * W_8192^p
                const float2 gp = __ldg(&p.g[pp]);
                const float2 gm = __ldg(&p.g[pm]);
                const c2 u = c2{add2(a.re, bq.re), sub2(a.im, bq.im)};      // a + conj b
                const c2 w = c2{sub2(a.re, bq.re), add2(a.im, bq.im)};      // a - conj b
                const c2 t = cmul2s(w, c);
                const c2 xp = cadd(u, t);                                   // 2 X[p]
                const c2 xm = c2{sub2(u.re, t.re), sub2(t.im, u.im)};       // 2 X[4096-p] = conj(u - t)
                pst(P, waddr(pp), cmul2s(xp, gp));
                if (pp != 0 && pm != pp) pst(P, waddr(pm), cmul2s(xm, gm));
            }
        }
        team_barrier(team);

        // ---------------- D + E: backward transform, discard, placement ------------------------------------------
        const FrameOut2 oa = wide_frame_out<D>(2 * q, p.total_frames);
        const FrameOut2 ob = wide_frame_out<D>(2 * q + 1, p.total_frames);
        if constexpr (D <= 5) {
            constexpr int MS = M / 4;                                 // points per plane: 1024 >> D
            if constexpr (D == 0) {
                c2 v[32];
#pragma unroll
                for (int j = 0; j < 32; j++) v[brev5(j)] = pld(P, wbase + ((tb4 + lane + 32 * j) & 1023));
                __syncwarp();
                fft1024p<true>(v, P, wbase, tw1024, lane);
                if (warp) {
#pragma unroll
                    for (int k1 = 0; k1 < 32; k1++) {                 // W_4096^(-warp n')
                        const int n = lane + 32 * k1;
                        const float2 w = warp == 1 ? twq1[n] : warp == 2 ? twq2[n] : cmul(twq1[n], twq2[n]);
                        v[k1] = cmul_conj2s(v[k1], w);
                    }
                }
#pragma unroll
                for (int k1 = 0; k1 < 32; k1++) pst(P, wbase + lane + 32 * k1, v[k1]);
            } else {
                constexpr int R = MS / 32;                            // band bins of this plane per lane
                constexpr int G = 32 / R;
                constexpr int RB = ilog2(R);
                c2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int m = lane + 32 * j;
                    y[brev_n(j, RB)] = pld(P, wbase + ((tb4 + (m < MS / 2 ? m : m - MS)) & 1023));
                }
                __syncwarp();
                dit2<R, true>(y);                                     // pass 1 over n2 (m' = lane + 32 n2)
#pragma unroll
                for (int k2 = 1; k2 < R; k2++) y[k2] = cmul2s(y[k2], __ldg(&p.tw_sub[k2 * 32 + lane]));   // exp(+2 pi i lane k2 / MS)
#pragma unroll
                for (int k2 = 0; k2 < R; k2++) pst(P, wbase + lane * (R + 1) + k2, y[k2]);
                __syncwarp();
                const int k2p = lane / G, h = lane % G;
                c2 u[R];
#pragma unroll
                for (int n1 = 0; n1 < 32; n1++) {                     // u_j = sum_r x[j + R r] exp(+2 pi i h (j + R r) / 32)
                    const c2 x = pld(P, wbase + n1 * (R + 1) + k2p);
                    const float2 w = __ldg(&p.tw_sub[(R + n1) * 32 + lane]);
                    c2& acc = u[brev_n(n1 % R, RB)];
                    if (n1 < R) {
                        acc = cmul2s(x, w);
                    } else {
                        acc.re = fma2(bc(w.x), x.re, fma2(bc(-w.y), x.im, acc.re));
                        acc.im = fma2(bc(w.x), x.im, fma2(bc(w.y), x.re, acc.im));
                    }
                }
                __syncwarp();
                dit2<R, true>(u);                                     // u[q] = S_warp[k2p + R h + 32 q]
#pragma unroll
                for (int qq = 0; qq < R; qq++) {
                    const int n = k2p + R * h + 32 * qq;
                    c2 s = u[qq];
                    if (warp) {                                       // W_mfft^(-warp n') = W_4096^(-warp (n' << D))
                        const int e = n << D;
                        const float2 w = warp == 1 ? twq1[e] : warp == 2 ? twq2[e] : cmul(twq1[e], twq2[e]);
                        s = cmul_conj2s(s, w);
                    }
                    pst(P, wbase + n, s);
                }
            }
            team_barrier(team);
            // last radix-4 stage across the planes: y[n' + MS a] = sum_r i^(r a) S_r[n'], a = 0..2 kept
#pragma unroll
            for (int i = 0; i < (MS + kTeamThreads - 1) / kTeamThreads; i++) {
                const int n = tt + kTeamThreads * i;
                if (MS >= kTeamThreads || n < MS) {
                    const c2 s0 = pld(P, n), s1 = pld(P, kWPlane + n), s2 = pld(P, 2 * kWPlane + n), s3 = pld(P, 3 * kWPlane + n);
                    const c2 a0 = cadd(s0, s2), a1 = csub(s0, s2), a2 = cadd(s1, s3), a3 = csub(s1, s3);
                    store_pair(p.out, oa, ob, n, cadd(a0, a2), sgn);
                    store_pair(p.out, oa, ob, n + MS, c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)}, sgn);      // a1 + i a3
                    store_pair(p.out, oa, ob, n + 2 * MS, csub(a0, a2), sgn);
                }
            }
        } else {
            // d = 6: mfft = 64, warp 0 alone (R = 2 register scheme over the whole band)
            if (warp == 0) {
                constexpr int R = M / 32, G = 32 / R, RB = ilog2(R);
                auto ysrc = [&](int m) { return waddr((tb + (m < M / 2 ? m : m - M)) & (kWHalf - 1)); };
                c2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) y[brev_n(j, RB)] = pld(P, ysrc(lane + 32 * j));
                __syncwarp();
                dit2<R, true>(y);
#pragma unroll
                for (int k2 = 1; k2 < R; k2++) y[k2] = cmul2s(y[k2], __ldg(&p.tw_sub[k2 * 32 + lane]));
#pragma unroll
                for (int k2 = 0; k2 < R; k2++) pst(P, lane * (R + 1) + k2, y[k2]);
                __syncwarp();
                const int k2p = lane / G, h = lane % G;
                c2 u[R];
#pragma unroll
                for (int n1 = 0; n1 < 32; n1++) {
                    const c2 x = pld(P, n1 * (R + 1) + k2p);
                    const float2 w = __ldg(&p.tw_sub[(R + n1) * 32 + lane]);
                    c2& acc = u[brev_n(n1 % R, RB)];
                    if (n1 < R) {
                        acc = cmul2s(x, w);
                    } else {
                        acc.re = fma2(bc(w.x), x.re, fma2(bc(-w.y), x.im, acc.re));
                        acc.im = fma2(bc(w.x), x.im, fma2(bc(w.y), x.re, acc.im));
                    }
                }
                dit2<R, true>(u);
#pragma unroll
                for (int qq = 0; qq < R; qq++) {
                    const int idx = k2p + R * h + 32 * qq;
                    if (idx < 3 * M / 4) store_pair(p.out, oa, ob, idx, u[qq], sgn);
                }
            }
        }
        team_barrier(team);                                           // the planes are rewritten by the next pair
    }
}

template <int D, int T>
static cudaError_t setup_wp() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_wide_pair_kernel<D, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide_smem_bytes<T>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_wide_pair_kernel<D, T>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <int D, int T>
static cudaError_t launch_wp(const WideParams& p, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long grid = sm_count;                        // persistent: one CTA per SM
    const long long want = (n_pairs + T - 1) / T;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_wide_pair_kernel<D, T><<<(unsigned)grid, T * kTeamThreads, wide_smem_bytes<T>(), stream>>>(p);
    return cudaGetLastError();
}

// teams per CTA for each decimation
#define R2IQ_WIDE_FOR_ALL(X) X(0, 3) X(1, 3) X(2, 3) X(3, 3) X(4, 3) X(5, 3) X(6, 3)

cudaError_t setup_wide_kernels() {
    cudaError_t e;
#define X(D, T) \
    if ((e = setup_wp<D, T>()) != cudaSuccess) return e;
    R2IQ_WIDE_FOR_ALL(X)
#undef X
#ifdef R2IQ_AB_VARIANTS
    if ((e = setup_wide_gen1_kernels()) != cudaSuccess) return e;
#endif
    return cudaSuccess;
}

cudaError_t launch_wide(int d, const WideParams& p, int sm_count, cudaStream_t stream) {
#ifdef R2IQ_AB_VARIANTS
    if (kernel_generation() != 0) return launch_wide_gen1(d, p, sm_count, stream);   // the four-warps-per-frame scalar kernels
#endif
#define X(D, T) \
    if (d == D) return launch_wp<D, T>(p, sm_count, stream);
    R2IQ_WIDE_FOR_ALL(X)
#undef X
    return cudaErrorInvalidValue;
}

const char* wide_kernel_name(int d) {
    static char names[kNDec][48];
    if (d < 0 || d >= kNDec) return "";
    if (!names[d][0]) snprintf(names[d], sizeof(names[d]), "r2iq_wide_pair_kernel<%d, %d>", d, 3);
    return names[d];
}

}  // namespace r2iq
