// r2iq_wide_pair.cu -- FFTN = 8192 (the value the reference tree ships with, Core/config.h:49) on the packed
// two-frame lanes of the FFTN 2048 kernels.  halfFft = 4096, hop 6144 samples, 11 frames per 65536-sample block,
// inverse sizes mfft = 4096 >> d; worker-loop semantics of Core/fft_mt_r2iq_impl.hpp:15-151.
//
// The decomposition is that of the first wide kernel (r2iq_wide.cu, comparison builds): 4096 = 4 x 1024, a radix-4
// stage across four shared-memory planes and one 1024-point register FFT per plane and warp, so a TEAM of four warps
// owns a transform.  What this generation does differently, each from a measurement (ncu source page of its
// predecessors: the shared-memory data pipe is the busiest unit at 72 %, the split phase cost a quarter of the time):
//   * a team transforms a PAIR of consecutive frames at once, every value an (A, B) pair in one 64-bit register,
//     every butterfly one FFMA2 / FADD2 / FMUL2 (dft_regs2.cuh): half the issue slots per sample;
//   * ONE persistent CTA per SM of three teams; teams synchronise among their own 128 threads with named barriers and
//     take frame pairs from the device-wide counter -- a team waiting at a barrier leaves the SM to the other teams;
//   * no separate split phase: the backward transform reads bin k from its own plane and bin 4096-k from the mirror
//     plane and applies Y = P[k] Z[k] + Q[k] conj(Z[4096-k]) (split, filter and band edge in two table values, as
//     the 2048 family) on the way into its registers; the forward transform stores only the rows of its plane that
//     some backward transform will read;
//   * d = 0: the tune rotation costs nothing -- lane l of the backward transform plays residue (l - tunebin/4) mod 32,
//     which only changes the inter-pass twiddle table and the tile row the lane writes (as r2iq_pair.cu, d = 0);
//   * the first half of the NEXT pair's input words is fetched into registers before the last phase of the current
//     pair and the second half while the first is being transformed, so phase A no longer waits on HBM;
//   * all twiddle tables live in shared memory (W_4096^2n and ^3n are products of W_4096^n: no table, no load).
// Phases per pair (B = team barrier):
//   A   128 threads: de-randomise + int16 -> float (integer pipe, dual_frame.cuh) and radix-4 decimation in frequency
//       over the four quarters of both frames, twiddle W_4096^(q n), result to plane q                             B
//   B   warp q: 1024-point forward transform of plane q (two radix-32 register passes), needed rows back to plane q  B
//   D1  warp r: band bins of plane r (the tune bin is a multiple of 4, Core/fft_mt_r2iq.cpp:101-109, so the band's
//       bins 4m'+r are plane r rotated by tunebin/4) and their mirrors -> P/Q -> first register pass                 B
//   D2  warp r: transposition through its own plane, second pass, W_mfft^(-r n'), S_r to plane r                    B
//   E   128 threads: last radix-4 stage across the planes, overlap-save discard, placement of both frames          B
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
constexpr int kTeams = 3;
constexpr int kTeamBytes = 2 * 4 * kWPlane * 8;     // four planes of (re_A, re_B) then four of (im_A, im_B)

// tables in shared memory: W_1024 inter-pass twiddles, transposed [lane][34] (one LDS.128 per two entries) | W_4096^n |
// d = 0: backward inter-pass twiddles of this tune bin, d >= 1: the pass-2 weights of the R-scheme (build_tw_aux, second part)
constexpr int kWTabPitch = 34;
constexpr int kWTab1 = 32 * kWTabPitch;             // start of W_4096^n
constexpr int kWTabX = kWTab1 + 1024;               // start of the per-decimation table
template <int D>
__host__ __device__ constexpr int wide_tab_float2() {
    return kWTabX + 32 * kWTabPitch;
}
// entry n of a lane's row in a transposed table: one LDS.128 serves n and n + 1
__device__ __forceinline__ float2 wtab_at(const float2* row, int n) {
    const float4 w = *reinterpret_cast<const float4*>(row + (n & ~1));
    return (n & 1) ? make_float2(w.z, w.w) : make_float2(w.x, w.y);
}
template <int D>
__host__ __device__ constexpr int wide_smem_bytes() {
    return wide_tab_float2<D>() * 8 + kTeams * kTeamBytes;
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
__device__ __forceinline__ float2 csqr(float2 w) { return make_float2(fmaf(w.x, w.x, -w.y * w.y), 2.0f * w.x * w.y); }
__device__ __forceinline__ void team_barrier(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(kTeamThreads) : "memory");
}
// split + filter + band edge: Y = P a + Q conj(b), a = Z[k], b = Z[4096 - k]; pq = (P.re, P.im, Q.re, Q.im)
__device__ __forceinline__ c2 split_pq(c2 a, c2 b, float4 pq) {
    c2 y;
    y.re = fma2(bc(pq.w), b.im, fma2(bc(pq.z), b.re, fma2(bc(-pq.y), a.im, mul2(bc(pq.x), a.re))));
    y.im = fma2(bc(pq.w), b.re, fma2(bc(-pq.z), b.im, fma2(bc(pq.y), a.re, mul2(bc(pq.x), a.im))));
    return y;
}
// slot of bin 4096 - (4 mm + r) inside plane (4 - r) & 3
__device__ __forceinline__ int mirror_slot(int r, int mm) { return r == 0 ? ((1024 - mm) & 1023) : 1023 - mm; }

template <int D>
__device__ __forceinline__ FrameOut2 wide_frame_out(long long f, long long total) {
    constexpr int M = kWHalf >> D;
    const uint32_t b = (uint32_t)f / (uint32_t)kWFrames;      // frame indices fit 32 bits (dual_frame.cuh: frame_src)
    const int k = (int)((uint32_t)f - b * (uint32_t)kWFrames);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = (long long)((unsigned long long)b * (unsigned)((kTransferSamples / 2) >> D)) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
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
// the same with the direct-phase fine-tune mixer: ph0 = the two frames' phasors at the thread's first sample, rot = the
// rotation from there to idx (a table entry: no chain of products); sideband conjugation first (RadioHandler.cpp:33-37)
__device__ __forceinline__ void store_pair_mixed(float2* out, const FrameOut2& oa, const FrameOut2& ob, int idx, c2 v, c2 ph0,
                                                 float2 rot, float imag_sign) {
    v = direct_mix(v, ph0, rot, imag_sign);
    float ar, br, ai, bi;
    upk(v.re, ar, br);
    upk(v.im, ai, bi);
    if (idx >= oa.lo) out[oa.base + idx] = make_float2(ar, ai);
    if (ob.live && idx >= ob.lo) out[ob.base + idx] = make_float2(br, bi);
}

// Phase E stores: sample n + MS a with n < MS and lo in {0, MS} -- the a = 0 third is kept iff lo == 0 (frames other
// than the first of a block), the other two always: warp-uniform predicates, computed once per pair
struct KeepPair {
    bool a, b;
};
__device__ __forceinline__ void store_pair_rows(float2* outA, float2* outB, KeepPair keep, int idx, c2 v, uint32_t sgn) {
    float ar, br, ai, bi;
    upk(v.re, ar, br);
    upk(v.im, ai, bi);
    if (keep.a) outA[idx] = make_float2(ar, flip(ai, sgn));
    if (keep.b) outB[idx] = make_float2(br, flip(bi, sgn));
}
__device__ __forceinline__ void store_pair_rows_mixed(float2* outA, float2* outB, KeepPair keep, int idx, c2 v, c2 ph0, float2 rot,
                                                      float imag_sign) {
    v = direct_mix(v, ph0, rot, imag_sign);
    float ar, br, ai, bi;
    upk(v.re, ar, br);
    upk(v.im, ai, bi);
    if (keep.a) outA[idx] = make_float2(ar, ai);
    if (keep.b) outB[idx] = make_float2(br, bi);
}

// Input words of one half of a frame pair: thread tt's points n = tt + 128 (j0 + jj), jj < 4, of the four quarters.
struct HalfWords {
    uint32_t a[4][4], b[4][4];   // [jj][quarter], frame A / frame B
};
__device__ __forceinline__ void load_half(const WideParams& p, long long pair, int tt, int j0, HalfWords& w) {
    const long long fa = 2 * pair;
    const uint32_t blk = (uint32_t)fa / (uint32_t)kWFrames;
    const int kf = (int)((uint32_t)fa - blk * (uint32_t)kWFrames);
    const uint32_t* srcA = p.in_words + (size_t)blk * (kTransferSamples / 2) + kf * kWHopWords + tt + 128 * j0;
    // frame B starts one hop later (2048 words later across a block edge); the phantom B of an odd total reads A
    const int offB = fa + 1 >= p.total_frames ? 0 : (kf == kWFrames - 1 ? kWHalf / 2 : kWHopWords);
    const uint32_t* srcB = srcA + offB;
#pragma unroll
    for (int jj = 0; jj < 4; jj++)
#pragma unroll
        for (int qq = 0; qq < 4; qq++) {
            w.a[jj][qq] = __ldg(srcA + 128 * jj + 1024 * qq);
            w.b[jj][qq] = __ldg(srcB + 128 * jj + 1024 * qq);
        }
}
// Both frames of a pair (2 x 128 lines of 128 bytes) into L2, one line of each frame per thread: issued a phase and a half
// before load_half reads them, so that the loads wait for an L2 hit and not for HBM (13 % of the d = 5 kernel's stall
// samples sat on the first use of the loaded words: long scoreboard, tools/ncu_hot.py).  No registers are held.
__device__ __forceinline__ void prefetch_pair_l2(const WideParams& p, long long pair, int tt) {
    const long long fa = 2 * pair;
    const uint32_t blk = (uint32_t)fa / (uint32_t)kWFrames;
    const int kf = (int)((uint32_t)fa - blk * (uint32_t)kWFrames);
    const uint32_t* srcA = p.in_words + (size_t)blk * (kTransferSamples / 2) + kf * kWHopWords + 32 * tt;
    const int offB = fa + 1 >= p.total_frames ? 0 : (kf == kWFrames - 1 ? kWHalf / 2 : kWHopWords);
    asm volatile("prefetch.global.L2 [%0];" ::"l"(srcA));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(srcA + offB));
}
// phase A on one half: radix-4 decimation in frequency over the quarters, twiddles, planes
__device__ __forceinline__ void quarter_stage(const HalfWords& w, const Planes& P, const float2* twq1, uint32_t rand_mul, int tt, int j0) {
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
        const int n = tt + 128 * (j0 + jj);
        c2 z0, z1, z2, z3;
        unpack_words(w.a[jj][0], w.b[jj][0], rand_mul, z0);
        unpack_words(w.a[jj][1], w.b[jj][1], rand_mul, z1);
        unpack_words(w.a[jj][2], w.b[jj][2], rand_mul, z2);
        unpack_words(w.a[jj][3], w.b[jj][3], rand_mul, z3);
        const c2 a0 = cadd(z0, z2), a1 = csub(z0, z2), a2 = cadd(z1, z3), a3 = csub(z1, z3);
        const c2 u0 = cadd(a0, a2), u2 = csub(a0, a2);
        const c2 u1 = c2{add2(a1.re, a3.im), sub2(a1.im, a3.re)};      // a1 - i a3
        const c2 u3 = c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)};      // a1 + i a3
        const float2 w1 = twq1[n], w2 = csqr(w1);
        pst(P, 0 * kWPlane + n, u0);
        pst(P, 1 * kWPlane + n, cmul2s(u1, w1));
        pst(P, 2 * kWPlane + n, cmul2s(u2, w2));
        pst(P, 3 * kWPlane + n, cmul2s(u3, cmul(w1, w2)));               // W_4096^(3 n)
    }
}

// R-scheme backward transform of an S = 32 R point sequence on one warp, second half: y[k2] (pass-1 outputs times
// their twiddles, one column per lane) -> u[q] = s[k2p + R h + 32 q]; scratch = 32 x (R + 1) slots at `base`
template <int R>
__device__ __forceinline__ void rscheme_pass2(c2 (&y)[R], c2 (&u)[R], const Planes& P, int base, const float2* wts, int lane) {
    constexpr int G = 32 / R;
    constexpr int RB = ilog2(R);
#pragma unroll
    for (int k2 = 0; k2 < R; k2++) pst(P, base + lane * (R + 1) + k2, y[k2]);
    __syncwarp();
    const int k2p = lane / G;
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) {                     // u_j = sum_r x[j + R r] exp(+2 pi i h (j + R r) / 32)
        const c2 x = pld(P, base + n1 * (R + 1) + k2p);
        const float2 w = wtab_at(wts + lane * kWTabPitch, n1);
        c2& acc = u[brev_n(n1 % R, RB)];
        if (n1 < R) {
            acc = cmul2s(x, w);
        } else {
            acc.re = fma2(bc(w.x), x.re, fma2(bc(-w.y), x.im, acc.re));
            acc.im = fma2(bc(w.x), x.im, fma2(bc(w.y), x.re, acc.im));
        }
    }
    __syncwarp();
    dit2<R, true>(u);
}

}  // namespace

#ifndef R2IQ_WIDE_HELPERS_ONLY   // r2iq_wide_channels.cu includes this file for the helpers above only
template <int D, int kMix>
__global__ void __launch_bounds__(kTeams* kTeamThreads, 1) r2iq_wide_pair_kernel(const WideParams p, const WideNco nco) {
    constexpr int M = kWHalf >> D;
    extern __shared__ __align__(128) unsigned char wsmem[];
    __shared__ long long s_pair[kTeams];
    const int team = threadIdx.x / kTeamThreads;
    const int tt = threadIdx.x % kTeamThreads;
    const int lane = tt & 31;
    const int warp = tt >> 5;                          // plane this warp transforms
    float2* tab = reinterpret_cast<float2*>(wsmem);
    const float2* tw1024 = tab + lane * kWTabPitch;    // this lane's row: entry n at tw1024[n]
    const float2* twq1 = tab + kWTab1;                 // W_4096^n
    const float2* tabx = tab + kWTabX;                 // d = 0: tw_inv; d >= 1: pass-2 weights
    Planes P;
    P.re = reinterpret_cast<float2*>(wsmem + wide_tab_float2<D>() * 8 + (size_t)team * kTeamBytes);
    P.im = P.re + 4 * kWPlane;
    const int wbase = warp * kWPlane;
    const int mbase = ((4 - warp) & 3) * kWPlane;      // the plane that holds this plane's mirror bins
    const int tb = p.tunebin;
    const int tb4 = tb >> 2;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;
    const uint32_t row_mask = warp == 0 ? p.row_mask[0] : warp == 1 ? p.row_mask[1] : warp == 2 ? p.row_mask[2] : p.row_mask[3];
    const float4* pq = p.pq + warp * 1024;
    // d = 3: the lane's (P, Q) of its four band bins are the same for every frame pair -- kept in registers instead of one
    // L1 / L2 round trip per pair at the head of the short phase D: 650 -> 660 Gsamples/s.  (d = 4: no difference, d = 5:
    // 798 -> 775 -- the registers cost more in phase B than the loads did.)
    constexpr bool kPqRegs = D == 3;
    constexpr int kPqN = kPqRegs ? (kWHalf >> D) / 4 / 32 : 1;
    float4 pq_reg[kPqN];
    if constexpr (kPqRegs) {
#pragma unroll
        for (int j = 0; j < kPqN; j++) {
            const int m = lane + 32 * j, ms = (kWHalf >> D) / 4;
            pq_reg[j] = __ldg(pq + ((tb4 + (m < ms / 2 ? m : m - ms)) & 1023));
        }
    }

    for (int i = threadIdx.x; i < 1024; i += kTeams * kTeamThreads) {
        tab[(i & 31) * kWTabPitch + (i >> 5)] = __ldg(p.tw1024 + i);
        tab[kWTab1 + i] = __ldg(p.twq + i);
    }
    // d = 4, 5 (1024 >> d = 64, 32 points per plane): the per-plane backward transform runs across the lanes
    // (dual_frame.cuh: warp_idft32), its per-lane twiddles -- pitch 10: conflict-free 16-byte reads -- replace the weights
    constexpr bool kCrossLane = D == 4 || D == 5;
    constexpr int kXPitch = 10;
    if constexpr (kCrossLane) {
        constexpr int RX = (kWHalf >> D) / 4 / 32;
        for (int i = threadIdx.x; i < 32 * (3 + RX); i += kTeams * kTeamThreads)
            tab[kWTabX + (i & 31) * kXPitch + (i >> 5)] = warp_idft_twiddle(i & 31, i >> 5, (kWHalf >> D) / 4);
    } else {
        // d >= 1: only the [32][32] weights behind the R unused twiddle rows of tw_sub (their twiddles are folded in)
        constexpr int kSkip = D == 0 ? 0 : (D == 6 ? 64 : (kWHalf >> D) / 4) / 32 * 32;
        for (int i = threadIdx.x; i < 1024; i += kTeams * kTeamThreads)     // [n][lane] -> [lane][n]
            tab[kWTabX + (i & 31) * kWTabPitch + (i >> 5)] = __ldg((D == 0 ? p.tw_inv : p.tw_sub + kSkip) + i);
    }
    const long long n_pairs = (p.total_frames + 1) / 2;
    if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);
    __syncthreads();

    long long q = s_pair[team];
    HalfWords w0;
    load_half(p, q < n_pairs ? q : 0, tt, 0, w0);
    while (q < n_pairs) {
        // ---------------- A: radix-4 DIF over the four quarters of both frames -------------------------------
        {
            HalfWords w1;
            load_half(p, q, tt, 4, w1);
            quarter_stage(w0, P, twq1, p.rand_mul, tt, 0);
            quarter_stage(w1, P, twq1, p.rand_mul, tt, 4);
        }
        team_barrier(team);
        if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);   // every thread has read q; read again after barrier D2

        // ---------------- B: warp q transforms plane q; rows some backward transform reads go back ------------
        {
            c2 v[32];
#pragma unroll
            for (int j = 0; j < 32; j++) v[brev5(j)] = pld(P, wbase + lane + 32 * j);
            __syncwarp();
            dit2<32, false>(v);
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) pst(P, wbase + lane * 33 + k2, v[k2]);
            __syncwarp();
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = pld(P, wbase + n1 * 33 + lane);
            __syncwarp();
            // inter-pass twiddles W_1024^(n1 lane) folded into the first stage of the second pass (dft_regs2.cuh)
            dit2_tw<32, false>(v, [&](int i) { return wtab_at(tw1024, brev5(i)); });
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++)                                         // bin 4 (lane + 32 k1) + warp
                if (D == 0 || ((row_mask >> k1) & 1u)) pst(P, wbase + lane + 32 * k1, v[k1]);
        }
        team_barrier(team);
        // measured per decimation (B200, 8192 blocks): d = 3 665 -> 674, d = 5 799 -> 819, d = 6 725 -> 730 Gsamples/s; d = 0, 2, 4
        // lose 0.5 - 0.9 % (two more LSU instructions per thread in kernels bound by that pipe), d = 1 is even
        if constexpr (D == 3 || D >= 5) {
            const long long qpf = s_pair[team];       // the next pair of this team (written after barrier A)
            if (qpf < n_pairs) prefetch_pair_l2(p, qpf, tt);
        }

        const FrameOut2 oa = wide_frame_out<D>(2 * q, p.total_frames);
        const FrameOut2 ob = wide_frame_out<D>(2 * q + 1, p.total_frames);
        long long qn;
        if constexpr (D == 0) {
            // ---------------- D1: split + filter into registers, first backward pass ---------------------------
            c2 t[32];
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) {
                const int mm = lane + 32 * k1;
                const c2 a = pld(P, wbase + mm);      // (keeping phase B's v[k1] alive across the barrier instead: 291 vs 325 Gsamples/s)
                const c2 b = pld(P, mbase + mirror_slot(warp, mm));
                t[brev5(k1)] = split_pq(a, b, __ldg(pq + mm));
            }
            dit2<32, true>(t);
            team_barrier(team);                       // every mirror read of the team is done: the planes are free
            // ---------------- D2: transposition, second pass, W_4096^(-warp n), S_warp to the plane ------------
            const int n1_inv = (lane - tb4) & 31;
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) pst(P, wbase + n1_inv * 33 + k2, t[k2]);
            __syncwarp();
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) t[brev5(n1)] = pld(P, wbase + n1 * 33 + lane);
            __syncwarp();
            // inter-pass twiddles on the reading side (tw_inv row n = twiddle of the element read from row n, written by the
            // lane that plays residue n), folded into the first stage
            dit2_tw<32, true, false>(t, [&](int i) { return wtab_at(tabx + lane * kWTabPitch, brev5(i)); });
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) pst(P, wbase + lane + 32 * k1, t[k1]);   // W_4096^(-warp n) follows in phase E
        } else if constexpr (D <= 5) {
            constexpr int MS = M / 4;                                 // points per plane: 1024 >> D
            constexpr int R = MS / 32;                                // band bins of this plane per lane
            constexpr int G = 32 / R;
            constexpr int RB = ilog2(R);
            c2 y[R];
#pragma unroll
            for (int j = 0; j < R; j++) {
                const int m = lane + 32 * j;
                const int mm = (tb4 + (m < MS / 2 ? m : m - MS)) & 1023;
                const c2 a = pld(P, wbase + mm);
                const c2 b = pld(P, mbase + mirror_slot(warp, mm));
                y[brev_n(j, RB)] = split_pq(a, b, kPqRegs ? pq_reg[kPqRegs ? j : 0] : __ldg(pq + mm));
            }
            dit2<R, true>(y);                                         // pass 1 over n2 (m' = lane + 32 n2); its twiddles are part of the weights
            c2 u[R];
            if constexpr (kCrossLane) {
                // S_warp[R k + e] = sum_m' [W_MS^(m' e) y[e]] W_32^(m' k) over the lanes: lane l ends with samples R brev5(l) + e
                const float2* xw = tabx + lane * kXPitch;
                const float4 w01 = *reinterpret_cast<const float4*>(xw), w23 = *reinterpret_cast<const float4*>(xw + 2);
                const float2 wsh[4] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w), make_float2(w23.x, w23.y),
                                       make_float2(w23.z, w23.w)};
#pragma unroll
                for (int e = 0; e < R; e++) u[e] = warp_idft32(e ? cmul2s(y[e], xw[3 + e]) : y[e], wsh, lane);
                team_barrier(team);                   // every band read of the team is done: the planes are free
#pragma unroll
                for (int e = 0; e < R; e++) {
                    const int n = R * brev5(lane) + e;
                    pst(P, wbase + n, u[e]);
                }
            } else {
            team_barrier(team);                       // every band read of the team is done: the planes are free
            rscheme_pass2<R>(y, u, P, wbase, tabx, lane);    // u[q] = S_warp[k2p + R h + 32 q]
            const int k2p = lane / G, h = lane % G;
#pragma unroll
            for (int qq = 0; qq < R; qq++) {
                // lanes (k2p, h) and (k2p, h + 16 / R) are 16 samples apart = the same bank: bit 4 of n moves them apart
                // (phase E reads 16 consecutive n per half warp: still one wavefront)
                const int n = k2p + R * h + 32 * qq;
                pst(P, wbase + (n ^ (((n >> 4) & 1) << (RB - 1))), u[qq]);
            }
            }
        }
        if constexpr (D <= 5) {
            team_barrier(team);
            qn = s_pair[team];
            load_half(p, qn < n_pairs ? qn : q, tt, 0, w0);           // in flight during E (past the end: this pair again, unused)
            // ---------------- E: last radix-4 stage across the planes: y[n' + MS a] = sum_r i^(r a) S_r[n'], a = 0..2 kept
            constexpr int MS = M / 4;
            c2 ph0;
            if constexpr (kMix == 2) ph0 = direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + tt, ob.base + tt);
            const KeepPair head{oa.lo == 0, ob.live && ob.lo == 0}, rest{true, ob.live};
            float2* const outA = p.out + oa.base;
            float2* const outB = p.out + ob.base;
#pragma unroll
            for (int i = 0; i < (MS + kTeamThreads - 1) / kTeamThreads; i++) {
                const int n = tt + kTeamThreads * i;
                if (MS >= kTeamThreads || n < MS) {
                    // S_r[n] W_mfft^(-r n) = S_r[n] conj(W_4096^(r (n << D))): applied here, where all 128 threads share the work
                    // (in phase D2 warp 3 carried 32 x (load + 7 scalar + 4 packed) and warp 0 nothing: the team waited for it)
                    const float2 w1 = twq1[n << D], w2 = csqr(w1), w3 = cmul(w1, w2);
                    constexpr int kSwz = (D >= 1 && D <= 3) ? ilog2(MS / 32) - 1 : -1;   // the R-scheme's bank swizzle (phase D2)
                    const int ns = kSwz >= 0 ? n ^ (((n >> 4) & 1) << (kSwz >= 0 ? kSwz : 0)) : n;
                    const c2 s0 = pld(P, ns), s1 = cmul_conj2s(pld(P, kWPlane + ns), w1), s2 = cmul_conj2s(pld(P, 2 * kWPlane + ns), w2),
                             s3 = cmul_conj2s(pld(P, 3 * kWPlane + ns), w3);
                    const c2 a0 = cadd(s0, s2), a1 = csub(s0, s2), a2 = cadd(s1, s3), a3 = csub(s1, s3);
                    if constexpr (kMix == 2) {
                        // sample n + MS a is 32 (4 i + (MS / 32) a) samples after sample tt: entries of nco_rot (< 96)
                        store_pair_rows_mixed(outA, outB, head, n, cadd(a0, a2), ph0, __ldg(nco.nco_rot + 4 * i), p.imag_sign);
                        store_pair_rows_mixed(outA, outB, rest, n + MS, c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)}, ph0,
                                              __ldg(nco.nco_rot + 4 * i + MS / 32), p.imag_sign);
                        store_pair_rows_mixed(outA, outB, rest, n + 2 * MS, csub(a0, a2), ph0, __ldg(nco.nco_rot + 4 * i + 2 * (MS / 32)),
                                              p.imag_sign);
                    } else {
                        store_pair_rows(outA, outB, head, n, cadd(a0, a2), sgn);
                        store_pair_rows(outA, outB, rest, n + MS, c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)}, sgn);      // a1 + i a3
                        store_pair_rows(outA, outB, rest, n + 2 * MS, csub(a0, a2), sgn);
                    }
                }
            }
        } else {
            // d = 6: mfft = 64, warp 0 alone (R = 2 register scheme over the whole band)
            qn = s_pair[team];
            load_half(p, qn < n_pairs ? qn : q, tt, 0, w0);
            if (warp == 0) {
                constexpr int R = M / 32, G = 32 / R, RB = ilog2(R);
                c2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int m = lane + 32 * j;
                    const int k = (tb + (m < M / 2 ? m : m - M)) & (kWHalf - 1);
                    const c2 a = pld(P, waddr(k));
                    const c2 b = pld(P, waddr((kWHalf - k) & (kWHalf - 1)));
                    y[brev_n(j, RB)] = split_pq(a, b, __ldg(p.pq + (k & 3) * 1024 + (k >> 2)));
                }
                __syncwarp();
                dit2<R, true>(y);
                c2 u[R];
                rscheme_pass2<R>(y, u, P, 0, tabx, lane);
                const int k2p = lane / G, h = lane % G;
                c2 ph0;
                if constexpr (kMix == 2) ph0 = direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + k2p + R * h, ob.base + k2p + R * h);
#pragma unroll
                for (int qq = 0; qq < R; qq++) {
                    const int idx = k2p + R * h + 32 * qq;
                    if (idx < 3 * M / 4) {
                        if constexpr (kMix == 2) store_pair_mixed(p.out, oa, ob, idx, u[qq], ph0, __ldg(nco.nco_rot + qq), p.imag_sign);
                        else store_pair(p.out, oa, ob, idx, u[qq], sgn);
                    }
                }
            }
        }
        team_barrier(team);                                           // the planes are rewritten by the next pair
        q = qn;
    }
}

template <int D, int kMix>
static cudaError_t setup_wp() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_wide_pair_kernel<D, kMix>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide_smem_bytes<D>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_wide_pair_kernel<D, kMix>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <int D, int kMix>
static cudaError_t launch_wp(const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long grid = sm_count;                        // persistent: one CTA per SM
    const long long want = (n_pairs + kTeams - 1) / kTeams;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_wide_pair_kernel<D, kMix><<<(unsigned)grid, kTeams * kTeamThreads, wide_smem_bytes<D>(), stream>>>(p, nco);
    return cudaGetLastError();
}

// One translation unit (one device module) per decimation and mixer form: their
// placement is then independent of their siblings (loop bodies of 45..60 KB against a 32 KB instruction cache shared by
// three free-running teams; measured round 2: placement inside a module, 0..20 KB of padding, changes nothing), and the
// fourteen kernels compile in parallel.  r2iq_wide_pair_d<D>[m].cu define R2IQ_WIDE_TU_D (and R2IQ_WIDE_TU_MIX) and
// include this file; compiled on its own, this file holds the dispatchers only.
#ifdef R2IQ_WIDE_TU_D
#ifdef R2IQ_WIDE_TU_MIX
#define R2IQ_WIDE_TU_FN(name, d) name##_mix_d##d
constexpr int kTuMix = 2;
#else
#define R2IQ_WIDE_TU_FN(name, d) name##_d##d
constexpr int kTuMix = 0;
#endif
#define R2IQ_WIDE_TU_FN2(name, d) R2IQ_WIDE_TU_FN(name, d)
cudaError_t R2IQ_WIDE_TU_FN2(setup_wide, R2IQ_WIDE_TU_D)() { return setup_wp<R2IQ_WIDE_TU_D, kTuMix>(); }
cudaError_t R2IQ_WIDE_TU_FN2(launch_wide, R2IQ_WIDE_TU_D)(const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream) {
    return launch_wp<R2IQ_WIDE_TU_D, kTuMix>(p, nco, sm_count, stream);
}
#else
#define R2IQ_WIDE_FOR_ALL(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6)
#define X(D)                                                                                  \
    cudaError_t setup_wide_d##D();                                                            \
    cudaError_t setup_wide_mix_d##D();                                                        \
    cudaError_t launch_wide_d##D(const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream);     \
    cudaError_t launch_wide_mix_d##D(const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream);
R2IQ_WIDE_FOR_ALL(X)
#undef X

cudaError_t setup_wide_kernels() {
    cudaError_t e;
#define X(D)                                                    \
    if ((e = setup_wide_d##D()) != cudaSuccess) return e;      \
    if ((e = setup_wide_mix_d##D()) != cudaSuccess) return e;
    R2IQ_WIDE_FOR_ALL(X)
#undef X
#ifdef R2IQ_AB_VARIANTS
    if ((e = setup_wide_gen1_kernels()) != cudaSuccess) return e;
#endif
    return cudaSuccess;
}

cudaError_t launch_wide(int d, const WideParams& p, const WideNco& nco, int sm_count, cudaStream_t stream) {
#ifdef R2IQ_AB_VARIANTS
    // the four-warps-per-frame scalar kernels have no mixer epilogue
    if (kernel_generation() != 0 && !nco.nco_rot) return launch_wide_gen1(d, p, sm_count, stream);
#endif
    // direct-phase mixer in the store epilogue when nco_rot is set (the reference-bit mixer is a separate pass)
#define X(D) \
    if (d == D) return nco.nco_rot ? launch_wide_mix_d##D(p, nco, sm_count, stream) : launch_wide_d##D(p, nco, sm_count, stream);
    R2IQ_WIDE_FOR_ALL(X)
#undef X
    return cudaErrorInvalidValue;
}

const char* wide_kernel_name(int d, bool direct_mixer) {
    static char names[2][kNDec][48];
    if (d < 0 || d >= kNDec) return "";
    char* n = names[direct_mixer ? 1 : 0][d];
    if (!n[0]) snprintf(n, sizeof(names[0][0]), "r2iq_wide_pair_kernel<%d, %d>", d, direct_mixer ? 2 : 0);
    return n;
}
#endif
#endif   // R2IQ_WIDE_HELPERS_ONLY

}  // namespace r2iq
