// r2iq_wide_channels.cu -- channelizer for FFTN = 8192 (the size the reference tree ships, Core/config.h:49) at d = 4, 5
// (mfft = 256, 128: the "many narrow channels" case of BASELINE config 4).
//
// One single-channel launch per channel (r2iq_capi.cu) repeats the whole 4096-point forward transform for every channel.
// Here a TEAM of four warps transforms a frame pair ONCE with the phases A and B of r2iq_wide_pair.cu, forms the real-FFT
// split once -- X'[k] = (a + conj b) + c[k] (a - conj b), a = Z[k], b = Z[4096 - k], c = -i W_8192^k -- and leaves it in
// its four planes (bin 4 m + r in plane r, slot m).  Then every WARP of the team serves whole channels (warp w the channels
// w, w + 4, ...), without a barrier or a byte of shared-memory traffic beyond the gather:
//   * it gathers the 32 (64) band bins of each of the four planes around the channel's tune bin (a multiple of 4, so the
//     band's bins 4 m' + r are plane r shifted by tunebin / 4; zero outside [0, 4096): Core/fft_mt_r2iq_impl.hpp:91-96),
//     multiplies by its per-lane filter constants 0.5 H_d and runs the four 32-point per-plane backward transforms across
//     its lanes (dual_frame.cuh: warp_idft32) -- four independent shuffle chains in flight;
//   * S_0..S_3 of its samples n = R brev5(lane) + e are then in its own registers: the last radix-4 stage across the planes
//     (twiddles W_4096^(-r (n << d)), per-lane constants) and the stores of the kept three quarters of both frames follow
//     in place.
// ~330 instructions per channel and frame pair instead of a whole single-channel kernel (~2300 per thread): 64 channels x
// 8192 blocks in 2.43 ms at d = 5 and 4.01 ms at d = 4 (64 single-channel launches: 44 / 46 ms; the first form of this
// kernel -- warp r on plane r of one channel, S_r exchanged through shared memory between two team barriers per channel,
// a dependent chain of ~1200 cycles per channel: 4.21 / 6.86 ms).
// Worker-loop semantics per channel: Core/fft_mt_r2iq_impl.hpp:76-138.
#define R2IQ_WIDE_HELPERS_ONLY
#include "r2iq_wide_pair.cu"

namespace r2iq {

namespace {
constexpr int kWcXPitch = 10;
constexpr int kWcListed = 256;
template <int D>
struct WcGeo {
    static constexpr int M = kWHalf >> D;             // 128 (d = 5), 256 (d = 4)
    static constexpr int MS = M / 4;                  // points per plane
    static constexpr int R = MS / 32;                 // band bins of a plane per lane
    // tw1024^T | lane twiddles | tune bins | W_4096^n (8 KB: phase A reads all of it, the channel loop only its own entries)
    static constexpr int kTabFloat2 = kWTab1 + 32 * kWcXPitch + kWcListed / 2 + 1024;
    static constexpr int kSmemBytes = kTabFloat2 * 8 + kTeams * kTeamBytes;
};
}  // namespace

template <int D>
__global__ void __launch_bounds__(kTeams* kTeamThreads, 1) r2iq_wide_channels_kernel(const WideChannelParams p) {
    using G = WcGeo<D>;
    constexpr int M = G::M;
    constexpr int MS = G::MS;
    constexpr int R = G::R;
    constexpr int RB = ilog2(R);
    extern __shared__ __align__(128) unsigned char wcsmem[];
    __shared__ long long s_pair[kTeams];
    const int team = threadIdx.x / kTeamThreads;
    const int tt = threadIdx.x % kTeamThreads;
    const int lane = tt & 31;
    const int warp = tt >> 5;                          // plane this warp transforms
    float2* tab = reinterpret_cast<float2*>(wcsmem);
    const float2* tw1024 = tab + lane * kWTabPitch;
    const float2* twq1 = tab + kWTab1 + 32 * kWcXPitch + kWcListed / 2;   // W_4096^n
    const float2* xw = tab + kWTab1 + lane * kWcXPitch;
    int* listed = reinterpret_cast<int*>(tab + kWTab1 + 32 * kWcXPitch);
    unsigned char* tbase = wcsmem + G::kTabFloat2 * 8 + (size_t)team * kTeamBytes;
    Planes P;
    P.re = reinterpret_cast<float2*>(tbase);
    P.im = P.re + 4 * kWPlane;
    const int wbase = warp * kWPlane;
    const int mbase = ((4 - warp) & 3) * kWPlane;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;

    for (int i = threadIdx.x; i < 1024; i += kTeams * kTeamThreads) {
        tab[(i & 31) * kWTabPitch + (i >> 5)] = __ldg(p.tw1024 + i);
        tab[kWTab1 + 32 * kWcXPitch + kWcListed / 2 + i] = __ldg(p.twq + i);
    }
    for (int i = threadIdx.x; i < 32 * (3 + R); i += kTeams * kTeamThreads)
        tab[kWTab1 + (i & 31) * kWcXPitch + (i >> 5)] = warp_idft_twiddle(i & 31, i >> 5, MS);
    const bool in_smem = p.n_channels <= kWcListed;
    if (in_smem)
        for (int i = threadIdx.x; i < p.n_channels; i += kTeams * kTeamThreads) listed[i] = __ldg(p.tunebins + i);
    const long long n_pairs = (p.total_frames + 1) / 2;
    if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);
    __syncthreads();

    const float4 w01 = *reinterpret_cast<const float4*>(xw), w23 = *reinterpret_cast<const float4*>(xw + 2);
    const float2 wsh[4] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w), make_float2(w23.x, w23.y), make_float2(w23.z, w23.w)};
    // only the WideParams fields load_half / quarter_stage read
    WideParams wp;
    wp.in_words = p.in_words;
    wp.total_frames = p.total_frames;

    long long q = s_pair[team];
    HalfWords w0;
    load_half(wp, q < n_pairs ? q : 0, tt, 0, w0);
    while (q < n_pairs) {
        // ---------------- A: radix-4 DIF over the four quarters of both frames (r2iq_wide_pair.cu) ------------------
        {
            HalfWords w1;
            load_half(wp, q, tt, 4, w1);
            quarter_stage(w0, P, twq1, p.rand_mul, tt, 0);
            quarter_stage(w1, P, twq1, p.rand_mul, tt, 4);
        }
        team_barrier(team);
        if (tt == 0) s_pair[team] = (long long)atomicAdd(p.work_counter, 1ULL);   // read again after the channel loop

        // ---------------- B: warp q transforms plane q; all rows back to the plane (mirror exchange) -------------------
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
        dit2_tw<32, false>(v, [&](int i) { return wtab_at(tw1024, brev5(i)); });
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) pst(P, wbase + lane + 32 * k1, v[k1]);
        team_barrier(team);

        // ---------------- split, once for all channels: bin k = 4 mm + warp, its mirror in plane (4 - warp) & 3 ---------
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) {
            const int mm = lane + 32 * k1;
            const c2 a = v[k1];
            const c2 b = pld(P, mbase + mirror_slot(warp, mm));
            const float2 c = __ldg(p.ctab + 4 * mm + warp);
            const u64 ore = sub2(a.re, b.re), oim = add2(a.im, b.im);      // a - conj b
            c2 x;
            x.re = fma2(bc(c.x), ore, fma2(bc(-c.y), oim, add2(a.re, b.re)));
            x.im = fma2(bc(c.x), oim, fma2(bc(c.y), ore, sub2(a.im, b.im)));
            v[k1] = x;
        }
        team_barrier(team);                           // every mirror read of the team is done: the planes take the split
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) pst(P, wbase + lane + 32 * k1, v[k1]);

        const FrameOut2 oa = wide_frame_out<D>(2 * q, p.total_frames);
        const FrameOut2 ob = wide_frame_out<D>(2 * q + 1, p.total_frames);
        const KeepPair head{oa.lo == 0, ob.live && ob.lo == 0}, rest{true, ob.live};
        // ---------------- per channel ------------------------------------------------------------------------------------
        // A WARP serves a whole channel (warp w the channels w, w + 4, ...): it gathers the band bins of all four planes
        // (read-only since the barrier above), runs the four per-plane transforms across its lanes -- four independent
        // shuffle chains in flight -- and has S_0..S_3 of its samples n = R brev5(lane) + e in its own registers, so the last
        // radix-4 stage needs no exchange and the channel loop no barrier.  (Until r02f warp r transformed plane r of ONE
        // channel, S_r met in a shared-memory exchange area between two team barriers per channel: one dependent chain of
        // ~1200 cycles per channel and team, barrier + scoreboard stalls 2.4 + 1.7 cycles per issue,
        // profiles/r02e_chan_kernels.json.)
        team_barrier(team);                           // every plane holds the split
        // (per pair, not once per kernel: these registers would otherwise be live across phase B's 32-value register pass)
        // this lane's R band positions inside a plane: slot offset s' in [-MS/2, MS/2); in plane r that is bin offset s = 4 s' + r,
        // filter index s mod M.  In the channel loop a warp serves whole channels: all four planes, R bins each.
        int sp[R];
        float2 g[4][R];
#pragma unroll
        for (int j = 0; j < R; j++) {
            const int m = lane + 32 * j;
            sp[j] = m < MS / 2 ? m : m - MS;
#pragma unroll
            for (int r = 0; r < 4; r++) g[r][j] = __ldg(p.gh + ((4 * sp[j] + r) & (M - 1)));
        }
        // twiddles of the last radix-4 stage for this lane's samples n = R brev5(lane) + e: W_4096^(r (n << d)), r = 1, 2, 3
        float2 wq[R][3];
#pragma unroll
        for (int e = 0; e < R; e++) {
            const float2 w1 = twq1[(R * brev5(lane) + e) << D];
            wq[e][0] = w1;
            wq[e][1] = csqr(w1);
            wq[e][2] = cmul(w1, wq[e][1]);
        }
        for (int ch = warp; ch < p.n_channels; ch += 4) {
            const int tb4 = (in_smem ? listed[ch] : __ldg(p.tunebins + ch)) >> 2;
            c2 S[4][R];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                c2 y[R];
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int slot = tb4 + sp[j];
                    const bool inside = (unsigned)slot < 1024u;               // bin 4 slot + r inside [0, 4096)
                    const c2 x = pld(P, r * kWPlane + (slot & 1023));
                    y[brev_n(j, RB)] = cmul2s(x, inside ? g[r][j] : make_float2(0.0f, 0.0f));
                }
                dit2<R, true>(y);                         // y[e] = sum_j Y[lane + 32 j] W_R^(j e)
#pragma unroll
                for (int e = 0; e < R; e++) S[r][e] = y[e];
            }
#pragma unroll
            for (int e = 0; e < R; e++) {
#pragma unroll
                for (int r = 0; r < 4; r++)               // S_r[R brev5(lane) + e]
                    S[r][e] = warp_idft32(e ? cmul2s(S[r][e], xw[3 + e]) : S[r][e], wsh, lane);
            }
            float2* const outA = p.out + (size_t)ch * p.out_stride + oa.base;
            float2* const outB = p.out + (size_t)ch * p.out_stride + ob.base;
            c2 row[3][R];
#pragma unroll
            for (int e = 0; e < R; e++) {
                // last radix-4 stage across the planes: y[n + MS a] = sum_r i^(r a) S_r[n] W_mfft^(-r n), a = 0..2 kept
                const c2 s0 = S[0][e], s1 = cmul_conj2s(S[1][e], wq[e][0]), s2 = cmul_conj2s(S[2][e], wq[e][1]),
                         s3 = cmul_conj2s(S[3][e], wq[e][2]);
                const c2 a0 = cadd(s0, s2), a1 = csub(s0, s2), a2 = cadd(s1, s3), a3 = csub(s1, s3);
                row[0][e] = cadd(a0, a2);
                row[1][e] = c2{sub2(a1.re, a3.im), add2(a1.im, a3.re)};      // a1 + i a3
                row[2][e] = csub(a0, a2);
            }
            if constexpr (R == 2) {
                // d = 4: lane l holds samples 2 brev5(l), + 1 -- 8-byte stores at a 16-byte stride from lanes in bit-reversed
                // order are 8 L1 wavefronts each (data-pipe rule 1).  24 SHFL put sample pair l into lane l:
                // six 16-byte stores of 512 consecutive bytes, 4 wavefronts each, instead of twelve at 8.
                const int src = brev5(lane);
#pragma unroll
                for (int a = 0; a < 3; a++) {
                    const KeepPair keep = a == 0 ? head : rest;
                    float4 va, vb;
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        float ar, br, ai, bi;
                        upk(row[a][e].re, ar, br);
                        upk(row[a][e].im, ai, bi);
                        ar = __shfl_sync(0xffffffffu, ar, src);
                        br = __shfl_sync(0xffffffffu, br, src);
                        ai = __shfl_sync(0xffffffffu, ai, src);
                        bi = __shfl_sync(0xffffffffu, bi, src);
                        if (e == 0) { va.x = ar; va.y = flip(ai, sgn); vb.x = br; vb.y = flip(bi, sgn); }
                        else { va.z = ar; va.w = flip(ai, sgn); vb.z = br; vb.w = flip(bi, sgn); }
                    }
                    // 16-byte aligned: sample index 2 lane + MS a, frame bases and the channel stride are even (header contract)
                    if (keep.a) *reinterpret_cast<float4*>(outA + 2 * lane + MS * a) = va;
                    if (keep.b) *reinterpret_cast<float4*>(outB + 2 * lane + MS * a) = vb;
                }
            } else {
#pragma unroll
                for (int e = 0; e < R; e++) {
                    const int n = R * brev5(lane) + e;
                    store_pair_rows(outA, outB, head, n, row[0][e], sgn);
                    store_pair_rows(outA, outB, rest, n + MS, row[1][e], sgn);
                    store_pair_rows(outA, outB, rest, n + 2 * MS, row[2][e], sgn);
                }
            }
        }
        const long long qn = s_pair[team];
        load_half(wp, qn < n_pairs ? qn : q, tt, 0, w0);
        team_barrier(team);                           // the planes are rewritten by the next pair
        q = qn;
    }
}

template <int D>
static cudaError_t setup_wc() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_wide_channels_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, WcGeo<D>::kSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_wide_channels_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <int D>
static cudaError_t launch_wc(const WideChannelParams& p, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long grid = sm_count;                        // persistent: one CTA per SM
    const long long want = (n_pairs + kTeams - 1) / kTeams;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_wide_channels_kernel<D><<<(unsigned)grid, kTeams * kTeamThreads, WcGeo<D>::kSmemBytes, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t setup_wide_channel_kernels() {
    cudaError_t e = setup_wc<4>();
    if (e != cudaSuccess) return e;
    return setup_wc<5>();
}

cudaError_t launch_wide_channels(int d, const WideChannelParams& p, int sm_count, cudaStream_t stream) {
    if (d == 4) return launch_wc<4>(p, sm_count, stream);
    if (d == 5) return launch_wc<5>(p, sm_count, stream);
    return cudaErrorInvalidValue;
}

}  // namespace r2iq
