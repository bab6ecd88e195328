// r2iq_channels.cu -- wideband channelizer: many tuned DDC channels from ONE ADC stream (BASELINE config 4).
//
// The reference runs one fft_mt_r2iq per tuned channel, so C channels cost C forward FFTs of every frame
// (Core/fft_mt_r2iq_impl.hpp:84-98).  Here a warp transforms a frame PAIR once (packed two-frame lanes, TMA-staged
// input, as r2iq_pair.cu), forms the real-FFT split once, and serves every channel from it with the split / filter /
// edge zero-fill arithmetic of the single-channel kernels (impl.hpp:76-96); per channel the result is the stream a
// separate operator tuned to that bin would produce.  Output is channel-major: channel c writes out[c*out_stride + ...].
//   d = 5, 6 (mfft 32, 16): r2iq_channels_lane_kernel -- a LANE per channel, backward transform in its registers, one TMA
//                           bulk store per channel run;
//   d = 2, 3, 4           : r2iq_channels_pair_kernel -- channels one after the other, backward transform across the lanes;
//   d = 0, 1 (and FFTN 8192): one single-channel launch per channel (r2iq_capi.cu: launch_channels_impl).
// (The first-generation kernel -- one scalar frame per warp, groups of 2^d channels cut, staged and transformed back
// together -- is gone: 8.96 / 4.21 / 2.85 ms for 64 channels x 2048 blocks at d = 2 / 3 / 4 against 2.97 / 1.58 / 0.96.)
#include "dual_frame.cuh"
#include "r2iq_common.cuh"
#include "work_dispenser.cuh"

namespace r2iq {

// ---------------------------------------------------------------------------------------------------------------
// d = 5 and 6 (mfft = 32 / 16: the "many narrow channels" case): ONE LANE PER CHANNEL on the packed two-frame lanes of
// the current single-channel kernels (r2iq_pair.cu).  Per warp and frame pair:
//   1. the pair arrives by one TMA bulk copy; forward 1024-point transform, two frames per lane (dual_frame.cuh);
//   2. the real-FFT split is channel independent: 2 X[k] = (Z[k] + conj Z[1024-k]) + c[k] (Z[k] - conj Z[1024-k]) is
//      formed ONCE for all 1024 bins and left in shared memory as (re_A, re_B, im_A, im_B) quads;
//   3. 32 channels at a time, lane = channel: the lane gathers its mfft band bins around its own tune bin with one
//      LDS.128 each (zero outside [0, 1024): Core/fft_mt_r2iq_impl.hpp:91-96), multiplies by 0.5 H_d[m] and runs the
//      mfft-point backward transform entirely in its registers (no shuffles, no transposition);
//   4. the kept 3/4 of both frames is one contiguous run of the channel's output stream (frame k's samples follow
//      frame k-1's, also across block edges: impl.hpp:117-138), so the lane lays its run out in a shared-memory row
//      and hands it to the TMA engine as ONE bulk store of up to 384 bytes.  (Direct 16-byte stores from 32 lanes hit
//      32 different lines per instruction: ncu showed 40 % of all stall samples on the MIO queue behind them.)
// Per channel and frame that is 4 packed multiplies per bin plus one packed register DFT, against one whole band cut
// and staged multi-lane inverse per GROUP in the kernel above.
// ---------------------------------------------------------------------------------------------------------------
namespace {
using namespace dualf;
constexpr int kLaneWarps = 7;
template <int D>
struct LaneGeo {
    static constexpr int M = kHalfFft >> D;
    static constexpr int KEEP = 3 * M / 4;
    static constexpr int kRowBytes = 2 * KEEP * 8 + 16;                  // both frames' kept samples + 16 B of skew
    static constexpr int kWarpBytes = kTile * 16 + 32 * kRowBytes + 16;  // tile | 32 rows | mbarrier
    static constexpr int kSmemBytes = (1024 + 1024 + 32) * 8 + 1024 + kLaneWarps * kWarpBytes;
};
constexpr int kLaneTabFloat2 = 1024 + 1024 + 32;                         // tw_fwd | ctab | gh
constexpr int kLaneMaxListed = 256;                                      // tune bins kept in shared memory (1 KB)
// Slot of spectrum bin k in the quad array.  The lanes of a quarter-warp (one LDS.128 wavefront) read bins that differ
// by the channel spacing -- a multiple of 4, typically 16 -- which in natural order is one bank group for all of them
// (measured: 61 % of the kernel's shared-memory wavefronts were conflict replays).  XOR-folding bits 3..8 of k into the
// three bits that choose the 16-byte column spreads spacings of 4, 8, 16, 32 and 64 bins over all eight columns.
__device__ __forceinline__ int quad_slot(int k) { return k ^ (((k >> 3) ^ (k >> 6)) & 7); }
}  // namespace

template <int D>
__global__ void __launch_bounds__(kLaneWarps * 32, 1) r2iq_channels_lane_kernel(const ChannelParams p) {
    using LG = LaneGeo<D>;
    constexpr int M = LG::M;                  // 32 or 16
    constexpr int MB = ilog2(M);
    constexpr int KEEP = LG::KEEP;
    extern __shared__ __align__(128) unsigned char lane_smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float2* tab = reinterpret_cast<float2*>(lane_smem);
    const float2* ctab = tab + 1024;
    const float2* gh = tab + 2048;
    int* listed = reinterpret_cast<int*>(lane_smem + kLaneTabFloat2 * 8);
    unsigned char* wbase = lane_smem + kLaneTabFloat2 * 8 + 1024 + (size_t)warp * LG::kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    float4* X4 = reinterpret_cast<float4*>(wbase);                       // the split spectrum reuses the tile
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase);      // ... and so does the incoming pair
    float4* row = reinterpret_cast<float4*>(wbase + kTile * 16 + lane * LG::kRowBytes);
    const uint32_t mbar = smem_u32(wbase + LG::kWarpBytes - 16);
    const uint32_t dst = smem_u32(inA);

    for (int i = threadIdx.x; i < 1024; i += kLaneWarps * 32) {
        tab[i] = __ldg(p.tw_fwd + i);
        tab[1024 + i] = __ldg(p.ctab + i);
    }
    if (threadIdx.x < M) tab[2048 + threadIdx.x] = __ldg(p.gh + threadIdx.x);
    const bool in_smem = p.n_channels <= kLaneMaxListed;
    if (in_smem)
        for (int i = threadIdx.x; i < p.n_channels; i += kLaneWarps * 32) listed[i] = __ldg(p.tunebins + i);
    if (lane == 0) {
        mbar_init(mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    FusedParams fp;                 // frame_src only reads in_words
    fp.in_words = p.in_words;
    const long long n_pairs = (p.total_frames + 1) / 2;
    auto issue = [&](long long pair) {
        const long long fa = 2 * pair;
        const uint32_t* src = frame_src(fp, fa);
        uint32_t bytes = kFrameBytes;
        if (fa + 1 < p.total_frames) bytes += (uint32_t)(frame_src(fp, fa + 1) - src) * 4;
        mbar_expect_tx(mbar, bytes);
        tma_load_1d(dst, src, bytes, mbar);
    };
    unsigned long long g0 = 0, g1 = 0;
    if (lane == 0) {
        g0 = atomicAdd(p.work_counter, 1ULL);
        g1 = atomicAdd(p.work_counter, 1ULL);
    }
    long long q = (long long)__shfl_sync(0xffffffffu, g0, 0);
    unsigned long long pend = g1;
    if (lane == 0 && q < n_pairs) issue(q);
    uint32_t phase = 0;
    const float2* twf = tab + lane;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;

    while (q < n_pairs) {
        const long long qn = (long long)__shfl_sync(0xffffffffu, pend, 0);
        if (lane == 0) pend = atomicAdd(p.work_counter, 1ULL);
        const int offB = ((2 * q) % kFramesPerBlock == kFramesPerBlock - 1) ? kHalfFft / 2 : kHop / 2;
        const uint32_t* inB = inA + offB;
        c2 v[32];
        mbar_wait(mbar, phase);
        phase ^= 1;
#pragma unroll
        for (int j = 0; j < 32; j++) unpack_words(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
        __syncwarp();                             // every lane holds its words: the tile may be overwritten

        // ---- forward 1024-point transform (as r2iq_pair.cu) ---------------------------------------------------
        dit2<32, false>(v);
#pragma unroll
        for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], twf[32 * k2]);
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        __syncwarp();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
        __syncwarp();
        dit2<32, false>(v);                       // v[k1] = Z[32*k1 + lane]

        // ---- channel-independent split: v[k1] <- 2 X[32*k1 + lane] ---------------------------------------------
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
        __syncwarp();
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) {
            const int k = 32 * k1 + lane;
            const c2 a = v[k1];
            const c2 b = lds2(S, (kHalfFft - k) & (kHalfFft - 1));
            const float2 c = ctab[k];
            const u64 ux = add2(a.re, b.re), uy = sub2(a.im, b.im);      // a + conj b
            const u64 wx = sub2(a.re, b.re), wy = add2(a.im, b.im);      // a - conj b
            c2 x;
            x.re = fma2(bc(c.x), wx, fma2(bc(-c.y), wy, ux));
            x.im = fma2(bc(c.x), wy, fma2(bc(c.y), wx, uy));
            v[k1] = x;
        }
        __syncwarp();                             // every mirrored read is done: the tile becomes the quad array
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) {
            float ra, rb, ia, ib;
            upk(v[k1].re, ra, rb);
            upk(v[k1].im, ia, ib);
            X4[quad_slot(32 * k1 + lane)] = make_float4(ra, rb, ia, ib);
        }
        __syncwarp();

        const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
        const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
        const int len_a = KEEP - oa.lo;                                   // samples of the run that come from frame A
        const uint32_t run_bytes = (uint32_t)(len_a + (ob.live ? KEEP - ob.lo : 0)) * 8u;
        // ---- 32 channels per pass, lane = channel ----------------------------------------------------------------
        for (int c0 = 0; c0 < p.n_channels; c0 += 32) {
            const int ch = c0 + lane;
            const bool live = ch < p.n_channels;
            const int tb = !live ? 0 : in_smem ? listed[ch] : __ldg(p.tunebins + ch);
            c2 t[M];
#pragma unroll
            for (int m = 0; m < M; m++) {
                const int s = m < M / 2 ? m : m - M;
                const int k = tb + s;
                const bool inside = k >= 0 && k < kHalfFft;              // outside: zero fill (impl.hpp:91-96)
                const float4 x = X4[quad_slot(k & (kHalfFft - 1))];
                const float2 h = gh[m];
                const float hx = inside ? h.x : 0.0f, hy = inside ? h.y : 0.0f;
                const u64 xr = pk(x.x, x.y), xi = pk(x.z, x.w);
                c2 y;
                y.re = fma2(bc(hx), xr, mul2(bc(-hy), xi));
                y.im = fma2(bc(hx), xi, mul2(bc(hy), xr));
                t[brev_n(m, MB)] = y;
            }
            if (c0 + 32 >= p.n_channels) {
                // last pass: the quad array is dead once every lane has gathered -> the next pair may land in the tile
                __syncwarp();
                if (lane == 0 && qn < n_pairs) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(qn);
                }
            }
            dit2<M, true>(t);                     // t[n] = y[n] of this lane's channel, both frames
            // the row may still be being read by the previous pass's bulk store
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
#pragma unroll
            for (int n = 0; n < KEEP; n += 2) {
                float a0r, b0r, a0i, b0i, a1r, b1r, a1i, b1i;
                upk(t[n].re, a0r, b0r);
                upk(t[n].im, a0i, b0i);
                upk(t[n + 1].re, a1r, b1r);
                upk(t[n + 1].im, a1i, b1i);
                if (n >= oa.lo) row[(n - oa.lo) >> 1] = make_float4(a0r, flip(a0i, sgn), a1r, flip(a1i, sgn));
                if (ob.live && n >= ob.lo) row[(len_a + n - ob.lo) >> 1] = make_float4(b0r, flip(b0i, sgn), b1r, flip(b1i, sgn));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            if (live) {
                float2* run = p.out + (long long)ch * p.out_stride + oa.base + oa.lo;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(run), "r"(smem_u32(row)), "r"(run_bytes)
                             : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        q = qn;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");    // shared memory must outlive the engine's reads
}

template <int D>
static cudaError_t setup_ch_lane() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_channels_lane_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, LaneGeo<D>::kSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_channels_lane_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <int D>
static cudaError_t launch_ch_lane(const ChannelParams& p, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long grid = sm_count;                    // persistent: one CTA per SM
    const long long want = (n_pairs + kLaneWarps - 1) / kLaneWarps;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_channels_lane_kernel<D><<<(unsigned)grid, kLaneWarps * 32, LaneGeo<D>::kSmemBytes, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// d = 2, 3, 4 (mfft 256, 128, 64): the channelizer on the packed two-frame generation (round 2).
//   1. forward transform of a frame PAIR exactly as r2iq_pair.cu (TMA pair copy, integer-pipe conversion with the bias
//      folded into the first butterfly stage, two radix-32 register passes, inter-pass twiddles folded into the second);
//   2. the real-FFT split ONCE for all channels: X'[k] = (a + conj b) + c[k] (a - conj b), a = Z[k], b = Z[1024 - k],
//      c = -i W_2048^k (the factor 1/2 lives in gh), written back to the warp's tile de-interleaved by k mod R;
//   3. per channel the mfft-point backward transform is split as band bin m = R a + j, a ACROSS THE LANES and j in
//      registers, output sample n = n_lo + 32 n_hi, n_lo across the lanes: lane a gathers its R consecutive band bins
//      around the channel's tune bin (zero outside [0, 1024): Core/fft_mt_r2iq_impl.hpp:91-96), multiplies by gh[m] (R
//      per-lane constants), runs R 32-point transforms ACROSS THE LANES (dual_frame.cuh: warp_idft32; lane l ends with
//      n_lo = brev5(l)), then the twiddle exp(+2 pi i j n_lo / mfft) folded into an R-point register DFT over j
//      (dit2_tw) -- nothing but the gather touches shared memory, so the spectrum stays where it is for the next
//      channel.  Every store instruction covers 32 CONSECUTIVE samples, the kept range [lo, 3 mfft / 4) is a
//      warp-uniform set of rows n_hi for d <= 3, and the quarter of the register DFT that overlap-save discards is
//      never computed.
//      (Until r02e the split was the other way round -- an R-point register DFT, then the cross-lane transforms, lane l
//      ending with the R consecutive samples R brev5(l) ...: 16-byte stores at a 64 / 32-byte stride from lanes in
//      bit-reversed order cost one L1 wavefront per LANE; at d = 2 they were 42 % of an L1 data pipe that ran at 98 %,
//      MIO throttle 2.9 cycles per issue: profiles/r02e_chan_kernels.json.  4.45 / 2.26 / 1.01 -> 2.97 / 1.58 / 0.96 ms
//      for 64 channels x 2048 blocks at d = 2 / 3 / 4.)
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kCpWarps = 12;
constexpr int kCpTabPitch = 34;
constexpr int kCpXPitch = 12;                                            // per-lane cross-lane twiddles: 4 + (R - 1) <= 11 entries
constexpr int kCpTabFloat2 = 32 * kCpTabPitch + 1024 + 32 * kCpXPitch;    // tw_fwd transposed | ctab | cross-lane twiddles
constexpr int kCpMaxListed = 256;
constexpr int kCpWarpBytes = kTile * 16 + 16;                            // tile (spectrum; the next pair lands in its tail) | mbarrier
constexpr int kCpSmemBytes = kCpTabFloat2 * 8 + kCpMaxListed * 4 + kCpWarps * kCpWarpBytes;
constexpr int kCpLandOff = kTile * 16 - (kHop + kFftN) * 2;
}  // namespace

template <int D>
__global__ void __launch_bounds__(kCpWarps * 32, 1) r2iq_channels_pair_kernel(const ChannelParams p) {
    constexpr int M = kHalfFft >> D;
    constexpr int R = M / 32;
    constexpr int RB = ilog2(R);
    // the split spectrum is kept de-interleaved by k mod R (pitch 1024 / R + 16 / R: conflict-free for the split's writes
    // and for the gather, whose lanes are R bins apart)
    constexpr int kSlotPitch = kHalfFft / R + 16 / R;
    auto slot = [](int k) { return (k & (R - 1)) * kSlotPitch + (k >> RB); };
    extern __shared__ __align__(128) unsigned char cp_smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float2* tab = reinterpret_cast<float2*>(cp_smem);
    const float2* twf = tab + lane * kCpTabPitch;
    const float2* ctab = tab + 32 * kCpTabPitch;
    const float2* xw = tab + 32 * kCpTabPitch + 1024 + lane * kCpXPitch;
    int* listed = reinterpret_cast<int*>(cp_smem + kCpTabFloat2 * 8);
    unsigned char* wbase = cp_smem + kCpTabFloat2 * 8 + kCpMaxListed * 4 + (size_t)warp * kCpWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + kCpLandOff);
    const uint32_t mbar = smem_u32(wbase + kCpWarpBytes - 16);
    const uint32_t dst = smem_u32(inA);

    for (int i = threadIdx.x; i < 1024; i += kCpWarps * 32) {
        tab[(i & 31) * kCpTabPitch + (i >> 5)] = __ldg(p.tw_fwd + i);
        tab[32 * kCpTabPitch + i] = __ldg(p.ctab + i);
    }
    // entries 0..3: the lane's stage twiddles of warp_idft32; entries 4.. are exp(+2 pi i n_lo j / mfft) for the lane's output residue n_lo = brev5(lane)
    for (int i = threadIdx.x; i < 32 * (3 + R); i += kCpWarps * 32)
        tab[32 * kCpTabPitch + 1024 + (i & 31) * kCpXPitch + (i >> 5)] =
            warp_idft_twiddle((i >> 5) >= 4 ? (int)brev5(i & 31) : (i & 31), i >> 5, M);
    const bool in_smem = p.n_channels <= kCpMaxListed;
    if (in_smem)
        for (int i = threadIdx.x; i < p.n_channels; i += kCpWarps * 32) listed[i] = __ldg(p.tunebins + i);
    if (lane == 0) {
        mbar_init(mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    FusedParams fp;                 // frame_src only reads in_words
    fp.in_words = p.in_words;
    const long long n_pairs = (p.total_frames + 1) / 2;
    auto issue = [&](long long pair) {
        const long long fa = 2 * pair;
        const uint32_t* src = frame_src(fp, fa);
        uint32_t bytes = kFrameBytes;
        if (fa + 1 < p.total_frames) bytes += (uint32_t)(frame_src(fp, fa + 1) - src) * 4;
        mbar_expect_tx(mbar, bytes);
        tma_load_1d(dst, src, bytes, mbar);
    };
    unsigned long long g0 = 0, g1 = 0, pend = 0;
    if (lane == 0) {
        g0 = atomicAdd(p.work_counter, 1ULL);
        g1 = atomicAdd(p.work_counter, 1ULL);
        pend = atomicAdd(p.work_counter, 1ULL);
    }
    long long q = (long long)__shfl_sync(0xffffffffu, g0, 0);
    long long qn = (long long)__shfl_sync(0xffffffffu, g1, 0);
    if (lane == 0 && q < n_pairs) issue(q);
    uint32_t phase = 0;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;

    while (q < n_pairs) {
        const long long qnn = (long long)__shfl_sync(0xffffffffu, pend, 0);
        if (lane == 0) pend = atomicAdd(p.work_counter, 1ULL);
        const int offB = ((uint32_t)(2 * q) % (uint32_t)kFramesPerBlock == kFramesPerBlock - 1) ? kHalfFft / 2 : kHop / 2;
        const uint32_t* inB = inA + offB;
        {
            c2 v[32];
            mbar_wait(mbar, phase);
            phase ^= 1;
#pragma unroll
            for (int j = 0; j < 32; j++) unpack_words_biased(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
            __syncwarp();
            // ---- forward 1024-point transform (r2iq_pair.cu) ----------------------------------------------------
            dit2_biased<32, false>(v);
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
            __syncwarp();
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
            __syncwarp();
            dit2_tw<32, false>(v, [&](int i) {
                const int n1 = brev5(i);
                const float4 w = *reinterpret_cast<const float4*>(twf + (n1 & ~1));
                return (n1 & 1) ? make_float2(w.z, w.w) : make_float2(w.x, w.y);
            });                                       // v[k1] = Z[32 k1 + lane]
            // ---- real-FFT split, once for all channels ------------------------------------------------------------
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
            __syncwarp();
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) {
                const int k = 32 * k1 + lane;
                const c2 a = v[k1];
                const c2 b = lds2(S, (kHalfFft - k) & (kHalfFft - 1));
                const float2 c = ctab[k];
                const u64 ore = sub2(a.re, b.re), oim = add2(a.im, b.im);      // a - conj b
                c2 x;
                x.re = fma2(bc(c.x), ore, fma2(bc(-c.y), oim, add2(a.re, b.re)));
                x.im = fma2(bc(c.x), oim, fma2(bc(c.y), ore, sub2(a.im, b.im)));
                v[k1] = x;
            }
            __syncwarp();   // every mirror read is done: the tile takes the split spectrum
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) sts2(S, slot(32 * k1 + lane), v[k1]);
            __syncwarp();
        }
        // ---- per channel: gather, filter, backward transform across the lanes, store ------------------------------
        float2 gh[R];
#pragma unroll
        for (int j = 0; j < R; j++) gh[j] = __ldg(p.gh + R * lane + j);
        const float4 w01 = *reinterpret_cast<const float4*>(xw), w23 = *reinterpret_cast<const float4*>(xw + 2);
        const float2 wsh[4] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w), make_float2(w23.x, w23.y), make_float2(w23.z, w23.w)};
        const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
        const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
        for (int ch = 0; ch < p.n_channels; ch++) {
            const int tb = in_smem ? listed[ch] : __ldg(p.tunebins + ch);
            {
                c2 t[R];
                const int m0 = tb + R * lane - (lane < 16 ? 0 : M);        // m = R lane + j; the upper half are the negative bins
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int kk = m0 + j;
                    const bool inside = (unsigned)kk < (unsigned)kHalfFft;   // bins outside [0, halfFft) are zero
                    const c2 x = lds2(S, slot(kk & (kHalfFft - 1)));
                    const float2 g = inside ? gh[j] : make_float2(0.0f, 0.0f);
                    // T_j[n_lo] = sum_a Y[R a + j] exp(+2 pi i a n_lo / 32): lane l ends with n_lo = brev5(l)
                    t[brev_n(j, RB)] = warp_idft32(cmul2s(x, g), wsh, lane);
                }
                // y[n_lo + 32 n_hi] = sum_j exp(+2 pi i j n_hi / R) [exp(+2 pi i j n_lo / mfft) T_j[n_lo]]
                if constexpr (R >= 4) {
                    dit2_tw<R, true>(t, [&](int i) { return xw[3 + brev_n(i, RB)]; });
                } else {                                  // R = 2: one butterfly, a' = a + w b, b' = 2 a - a'
                    const float2 w = xw[4];
                    const u64 ax = fma2(bc(w.x), t[1].re, fma2(bc(-w.y), t[1].im, t[0].re));
                    const u64 ay = fma2(bc(w.x), t[1].im, fma2(bc(w.y), t[1].re, t[0].im));
                    t[1].re = fma2(bc(2.0f), t[0].re, neg2(ax));
                    t[1].im = fma2(bc(2.0f), t[0].im, neg2(ay));
                    t[0].re = ax;
                    t[0].im = ay;
                }
                float2* const outA = p.out + (size_t)ch * p.out_stride + (oa.base + (int)brev5(lane));
                float2* const outB = p.out + (size_t)ch * p.out_stride + (ob.base + (int)brev5(lane));
#pragma unroll
                for (int nh = 0; nh < (3 * R + 3) / 4; nh++) {
                    float ar, br, ai, bi;
                    upk(t[nh].re, ar, br);
                    upk(t[nh].im, ai, bi);
                    // d <= 3: mfft / 4 and 3 mfft / 4 are whole rows of 32 samples (warp-uniform); d = 4: half rows, per lane
                    const int n = R >= 4 ? 32 * nh : 32 * nh + (int)brev5(lane);
                    const bool in_a = n >= oa.lo && (R >= 4 || n < 3 * M / 4);
                    const bool in_b = ob.live && n >= ob.lo && (R >= 4 || n < 3 * M / 4);
                    if (in_a) outA[32 * nh] = make_float2(ar, flip(ai, sgn));
                    if (in_b) outB[32 * nh] = make_float2(br, flip(bi, sgn));
                }
            }
        }
        __syncwarp();   // every lane is done with the spectrum: the tail of the tile takes the next pair of frames
        if (lane == 0 && qn < n_pairs) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(qn);
        }
        q = qn;
        qn = qnn;
    }
}

template <int D>
static cudaError_t setup_ch_pair() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_channels_pair_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCpSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_channels_pair_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int D>
static cudaError_t launch_ch_pair(const ChannelParams& p, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long grid = sm_count;                        // persistent: one CTA per SM
    const long long want = (n_pairs + kCpWarps - 1) / kCpWarps;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_channels_pair_kernel<D><<<(unsigned)grid, kCpWarps * 32, kCpSmemBytes, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t setup_channel_kernels() {
    cudaError_t e;
    if ((e = setup_ch_pair<2>()) != cudaSuccess) return e;
    if ((e = setup_ch_pair<3>()) != cudaSuccess) return e;
    if ((e = setup_ch_pair<4>()) != cudaSuccess) return e;
    if ((e = setup_ch_lane<5>()) != cudaSuccess) return e;
    if ((e = setup_ch_lane<6>()) != cudaSuccess) return e;
    return cudaSuccess;
}

cudaError_t launch_channels(int d, const ChannelParams& p, int sm_count, cudaStream_t stream) {
    switch (d) {
        case 2: return launch_ch_pair<2>(p, sm_count, stream);
        case 3: return launch_ch_pair<3>(p, sm_count, stream);
        case 4: return launch_ch_pair<4>(p, sm_count, stream);
        case 5: return launch_ch_lane<5>(p, sm_count, stream);
        case 6: return launch_ch_lane<6>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
