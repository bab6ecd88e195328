// r2iq_pair.cu -- the default fused kernels for d = 0..5, FFTN 2048 (third generation of the dual-frame family).
//
// Two frames per warp on packed fp32x2 lanes as in r2iq_dual_d0.cu / r2iq_dual_dn.cu (same mathematics, same register
// passes).  What changed, each from a measurement of the round-1 kernels (profiles/r01_kernels.json, ncu source page):
//   * ONE persistent CTA per SM of W free-running warps (W = 8 at 255 registers for d = 0 and 1, W = 12 at 168 for
//     d >= 2); every warp takes its frame pairs one at a time from the device-wide counter and never meets a block
//     barrier after the prologue.
//   * The twiddle and filter tables live in SHARED memory, copied once per CTA.  The round-1 kernels read them with
//     LDG: 32 KB of tables against the ~26 KB of L1 that the shared-memory carve-out leaves -- l1tex sector hit rate
//     58 % at d = 0, as many L2->L1 sectors for tables as for the input itself, `long_scoreboard` the third-largest
//     stall.
//   * A frame pair is fetched by ONE TMA bulk copy: frames A and B of a pair are consecutive (hop 1536 or, across a
//     block edge, 1024 samples), so the pair is 7168 (6144) contiguous bytes instead of 2 x 4096 -- 12.5 % less
//     L2->SM traffic and the shared memory that makes room for the tables.
//   * int16 -> float on the integer pipe (dual_frame.cuh: unpack_words) instead of the conversion unit.
// Later in round 2 the loop bodies went on an instruction diet (the kernels are bound by issue slots, the FMA pipe and the
// shared-memory data pipe together, never by HBM; every instruction taken out showed up in the time -- DESIGN.md section 4):
//   * the kernel parameter block narrowed to what the kernel reads (PairParams; decided per decimation by measurement);
//   * 32-bit frame arithmetic, warp-uniform store predicates, kept rows as a compile-time set;
//   * inter-pass twiddles folded into the first butterfly stage of the following pass (dft_regs2.cuh: dit2_tw), backward
//     pass-1 twiddles folded into the pass-2 weights, the conversion's bias folded into the first stage (dit2_biased);
//   * (P, Q) in band order as one float4 table, twiddle tables transposed in shared memory (one LDS.128 per two entries);
//   * d = 4, 5: the backward transform across the lanes by shuffle butterflies (dual_frame.cuh: warp_idft32);
//   * d = 5: no second forward pass -- the one band bin and the one mirror bin of every spectrum column evaluated directly.
// Reference lines replaced: Core/fft_mt_r2iq_impl.hpp:40-54,84-138 (see r2iq_kernels.cu).
#include "dual_frame.cuh"
#include "r2iq_kernels.cuh"

namespace r2iq {

using namespace dualf;

namespace {

constexpr int kPairBytes = (kHop + kFftN) * 2;   // 7168: frames A and B of a pair, hop 1536 samples apart

// Shared memory of one CTA: [tables][W per-warp areas].
//   per warp, W = 8 : tile (2 planes x 8448 B) | pair buffer 7168 B | mbarrier 16 B
//   per warp, W = 12: tile | mbarrier; the pair lands in the tail of the tile once the tile is free
template <int W>
__host__ __device__ constexpr int pair_warp_bytes() {
    return kTile * 16 + (W == 8 ? kPairBytes : 0) + 16;
}
// tables: d = 0: tw_fwd, tw_inv, gp, gq (4 x 8 KB); d >= 1: tw_fwd (8 KB) + tw_aux ((R + 32) x 256 B); (P, Q) stay in
// global memory there, as ONE float4 table in band order behind `gp` (R <= 16 coalesced 16-byte loads per frame pair)
// d = 0 keeps its four tables TRANSPOSED, [lane][34]: rows k and k+1 of a lane are 16 contiguous bytes, one LDS.128
// (pitch 34 float2 = 272 B keeps the eight lanes of a quarter-warp on distinct bank groups)
constexpr int kTabPitch = 34;
template <int D>
__host__ __device__ constexpr int pair_table_float2() {
    // d = 5: [32 lanes][33] float4 coefficients of the directly evaluated bins + [32][kTabPitch] cross-lane twiddles
    return D == 0 ? 4 * 32 * kTabPitch : D == 1 ? 2048 : D >= 5 ? 32 * 33 * 2 + 32 * kTabPitch : 2 * 32 * kTabPitch;
}
template <int D, int W>
__host__ __device__ constexpr int pair_smem_bytes() {
    return pair_table_float2<D>() * 8 + W * pair_warp_bytes<W>();
}
constexpr int kLandOff = kTile * 16 - kPairBytes;   // W = 12: the pair occupies the last 7168 B of the tile

// backward transposition of the d >= 1 kernels: two planes of 32 x (R + 1) elements packed at the head of the tile
// (pitch R + 1, conflict-free), so that the tail of the tile is free for the next pair while it runs
template <int R>
__device__ __forceinline__ void sts2r(u64* S, int idx, c2 v) {
    float a, b, c, d;
    upk(v.re, a, b);
    upk(v.im, c, d);
    reinterpret_cast<float2*>(S)[idx] = make_float2(a, b);
    reinterpret_cast<float2*>(S)[32 * (R + 1) + idx] = make_float2(c, d);
}
template <int R>
__device__ __forceinline__ c2 lds2r(const u64* S, int idx) {
    const float2 r = reinterpret_cast<const float2*>(S)[idx];
    const float2 i = reinterpret_cast<const float2*>(S)[32 * (R + 1) + idx];
    c2 v;
    v.re = pk(r.x, r.y);
    v.im = pk(i.x, i.y);
    return v;
}

}  // namespace

// What the kernels really read, as their parameter block: FusedParams (the launchers' common currency, 152 bytes with
// the tables of the older generations and all mixer forms) narrowed to 88 bytes, the direct-phase mixer's three fields
// as a second parameter.  Code generation follows the size of the parameter block: with the wide block the d = 0 kernel
// carries 115 more integer instructions per frame pair (3443 vs 3315) and d = 1 73 more -- at their register limits
// ptxas re-derives addresses instead of keeping them.
struct PairParams {
    const uint32_t* in_words;
    float2* out;
    long long total_frames;
    const float2* tw_fwd;
    const float2* gp;
    const float2* gq;
    const float2* tw_inv;
    const float2* tw_aux;
    int tunebin;
    uint32_t row_mask;
    uint32_t rand_mul;
    float imag_sign;
    unsigned long long* work_counter;
};
struct PairNco {
    const float2* nco_rot;
    double nco_phase0;
    double nco_fc;
};

// Params: the parameter block (FusedParams or PairParams), NcoParams: where the direct-phase mixer's fields live (the same
// FusedParams object, or a PairNco)
template <int D, int W, int kMix, class Params, class NcoParams>
__device__ __forceinline__ void pair_body(const Params& p, const NcoParams& nco) {
    constexpr bool kLandInTile = W == 12;
    constexpr int M = kHalfFft >> D;
    constexpr int R = D == 6 ? 1 : 32 >> D;   // band bins per lane (d = 1..5; d = 6: one bin on each of 16 lanes)
    constexpr int G = 32 / R;             // lanes sharing one 32-point DFT in pass 2 of the backward transform
    constexpr int RB = ilog2(R);
    // backward transform by shuffle butterflies (warp_idft32) for mfft = 64, 32: 1021 -> 1117 and 1048 -> 1152 Gsamples/s at
    // d = 4, 5; at d = 3 (four transforms, 80 SHFL, stores of 32-byte runs) it loses, 992 -> 963
    constexpr bool kCrossLane = D == 4 || D == 5;
    // d = 5 (mfft = 32): every column of the spectrum holds ONE band bin and ONE mirror bin, so the second forward pass is
    // not run at all -- lane c evaluates those two bins directly from its column of the transposition, 2 x 32 complex
    // multiply-accumulates with per-lane coefficients exp(-2 pi i n1 k / 1024) (inter-pass twiddle included; host table per
    // tune bin) = 256 packed instructions instead of 124 + 388, and hands them to the lanes that need them by SHFL:
    // no spectrum in shared memory, no band cut.
    // d = 6 (mfft = 16) the same way: a column holds at most one band bin and one mirror bin (zero coefficients where it
    // holds none), and the 16-point backward transform runs inside each half warp.
    constexpr bool kDirect = D >= 5;
    // the conversion's bias removed in the first butterfly stage (dit2_biased): +1.0 ... +1.5 % at d = 1..3; -0.5 % at d = 0
    constexpr bool kBiased = D != 0;
    static_assert(D == 0 || 2 * 32 * (R + 1) * 8 <= kLandOff, "backward transposition overlaps the landing zone");
    static_assert(D != 0 || !kLandInTile, "d = 0 uses the whole tile until the last pass: W = 8 only");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float2* tab = reinterpret_cast<float2*>(smem_raw);
    constexpr int kWarpBytes = pair_warp_bytes<W>();
    unsigned char* wbase = smem_raw + pair_table_float2<D>() * 8 + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + (kLandInTile ? kLandOff : kTile * 16));
    const uint32_t mbar = smem_u32(wbase + kWarpBytes - 16);
    const uint32_t dst = smem_u32(inA);

    // ---- tables -> shared memory ------------------------------------------------------------------------------
    constexpr bool kTabT = D != 1;                    // tables transposed ([lane][34]); d = 1: [n][lane]
    constexpr int kTab2 = D >= 5 ? 32 * 33 * 2 : kTabT ? 32 * kTabPitch : 1024;   // second table (d = 5, 6: behind the coefficients)
    if constexpr (D == 0) {
        for (int i = threadIdx.x; i < 1024; i += W * 32) {
            const int o = (i & 31) * kTabPitch + (i >> 5);      // [row][lane] -> [lane][row]
            tab[o] = __ldg(p.tw_fwd + i);
            tab[32 * kTabPitch + o] = __ldg(p.tw_inv + i);
            tab[64 * kTabPitch + o] = __ldg(p.gp + i);
            tab[96 * kTabPitch + o] = __ldg(p.gq + i);
        }
    } else {
        // the forward twiddles and the pass-2 weights ([32][32] behind the R unused rows of tw_aux); transposed like
        // d = 0 (one LDS.128 per two entries) except at d = 1, where the aligned register quads cost spills
        // (measured: 656 -> 618 Gsamples/s at d = 1, +1.5 / +3.4 / +0.4 / +0.9 % at d = 2..5)
        if constexpr (kDirect) {
            // tw_aux carries the coefficient table [lane][n1] (float4: band bin, mirror bin); pitch 33 float4 in shared memory
            float4* cf = reinterpret_cast<float4*>(tab);
            for (int i = threadIdx.x; i < 1024; i += W * 32) cf[(i >> 5) * 33 + (i & 31)] = __ldg(reinterpret_cast<const float4*>(p.tw_aux) + i);
        } else {
        for (int i = threadIdx.x; i < 1024; i += W * 32) {
            const int o = kTabT ? (i & 31) * kTabPitch + (i >> 5) : i;
            tab[o] = __ldg(p.tw_fwd + i);
            if constexpr (!kCrossLane) tab[kTab2 + o] = __ldg(p.tw_aux + R * 32 + i);
        }
        }
        if constexpr (kCrossLane || kDirect) {   // the lane's twiddles of the cross-lane transform instead of the pass-2 weights
            if (threadIdx.x < 32 * (3 + R))
                tab[kTab2 + (threadIdx.x & 31) * kTabPitch + (threadIdx.x >> 5)] = warp_idft_twiddle(threadIdx.x & 31, threadIdx.x >> 5, M);
        }
    }
    if (lane == 0) {
        mbar_init(mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const long long n_pairs = (p.total_frames + 1) / 2;
    // one bulk copy per pair; B = A + hop (1536 samples, 1024 across a block edge).  The last pair of an odd total has
    // no B: only A is copied, B's lanes then transform stale bytes and are never stored (packed lanes are independent).
    auto issue = [&](long long pair) {
        const long long fa = 2 * pair;
        const uint32_t* src = frame_src(p, fa);
        uint32_t bytes = kFrameBytes;
        if (fa + 1 < p.total_frames) bytes += (uint32_t)(frame_src(p, fa + 1) - src) * 4;
        mbar_expect_tx(mbar, bytes);
        tma_load_1d(dst, src, bytes, mbar);
    };

    // ---- work: one pair at a time from the device-wide counter, one grab kept in flight ---------------------------
    // (nvcc compiles atomicAdd into its warp-aggregated form, whose SHFL waits for the atomic's round trip on the spot:
    // 4.3 % of the d = 0 kernel's stall samples sit on that instruction.  An inline-PTX atom that really stays in flight
    // changed nothing, 531 vs 532 Gsamples/s -- the other warps of the scheduler cover the wait: the kernels are bound by
    // issue slots and pipes, not by latency.  Measured round 2, tools/ncu_hot.py.)
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

    const float2* twf = tab + (kTabT ? lane * kTabPitch : lane);
    auto tw_at = [](const float2* t, int n) {         // entry n of this lane's table
        if constexpr (kTabT) {
            const float4 w = *reinterpret_cast<const float4*>(t + (n & ~1));   // one LDS.128 per two entries
            return (n & 1) ? make_float2(w.z, w.w) : make_float2(w.x, w.y);
        } else {
            return t[32 * n];
        }
    };
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;   // sideband inversion as a sign-bit flip

    while (q < n_pairs) {
        const long long qnn = (long long)__shfl_sync(0xffffffffu, pend, 0);
        if (lane == 0) pend = atomicAdd(p.work_counter, 1ULL);
        // word offset of frame B inside the pair buffer: 768, or 512 when A is the last frame of a block
        const int offB = ((uint32_t)(2 * q) % (uint32_t)kFramesPerBlock == kFramesPerBlock - 1) ? kHalfFft / 2 : kHop / 2;
        const uint32_t* inB = inA + offB;
        c2 v[32];
        mbar_wait(mbar, phase);
        phase ^= 1;
#pragma unroll
        for (int j = 0; j < 32; j++) {
            if constexpr (kBiased) unpack_words_biased(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
            else unpack_words(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
        }
        __syncwarp();   // every lane has consumed its words: the buffer may be overwritten
        if constexpr (!kLandInTile) {
            if (lane == 0 && qn < n_pairs) issue(qn);
        }

        // ---- forward 1024-point transform of z = x[2n] + i x[2n+1] ------------------------------
        // (128-bit stores of two neighbouring elements into a pitch-34 tile -- half the store instructions of the
        // transposition, no extra moves -- measured at -3.4 ... +0.9 % over the decimations: the tile keeps 64-bit accesses)
        // (both radix-32 passes are straight-line code on purpose: running them through one copy of the butterflies in
        // a two-trip loop takes 6 KB off a 33..55 KB loop body that shares a 32 KB instruction cache, and costs
        // 7..46 % of the throughput -- loop-carried register moves, a weaker static schedule; measured, round 2)
        if constexpr (kBiased) dit2_biased<32, false>(v);   // the conversion's bias leaves in the first butterfly stage
        else dit2<32, false>(v);
        if constexpr (kDirect) {
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
            __syncwarp();
            const float4* cf = reinterpret_cast<const float4*>(tab) + lane * 33;
            c2 zb, zm;                                // this column's band bin and mirror bin
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) {
                const c2 x = lds2(S, n1 * 33 + lane);
                const float4 c = cf[n1];
                if (n1 == 0) {
                    zb = cmul2s(x, make_float2(c.x, c.y));
                    zm = cmul2s(x, make_float2(c.z, c.w));
                } else {
                    zb.re = fma2(bc(c.x), x.re, fma2(bc(-c.y), x.im, zb.re));
                    zb.im = fma2(bc(c.x), x.im, fma2(bc(c.y), x.re, zb.im));
                    zm.re = fma2(bc(c.z), x.re, fma2(bc(-c.w), x.im, zm.re));
                    zm.im = fma2(bc(c.z), x.im, fma2(bc(c.w), x.re, zm.im));
                }
            }
            __syncwarp();   // every lane is done with the tile: its tail takes the next pair of frames
            if (lane == 0 && qn < n_pairs) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue(qn);
            }
            // band position m: bin tunebin + s(m) sits in column (tunebin + s(m)) mod 32, its mirror in column -(tunebin + s(m))
            // (d = 5: m = lane and s = m mod 32; d = 6: both half warps serve m = lane mod 16)
            const int mpos = lane & (M - 1);
            const int sband = mpos < M / 2 ? mpos : mpos - M;
            const int ca = (p.tunebin + sband) & 31, cb = (-(p.tunebin + sband)) & 31;
            c2 a, b;
            {
                float r0, r1, i0, i1;
                upk(zb.re, r0, r1);
                upk(zb.im, i0, i1);
                a.re = pk(__shfl_sync(0xffffffffu, r0, ca), __shfl_sync(0xffffffffu, r1, ca));
                a.im = pk(__shfl_sync(0xffffffffu, i0, ca), __shfl_sync(0xffffffffu, i1, ca));
                upk(zm.re, r0, r1);
                upk(zm.im, i0, i1);
                b.re = pk(__shfl_sync(0xffffffffu, r0, cb), __shfl_sync(0xffffffffu, r1, cb));
                b.im = pk(__shfl_sync(0xffffffffu, i0, cb), __shfl_sync(0xffffffffu, i1, cb));
            }
            const float4 pq = __ldg(reinterpret_cast<const float4*>(p.gp) + mpos);   // (P, Q) in band order
            c2 y;
            y.re = fma2(bc(pq.w), b.im, fma2(bc(pq.z), b.re, fma2(bc(-pq.y), a.im, mul2(bc(pq.x), a.re))));
            y.im = fma2(bc(pq.w), b.re, fma2(bc(-pq.z), b.im, fma2(bc(pq.y), a.re, mul2(bc(pq.x), a.im))));
            const float2* tw2 = tab + kTab2 + lane * kTabPitch;
            const float4 w01 = *reinterpret_cast<const float4*>(tw2), w23 = *reinterpret_cast<const float4*>(tw2 + 2);
            const float2 wsh[4] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w), make_float2(w23.x, w23.y), make_float2(w23.z, w23.w)};
            c2 o = warp_idft<D == 5 ? 5 : 4>(y, wsh, lane);   // output sample brev(lane) of the 32 (16) points
            const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
            const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
            const int idx0 = D == 5 ? brev5(lane) : (lane < 16 ? brev_n(lane, 4) : M);   // d = 6: the upper half warp stores nothing
            if constexpr (kMix == 2)
                o = direct_mix(o, direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + idx0, ob.base + idx0), make_float2(1.0f, 0.0f), p.imag_sign);
            float ar, br, ai, bi;
            upk(o.re, ar, br);
            upk(o.im, ai, bi);
            if (idx0 >= oa.lo && idx0 < 3 * M / 4) p.out[oa.base + idx0] = kMix == 2 ? make_float2(ar, ai) : make_float2(ar, flip(ai, sgn));
            if (ob.live && idx0 >= ob.lo && idx0 < 3 * M / 4) p.out[ob.base + idx0] = kMix == 2 ? make_float2(br, bi) : make_float2(br, flip(bi, sgn));
            q = qn;
            qn = qnn;
            continue;
        }
        // The inter-pass twiddles W_1024^(n1 lane) ride in the first stage of the second pass (dit2_tw, 32 packed
        // instructions fewer; the table is symmetric in (n1, k2), so the reader's lane takes the entry the writer's would
        // have).  Measured per decimation: +1.8 / +3.5 / +0.7 / +0.3 / -0.7 / -0.5 / +1.7 % at d = 0..6 -- d = 4, 5 keep
        // the multiplication on the writing side.
        constexpr bool kFoldTw = D != 4 && D != 5;
        if constexpr (!kFoldTw) {
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], tw_at(twf, k2));
        }
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        __syncwarp();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
        __syncwarp();
        if constexpr (kFoldTw) {
            dit2_tw<32, false>(v, [&](int i) {
                return tw_at(twf, brev5(i));
            });
        } else {
            dit2<32, false>(v);
        }                                         // v[k1] = Z[32*k1 + lane]

        if constexpr (D == 0) {
            // ---- real-FFT split + tune rotation + filter, in place ---------------------------------------
            const float2* twi = tab + 32 * kTabPitch + lane * kTabPitch;
            const float2* gp = tab + 64 * kTabPitch + lane * kTabPitch;
            const float2* gq = tab + 96 * kTabPitch + lane * kTabPitch;
            const int n1_inv = (lane - p.tunebin) & 31;
            c2 t[32];
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
            __syncwarp();
#pragma unroll
            for (int k1 = 0; k1 < 32; k1 += 2) {
                const float4 P = *reinterpret_cast<const float4*>(gp + k1);
                const float4 Q = *reinterpret_cast<const float4*>(gq + k1);
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const float Px = e ? P.z : P.x, Py = e ? P.w : P.y, Qx = e ? Q.z : Q.x, Qy = e ? Q.w : Q.y;
                    const c2 a = v[k1 + e];
                    const c2 b = lds2(S, (kHalfFft - (32 * (k1 + e) + lane)) & (kHalfFft - 1));
                    c2 y;
                    y.re = fma2(bc(Qy), b.im, fma2(bc(Qx), b.re, fma2(bc(-Py), a.im, mul2(bc(Px), a.re))));
                    y.im = fma2(bc(Qy), b.re, fma2(bc(-Qx), b.im, fma2(bc(Py), a.re, mul2(bc(Px), a.im))));
                    t[brev5(k1 + e)] = y;             // bit-reversed placement for the backward pass
                }
            }
            __syncwarp();
            // ---- backward 1024-point transform; lane l is residue (l - tunebin) mod 32 ---------------
            dit2<32, true>(t);
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) sts2(S, n1_inv * 33 + k2, t[k2]);
            __syncwarp();
#pragma unroll
            for (int n = 0; n < 32; n++) t[brev5(n)] = lds2(S, n * 33 + lane);
            __syncwarp();
            // inter-pass twiddles on the reading side, folded into the first stage (tw_inv is laid out for that: row n
            // holds the twiddle of the element the lane reads from row n)
            dit2_tw<32, true, false>(t, [&](int i) {   // row 0's twiddle is 1 only when the tune bin is below 32
                const int n = brev5(i);
                const float4 w = *reinterpret_cast<const float4*>(twi + (n & ~1));
                return (n & 1) ? make_float2(w.z, w.w) : make_float2(w.x, w.y);
            });                                       // t[k1] = y[lane + 32*k1] for both frames

            // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
            const FrameOut2 oa = frame_out2<0>(2 * q, p.total_frames);
            const FrameOut2 ob = frame_out2<0>(2 * q + 1, p.total_frames);
            c2 ph0;
            if constexpr (kMix == 2) ph0 = direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + lane, ob.base + lane);
            // lo is 0 or 256 = eight whole rows: every store's predicate is warp-uniform and computed once
            const bool headA = oa.lo == 0, liveB = ob.live, headB = liveB && ob.lo == 0;
            float2* const outA = p.out + (oa.base + lane);
            float2* const outB = p.out + (ob.base + lane);
#pragma unroll
            for (int k1 = 0; k1 < 24; k1++) {     // idx >= 768 is always discarded
                if constexpr (kMix == 2) t[k1] = direct_mix(t[k1], ph0, __ldg(nco.nco_rot + k1), p.imag_sign);
                float ar, br, ai, bi;
                upk(t[k1].re, ar, br);
                upk(t[k1].im, ai, bi);
                if (k1 >= 8 || headA) {
                    float2 va = kMix == 2 ? make_float2(ar, ai) : make_float2(ar, flip(ai, sgn));
                    outA[32 * k1] = va;
                }
                if (k1 >= 8 ? liveB : headB) {
                    float2 vb = kMix == 2 ? make_float2(br, bi) : make_float2(br, flip(bi, sgn));
                    outB[32 * k1] = vb;
                }
            }
        } else {
            const float2* tw2 = tab + kTab2 + (kTabT ? lane * kTabPitch : lane);   // exp(+2 pi i ((lane % G) n1 / 32 + n1 (lane / G) / M))
            // ---- spectrum to shared memory in natural order; only the rows the band cut will read -------
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++)
                if ((p.row_mask >> k1) & 1u) sts2(S, 32 * k1 + lane, v[k1]);
            __syncwarp();
            // ---- band cut in backward-transform order: m = lane + 32*j, bin k = tunebin + s(m) --------
            c2 y[R];
#pragma unroll
            for (int j = 0; j < R; j++) {
                const int m = lane + 32 * j;
                const int s = m < M / 2 ? m : m - M;                     // upper half of the band is below the tune bin
                const int kk = (p.tunebin + s) & (kHalfFft - 1);         // wrapped bins meet P = Q = 0: zero fill
                const c2 a = lds2(S, kk);
                const c2 b = lds2(S, (kHalfFft - kk) & (kHalfFft - 1));
                const float4 pq = __ldg(reinterpret_cast<const float4*>(p.gp) + m);   // (P, Q) in band order
                const float2 P = make_float2(pq.x, pq.y), Q = make_float2(pq.z, pq.w);
                c2 r;
                r.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
                r.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
                y[brev_n(j, RB)] = r;                                    // bit-reversed placement
            }
            __syncwarp();   // every lane is done with the spectrum: the tile is free
            if constexpr (kLandInTile) {
                // ... its tail for the next pair of frames (generic-proxy reads ordered before the async-proxy writes),
                // its head for the backward transposition
                if (lane == 0 && qn < n_pairs) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(qn);
                }
            }
            if constexpr (kCrossLane) {
                // ---- d = 3, 4, 5: backward transform across the lanes (warp_idft32), nothing goes through shared memory.
                // Output sample n = R k + e, band bin m = m' + 32 j:  x[R k + e] = sum_m' [W_M^(m' e) sum_j Y[m' + 32 j] W_R^(j e)]
                // W_32^(m' k) -- an R-point transform over j in registers, a twiddle, and R 32-point transforms over the
                // lanes; lane l ends with the R consecutive samples R brev5(l) .. R brev5(l) + R - 1.
                const float4 w01 = *reinterpret_cast<const float4*>(tw2), w23 = *reinterpret_cast<const float4*>(tw2 + 2);
                const float2 wsh[4] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w), make_float2(w23.x, w23.y),
                                       make_float2(w23.z, w23.w)};
                c2 o[R];                                  // o[e] = output sample R * brev5(lane) + e
                dit2<R, true>(y);                         // y[e] = sum_j Y[lane + 32 j] W_R^(j e)  (y was placed bit-reversed)
#pragma unroll
                for (int e = 0; e < R; e++) o[e] = warp_idft32(e ? cmul2s(y[e], tw2[3 + e]) : y[e], wsh, lane);
                const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
                const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
                const int idx0 = R * brev5(lane);
                const bool keepA = idx0 >= oa.lo && idx0 < 3 * M / 4;       // lo and 3 mfft / 4 are multiples of R
                const bool keepB = ob.live && idx0 >= ob.lo && idx0 < 3 * M / 4;
                float2* const outA = p.out + (oa.base + idx0);
                float2* const outB = p.out + (ob.base + idx0);
#pragma unroll
                for (int e = 0; e < R; e++) {
                    if constexpr (kMix == 2)
                        o[e] = direct_mix(o[e], direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + idx0 + e, ob.base + idx0 + e),
                                          make_float2(1.0f, 0.0f), p.imag_sign);
                    float ar, br, ai, bi;
                    upk(o[e].re, ar, br);
                    upk(o[e].im, ai, bi);
                    if (keepA) outA[e] = kMix == 2 ? make_float2(ar, ai) : make_float2(ar, flip(ai, sgn));
                    if (keepB) outB[e] = kMix == 2 ? make_float2(br, bi) : make_float2(br, flip(bi, sgn));
                }
            } else {
            // ---- backward mfft-point transform: n = lane + 32*n2, k = k2 + R*k1 ------------------------
            dit2<R, true>(y);                         // pass 1 over n2; its twiddles exp(+2 pi i lane k2 / M) are part of tw2
#pragma unroll
            for (int k2 = 0; k2 < R; k2++) sts2r<R>(S, lane * (R + 1) + k2, y[k2]);
            __syncwarp();
            const int k2p = lane / G;                 // which 32-point DFT this lane works on
            const int h = lane % G;                   // it computes outputs k1 = h + G*q
            c2 t[R];
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) {         // u_j = sum_r x[j + R*r] * exp(+2 pi i h (j + R*r) / 32)
                const c2 x = lds2r<R>(S, n1 * (R + 1) + k2p);
                const float2 w = tw_at(tw2, n1);
                const int j = n1 % R;
                c2& u = t[brev_n(j, RB)];
                if (n1 < R) {
                    u.re = fma2(bc(w.x), x.re, mul2(bc(-w.y), x.im));
                    u.im = fma2(bc(w.x), x.im, mul2(bc(w.y), x.re));
                } else {
                    u.re = fma2(bc(w.x), x.re, fma2(bc(-w.y), x.im, u.re));
                    u.im = fma2(bc(w.x), x.im, fma2(bc(w.y), x.re, u.im));
                }
            }
            __syncwarp();
            dit2<R, true>(t);                         // t[q] = y_out[k2p + R*h + 32*q] for both frames

            // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
            const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
            const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
            const int idx0 = k2p + R * h;
            c2 ph0;
            if constexpr (kMix == 2) ph0 = direct_phasors(nco.nco_phase0, nco.nco_fc, oa.base + idx0, ob.base + idx0);
            // d <= 3: mfft/4 and 3 mfft/4 are whole rows of 32 samples -- the kept range is a compile-time set of rows
            // and lo a warp-uniform choice; d = 4, 5: per-lane comparisons
            constexpr bool kRows = M / 4 >= 32;
            const bool headA = oa.lo == 0, liveB = ob.live, headB = liveB && ob.lo == 0;
            float2* const outA = p.out + (oa.base + idx0);
            float2* const outB = p.out + (ob.base + idx0);
#pragma unroll
            for (int q4 = 0; q4 < (kRows ? 3 * R / 4 : R); q4++) {
                const int idx = idx0 + 32 * q4;
                if constexpr (kMix == 2) t[q4] = direct_mix(t[q4], ph0, __ldg(nco.nco_rot + q4), p.imag_sign);
                float ar, br, ai, bi;
                upk(t[q4].re, ar, br);
                upk(t[q4].im, ai, bi);
                const bool keepA = kRows ? (q4 >= R / 4 || headA) : (idx >= oa.lo && idx < 3 * M / 4);
                const bool keepB = kRows ? (q4 >= R / 4 ? liveB : headB) : (liveB && idx >= ob.lo && idx < 3 * M / 4);
                if (keepA) {
                    float2 va = kMix == 2 ? make_float2(ar, ai) : make_float2(ar, flip(ai, sgn));
                    outA[32 * q4] = va;
                }
                if (keepB) {
                    float2 vb = kMix == 2 ? make_float2(br, bi) : make_float2(br, flip(bi, sgn));
                    outB[32 * q4] = vb;
                }
            }
            }   // R > 2
        }
        q = qn;
        qn = qnn;
    }
}

// The two forms of the parameter block (see PairParams): which one a decimation gets is a measurement --
// B200, 8192 blocks, narrow vs wide block: d = 0 516 vs 500 Gsamples/s, d = 2 790 vs 786, and the other way round at
// d = 1 (594 vs 607), d = 3 (905 vs 921), d = 4 (979 vs 997), d = 5 (998 vs 1013); d = 6 no difference.
template <int D>
constexpr bool kNarrowParams = D == 0 || D == 2;

template <int D, int W, int kMix>
__global__ void __launch_bounds__(W * 32, 1) r2iq_pair_kernel(const FusedParams p) {
    pair_body<D, W, kMix>(p, p);
}
template <int D, int W, int kMix>
__global__ void __launch_bounds__(W * 32, 1) r2iq_pair_kernel(const PairParams p, const PairNco nco) {
    pair_body<D, W, kMix>(p, nco);
}

template <int D, int W, int kMix>
static cudaError_t setup_pair_one() {
    const void* fn;
    if constexpr (kNarrowParams<D>) fn = (const void*)static_cast<void (*)(PairParams, PairNco)>(r2iq_pair_kernel<D, W, kMix>);
    else fn = (const void*)static_cast<void (*)(FusedParams)>(r2iq_pair_kernel<D, W, kMix>);
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, pair_smem_bytes<D, W>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <int D, int W, int kMix>
static cudaError_t launch_pair_one(const FusedParams& fp, int sm_count, cudaStream_t stream) {
    const long long n_pairs = (fp.total_frames + 1) / 2;
    long long grid = sm_count;                 // persistent: one CTA per SM
    const long long want = (n_pairs + W - 1) / W;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(fp.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    if constexpr (kNarrowParams<D>) {
        const PairParams p{fp.in_words, fp.out, fp.total_frames, fp.tw_fwd, fp.gp, fp.gq, fp.tw_inv, fp.tw_aux,
                           fp.tunebin, fp.row_mask, fp.rand_mul, fp.imag_sign, fp.work_counter};
        const PairNco nco{fp.nco_rot, fp.nco_phase0, fp.nco_fc};
        r2iq_pair_kernel<D, W, kMix><<<(unsigned)grid, W * 32, pair_smem_bytes<D, W>(), stream>>>(p, nco);
    } else {
        r2iq_pair_kernel<D, W, kMix><<<(unsigned)grid, W * 32, pair_smem_bytes<D, W>(), stream>>>(fp);
    }
    return cudaGetLastError();
}

// (D, W) combinations that are built
#define R2IQ_PAIR_FOR_ALL(X) \
    X(0, 8) X(1, 12) X(2, 12) X(3, 12) X(4, 12) X(5, 12) X(6, 12)

cudaError_t setup_pair_kernels() {
    cudaError_t e;
#define X(D, W)                                                    \
    if ((e = setup_pair_one<D, W, 0>()) != cudaSuccess) return e; \
    if ((e = setup_pair_one<D, W, 2>()) != cudaSuccess) return e;
    R2IQ_PAIR_FOR_ALL(X)
#undef X
    return cudaSuccess;
}

cudaError_t launch_pair(int d, int warps, const FusedParams& p, int sm_count, cudaStream_t stream) {
    // kMix: 0 = no mixer, 2 = direct-phase mixer in the store epilogue (the reference-bit mixer is a separate pass)
#define X(D, W)           \
    if (d == D && warps == W) \
        return p.nco_rot ? launch_pair_one<D, W, 2>(p, sm_count, stream) : launch_pair_one<D, W, 0>(p, sm_count, stream);
    R2IQ_PAIR_FOR_ALL(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace r2iq
