// r2iq_tok.cu -- the dual-frame kernels (two frames per warp on packed fp32x2 lanes, TMA-staged input) with an
// FMA TOKEN per scheduler: the default kernels for d = 0..5, FFTN 2048.
//
// Why.  A frame pair alternates between butterfly phases (a radix-32 register pass: ~400-770 packed FP32
// instructions, nothing else) and exchange phases (shared-memory transposition, conversion, stores: LSU and integer
// work, no FP32).  One warp in a butterfly phase saturates its scheduler's FMA pipe on its own (one FFMA2 per two
// cycles), one CTA in an exchange phase saturates the SM's shared-memory pipe.  The lock-step kernels of round 1
// (r2iq_dual_d0.cu ...) ran two or three identical CTAs per SM that start together and, because both resources are
// shared fairly, STAY in phase: both in butterflies (each at half speed), then both in exchanges -- the two halves
// of the machine took turns instead of overlapping (FMA pipe 54 %, L1/shared pipe 64 %, and a third CTA per SM
// bought nothing).  Here the warps that share a scheduler (warp w, w+4[, w+8] of one 8- or 12-warp CTA) pass a token
// around a ring through named barriers (producer/consumer bar.arrive / bar.sync, 64 threads each): a warp holds the
// token exactly while it runs a butterfly phase, so at any time one warp per scheduler issues FFMA2 at full rate and
// the others convert, transpose, load tables and store in the free issue slots.
//
// Work comes from the device-wide counter (work_dispenser.cuh idea) one RING at a time, so that the warps of a ring
// run the same number of trips and the token count balances.
// Reference lines replaced: Core/fft_mt_r2iq_impl.hpp:40-54,84-138 (see r2iq_kernels.cu).
#include "dual_frame.cuh"
#include "r2iq_kernels.cuh"

namespace r2iq {

using namespace dualf;

namespace {

// per-warp shared memory: transposition tile (two 8-byte planes) [+ two input frames] + one mbarrier
//   kWps = 2 (8 warps, 255 registers): input frames have their own buffers, the copy of the next pair is issued
//            as soon as this pair has been converted;
//   kWps = 3 (12 warps, 168 registers): the copy lands in the tile itself once the tile has been read for the last
//            time in a trip.
template <int kWps>
__host__ __device__ constexpr int tok_warp_bytes() {
    return kTile * 16 + (kWps == 2 ? 2 * kFrameBytes : 0) + 16;
}
constexpr int kLandOffDn = kTile * 8 - kFrameBytes;   // d >= 1, kWps = 3: a frame at the tail of each 8448-byte plane

__device__ __forceinline__ void tok_wait(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void tok_pass(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }

}  // namespace

// kVar bit 0: token on (off = the same free-running kernel without it: the control of the A/B measurement);
//      bit 1: the token is also held over the mixed segments (split + filter at d = 0, band cut and backward
//             transform at d >= 1) instead of only over the pure radix-32 passes;
//      bit 5 (kVarMix): fine-tune NCO in the store epilogue.
template <int D, int kWps, int kVar>
__global__ void __launch_bounds__(kWps * 128, 1) r2iq_tok_kernel(const FusedParams p) {
    constexpr bool kTok = (kVar & 1) != 0;
    constexpr bool kWide = (kVar & 2) != 0;
    constexpr bool kMix = (kVar & kVarMix) != 0;
    constexpr bool kLandInTile = kWps == 3;
    constexpr int M = kHalfFft >> D;
    constexpr int R = 32 >> D;            // band bins per lane (d >= 1)
    constexpr int G = 32 / (R ? R : 1);   // lanes sharing one 32-point DFT in pass 2 of the backward transform
    constexpr int RB = ilog2(R ? R : 1);
    static_assert(D == 0 || !kLandInTile || 32 * (R + 1) * 8 <= kLandOffDn, "backward transposition overlaps the landing zone");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ long long s_q[12][2];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int sched = warp & 3;           // warps w, w+4, w+8 share a scheduler: one ring each
    const int pos = warp >> 2;            // position in the ring
    const int tok_mine = 1 + pos * 4 + sched;
    const int tok_next = 1 + ((pos + 1) % kWps) * 4 + sched;
    // without the token (control variant) every warp is a ring of its own
    constexpr int kRing = kTok ? kWps : 1;
    const int rpos = kTok ? pos : 0;
    const int slot = kTok ? sched : warp;
    constexpr int kWarpBytes = tok_warp_bytes<kWps>();
    unsigned char* wbase = smem_raw + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    constexpr int kLand = kLandInTile ? (D == 0 ? 0 : kLandOffDn) : kTile * 16;
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + kLand);
    const uint32_t* inB = kLandInTile ? reinterpret_cast<const uint32_t*>(wbase + kTile * 8 + kLand) : inA + kFrameBytes / 4;
    const uint32_t mbar = smem_u32(wbase + kWarpBytes - 16);
    const uint32_t dstA = smem_u32(inA), dstB = smem_u32(inB);

    const long long n_pairs = (p.total_frames + 1) / 2;

    if (lane == 0) {
        mbar_init(mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        *reinterpret_cast<float2*>(wbase + kWarpBytes - 8) = make_float2(1.0f, 1.0f);   // see `one` below
    }
    auto issue = [&](long long pair) {
        const long long fa = 2 * pair;
        const long long fb = (fa + 1 < p.total_frames) ? fa + 1 : fa;   // odd tail: B repeats A, never stored
        mbar_expect_tx(mbar, 2 * kFrameBytes);
        tma_load_1d(dstA, frame_src(p, fa), kFrameBytes, mbar);
        tma_load_1d(dstB, frame_src(p, fb), kFrameBytes, mbar);
    };

    // ---- ring-level work dispenser: the ring leader (position 0, lane 0) grabs kWps consecutive pairs per trip and
    // keeps one grab in flight; the members learn the trip after next through shared memory, ordered by the block
    // barrier below (prologue) and by the token itself afterwards (the leader writes before it passes the token on).
    const bool leader = rpos == 0 && lane == 0;
    unsigned long long pend = 0;
    if (leader) {
        s_q[slot][0] = (long long)atomicAdd(p.work_counter, (unsigned long long)kRing);
        s_q[slot][1] = (long long)atomicAdd(p.work_counter, (unsigned long long)kRing);
        pend = atomicAdd(p.work_counter, (unsigned long long)kRing);
    }
    __syncthreads();
    long long base = s_q[slot][0], nbase = s_q[slot][1];
    __syncthreads();
    long long q = base + rpos;
    if (lane == 0 && q < n_pairs) issue(q);
    uint32_t phase = 0;
    bool have_token = pos == 0;           // the leader starts with it
    long long trip = 0;

    // `one` = packed (1.0f, 1.0f), re-read from shared memory after every acquire: the register passes are written
    // to depend on it (dit2_pinned), which pins them behind the barrier
    const uint32_t one_addr = mbar + 8;
    u64 one = bc(1.0f);
    auto acquire = [&]() {
        if constexpr (kTok) {
            if (!have_token) tok_wait(tok_mine);
            have_token = false;
            asm volatile("ld.volatile.shared.b64 %0, [%1];" : "=l"(one) : "r"(one_addr) : "memory");
        }
    };
    auto release = [&]() {
        if constexpr (kTok) tok_pass(tok_next);
    };
    constexpr int kPhases = (D == 0 && !kWide) ? 4 : 3;   // token holds per trip

    const float2* twf = p.tw_fwd + lane;
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;   // sideband inversion as a sign-bit flip

    while (base < n_pairs) {              // ring-uniform
        long long nnbase;
        if (leader) {
            s_q[slot][trip & 1] = (long long)pend;
            nnbase = (long long)pend;
            pend = atomicAdd(p.work_counter, (unsigned long long)kRing);
        }
        const long long qn = nbase + rpos;
        const bool idle = q >= n_pairs;   // tail: pass the token on, do nothing else
        if (idle) {
#pragma unroll 1
            for (int i = 0; i < kPhases; i++) {
                acquire();
                if (i == 0) nnbase = rpos == 0 ? __shfl_sync(0xffffffffu, nnbase, 0) : *(volatile long long*)&s_q[slot][trip & 1];
                release();
            }
            base = nbase;
            nbase = nnbase;
            q = qn;
            trip++;
            continue;
        }
        c2 v[32];
        mbar_wait(mbar, phase);
        phase ^= 1;
#pragma unroll
        for (int j = 0; j < 32; j++) unpack_words(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
        __syncwarp();   // every lane has consumed its words: the buffers may be overwritten
        if constexpr (!kLandInTile) {
            if (lane == 0 && qn < n_pairs) issue(qn);
        }

        // ---- forward 1024-point transform of z = x[2n] + i x[2n+1] ------------------------------
        acquire();
        nnbase = rpos == 0 ? __shfl_sync(0xffffffffu, nnbase, 0) : *(volatile long long*)&s_q[slot][trip & 1];
        dit2_pinned<32, false>(v, one);
#pragma unroll
        for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], __ldg(twf + 32 * k2));
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        release();          // after the stores: nothing of the pass can sink below the barrier
        __syncwarp();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
        __syncwarp();
        acquire();
        dit2_pinned<32, false>(v, one);           // v[k1] = Z[32*k1 + lane]

        if constexpr (D == 0) {
            // ---- real-FFT split + tune rotation + filter, in place; mirror bins through shared memory ----
            const float2* twi = p.tw_inv + lane;
            const float2* gp = p.gp + lane;
            const float2* gq = p.gq + lane;
            const int n1_inv = (lane - p.tunebin) & 31;
            c2 t[32];
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
            if constexpr (!kWide) release();
            __syncwarp();
            if constexpr (!kWide) acquire();      // the mirror loads below sit behind the barrier: no pin needed
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) {
                const c2 a = v[k1];
                const c2 b = lds2(S, (kHalfFft - (32 * k1 + lane)) & (kHalfFft - 1));
                const float2 P = __ldg(gp + 32 * k1);
                const float2 Q = __ldg(gq + 32 * k1);
                c2 y;
                y.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
                y.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
                t[brev5(k1)] = y;                     // bit-reversed placement for the backward pass
            }
            __syncwarp();
            // ---- backward 1024-point transform; lane l is residue (l - tunebin) mod 32 ---------------
            dit2<32, true>(t);
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) t[k2] = cmul2s(t[k2], __ldg(twi + 32 * k2));
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) sts2(S, n1_inv * 33 + k2, t[k2]);
            release();
            __syncwarp();
#pragma unroll
            for (int n = 0; n < 32; n++) t[brev5(n)] = lds2(S, n * 33 + lane);
            __syncwarp();
            if constexpr (kLandInTile) {
                // the tile has been read for the last time in this trip: the next pair of frames lands in it while
                // the last radix-32 pass and the stores run (generic-proxy reads ordered before the async-proxy writes)
                if (lane == 0 && qn < n_pairs) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(qn);
                }
            }
            acquire();
            dit2_pinned<32, true>(t, one);            // t[k1] = y[lane + 32*k1] for both frames
            release();

            // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
            const FrameOut2 oa = frame_out2<0>(2 * q, p.total_frames);
            const FrameOut2 ob = frame_out2<0>(2 * q + 1, p.total_frames);
#pragma unroll
            for (int k1 = 0; k1 < 24; k1++) {         // idx >= 768 is always discarded
                const int idx = lane + 32 * k1;
                float ar, br, ai, bi;
                upk(t[k1].re, ar, br);
                upk(t[k1].im, ai, bi);
                if (idx >= oa.lo) {
                    float2 va = make_float2(ar, flip(ai, sgn));
                    if constexpr (kMix) va = nco_mix_sample(va, oa.base + idx, p.nco_starts, p.nco_tab);
                    p.out[oa.base + idx] = va;
                }
                if (ob.live && idx >= ob.lo) {
                    float2 vb = make_float2(br, flip(bi, sgn));
                    if constexpr (kMix) vb = nco_mix_sample(vb, ob.base + idx, p.nco_starts, p.nco_tab);
                    p.out[ob.base + idx] = vb;
                }
            }
        } else {
            const float2* twi = p.tw_aux + lane;            // [R][32]  exp(+2 pi i lane k2 / M)
            const float2* tw2 = p.tw_aux + R * 32 + lane;   // [32][32] exp(+2 pi i (lane % G) n1 / 32)
            // ---- spectrum to shared memory in natural order; only the rows the band cut will read -------
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++)
                if ((p.row_mask >> k1) & 1u) sts2(S, 32 * k1 + lane, v[k1]);
            release();
            __syncwarp();
            acquire();                                // the band-cut loads below sit behind the barrier
            // ---- band cut in backward-transform order: m = lane + 32*j, bin k = tunebin + s(m) --------
            c2 y[R];
#pragma unroll
            for (int j = 0; j < R; j++) {
                const int m = lane + 32 * j;
                const int s = m < M / 2 ? m : m - M;                     // upper half of the band is below the tune bin
                const int kk = (p.tunebin + s) & (kHalfFft - 1);         // wrapped bins meet P = Q = 0: zero fill
                const c2 a = lds2(S, kk);
                const c2 b = lds2(S, (kHalfFft - kk) & (kHalfFft - 1));
                const float2 P = __ldg(p.gp + kk);
                const float2 Q = __ldg(p.gq + kk);
                c2 r;
                r.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
                r.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
                y[brev_n(j, RB)] = r;                                    // bit-reversed placement
            }
            __syncwarp();   // every lane is done with the spectrum: the tile is free for the transposition
            if constexpr (kLandInTile) {
                // ... and its tail for the next pair of frames (generic-proxy reads ordered before the async-proxy writes)
                if (lane == 0 && qn < n_pairs) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(qn);
                }
            }
            // ---- backward mfft-point transform: n = lane + 32*n2, k = k2 + R*k1 ------------------------
            dit2<R, true>(y);                         // pass 1 over n2
#pragma unroll
            for (int k2 = 1; k2 < R; k2++) y[k2] = cmul2s(y[k2], __ldg(twi + 32 * k2));   // exp(+2 pi i lane k2 / M)
#pragma unroll
            for (int k2 = 0; k2 < R; k2++) sts2(S, lane * (R + 1) + k2, y[k2]);
            __syncwarp();
            const int k2p = lane / G;                 // which 32-point DFT this lane works on
            const int h = lane % G;                   // it computes outputs k1 = h + G*q
            c2 t[R];
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) {         // u_j = sum_r x[j + R*r] * exp(+2 pi i h (j + R*r) / 32)
                const c2 x = lds2(S, n1 * (R + 1) + k2p);
                const float2 w = __ldg(tw2 + 32 * n1);
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
            release();

            // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
            const FrameOut2 oa = frame_out2<D>(2 * q, p.total_frames);
            const FrameOut2 ob = frame_out2<D>(2 * q + 1, p.total_frames);
            const int idx0 = k2p + R * h;
#pragma unroll
            for (int q4 = 0; q4 < R; q4++) {
                const int idx = idx0 + 32 * q4;
                float ar, br, ai, bi;
                upk(t[q4].re, ar, br);
                upk(t[q4].im, ai, bi);
                if (idx >= oa.lo && idx < 3 * M / 4) {
                    float2 va = make_float2(ar, flip(ai, sgn));
                    if constexpr (kMix) va = nco_mix_sample(va, oa.base + idx, p.nco_starts, p.nco_tab);
                    p.out[oa.base + idx] = va;
                }
                if (ob.live && idx >= ob.lo && idx < 3 * M / 4) {
                    float2 vb = make_float2(br, flip(bi, sgn));
                    if constexpr (kMix) vb = nco_mix_sample(vb, ob.base + idx, p.nco_starts, p.nco_tab);
                    p.out[ob.base + idx] = vb;
                }
            }
        }
        base = nbase;
        nbase = nnbase;
        q = qn;
        trip++;
    }
    // the token the last member passed on after its last butterfly phase is still pending at the leader: absorb it,
    // so that every bar.arrive has met its bar.sync when the CTA retires
    if constexpr (kTok) {
        if (pos == 0 && !have_token) tok_wait(tok_mine);
    }
}

template <int D, int kWps, int kVar>
static cudaError_t setup_tok_one() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_tok_kernel<D, kWps, kVar>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         4 * kWps * tok_warp_bytes<kWps>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_tok_kernel<D, kWps, kVar>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int D, int kWps, int kVar>
static cudaError_t launch_tok_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    constexpr int smem = 4 * kWps * tok_warp_bytes<kWps>();
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long want = (n_pairs + 4 * kWps - 1) / (4 * kWps);
    long long grid = sm_count;                 // persistent: one CTA per SM
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_tok_kernel<D, kWps, kVar><<<(unsigned)grid, kWps * 128, smem, stream>>>(p);
    return cudaGetLastError();
}

#define R2IQ_TOK_FOR_ALL(X) \
    X(0, 2, 0) X(0, 2, 1) X(0, 2, 3) X(0, 3, 0) X(0, 3, 1) X(0, 3, 3) \
    X(1, 2, 0) X(1, 2, 1) X(1, 2, 3) X(1, 3, 0) X(1, 3, 1) X(1, 3, 3) \
    X(2, 2, 0) X(2, 2, 1) X(2, 2, 3) X(2, 3, 0) X(2, 3, 1) X(2, 3, 3) \
    X(3, 2, 0) X(3, 2, 1) X(3, 2, 3) X(3, 3, 0) X(3, 3, 1) X(3, 3, 3) \
    X(4, 2, 0) X(4, 2, 1) X(4, 2, 3) X(4, 3, 0) X(4, 3, 1) X(4, 3, 3) \
    X(5, 2, 0) X(5, 2, 1) X(5, 2, 3) X(5, 3, 0) X(5, 3, 1) X(5, 3, 3)

cudaError_t setup_tok_kernels() {
    cudaError_t e;
#define X(D, W, V) \
    if ((e = setup_tok_one<D, W, V>()) != cudaSuccess) return e;
    R2IQ_TOK_FOR_ALL(X)
#undef X
    return cudaSuccess;
}

// wps = warps per scheduler (2 or 3), var = kVar bits 0..1
cudaError_t launch_tok(int d, int wps, int var, const FusedParams& p, int sm_count, cudaStream_t stream) {
#define X(D, W, V) \
    if (d == D && wps == W && var == V) return launch_tok_one<D, W, V>(p, sm_count, stream);
    R2IQ_TOK_FOR_ALL(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace r2iq
