// r2iq_dual_dn.cu -- dual-frame kernels for d = 2..5 (mfft = 256..32), generalising r2iq_dual_d1.cu.
//
// Forward transform as in the d = 0 / d = 1 kernels (two frames per warp on packed fp32x2 lanes, input
// staged by TMA).  With R = mfft/32 = 32 >> d band bins per lane:
//   * band cut from the natural-order spectrum in shared memory, in backward-transform order
//     (m = lane + 32*j, j < R), filtered in registers -- no staging buffer, no second launch phase;
//   * backward transform of size mfft = 32 * R on all 32 lanes, both frames packed:
//       n = lane + 32*n2, k = k2 + R*k1.   Pass 1: radix-R register DFT over n2, twiddle w_M^(lane*k2).
//       Pass 2 needs R 32-point DFTs from 32 lanes, G = 32/R lanes each: lane (k2, h) computes the
//       outputs k1 = h + G*q, q < R:  u_j = sum_r x[j + R*r] * w_32^(h*(j + R*r)), a 32-term
//       multiply-accumulate with per-lane weights from a 32x32 table, then a radix-R register DFT
//       over j.  Output index k = k2 + R*h + 32*q: every store instruction covers 32 consecutive samples.
// d = 6 (mfft = 16 < 32 lanes) stays with the first-generation kernel.
#include <cstdlib>

#include "dft_regs2.cuh"
#include "r2iq_kernels.cuh"
#include "work_dispenser.cuh"

namespace r2iq {

namespace {

constexpr int kDualWarps = 4;                 // warps per CTA
constexpr int kDualThreads = kDualWarps * 32;
// Input staging is a template switch so both variants stay measurable (profiles/README.md):
//   kTma = true : next pair of frames prefetched global->shared by TMA bulk copies; 25.1 KB of shared
//                 memory per warp, so 8 warps per SM (2 CTAs, 255 registers).
//   kTma = false: frames read straight into registers (32 coalesced 128-byte requests per frame);
//                 16.9 KB per warp, so 12 warps per SM (3 CTAs, 168 registers) hide the load latency.
constexpr int kTile = 32 * 33;                // elements per plane: 32x32 tile, pitch 33 (conflict-free for 8 B accesses)
constexpr int kFrameBytes = kFftN * 2;        // 4096 B of int16 per frame
// per-warp shared memory: transposition tile (two 8-byte planes) [+ two input frames + one mbarrier]
//   kStage = 2 : TMA bulk copies land IN the transposition tile, in the part the (small) backward transposition
//                 does not use, as soon as the band has been cut; 16.9 KB per warp, 12 warps per SM.
template <int kStage>
__host__ __device__ constexpr int warp_bytes() {
    return kTile * 16 + (kStage == 1 ? 2 * kFrameBytes : 0) + (kStage ? 16 : 0);
}
constexpr int kLandOff = kTile * 8 - kFrameBytes;   // 4352: a frame at the tail of each 8448-byte plane

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

__device__ __forceinline__ const uint32_t* frame_src(const FusedParams& p, long long f) {
    const long long b = f / kFramesPerBlock;
    const int k = (int)(f - b * kFramesPerBlock);
    return p.in_words + b * (kTransferSamples / 2) + k * (kHop / 2);
}

// Phase boundaries double as CTA barriers (a superset of the __syncwarp the data hazard needs): they
// keep the warps of a CTA within one phase of each other in the 50 KB unrolled loop body, so that
// instruction-cache lines fetched by one warp are still resident for the others.
// Variant bit 1 (kFree) replaces them by warp-level fences: the warps of a CTA then run unsynchronised.
#define PHASE_SYNC()                           \
    do {                                       \
        if constexpr (kFree) __syncwarp();     \
        else __syncthreads();                  \
    } while (0)

__device__ __forceinline__ int brev5(int v) { return (int)(__brev((unsigned)v) >> 27); }
__device__ __forceinline__ int brev_n(int v, int bits) { return bits == 0 ? 0 : (int)(__brev((unsigned)v) >> (32 - bits)); }

// Shared-memory exchanges use two 64-bit planes (re pairs, im pairs) rather than one 128-bit element:
// a v2.b64 access needs re and im in four consecutive registers, and ptxas paid ~400 MOVs per
// iteration to arrange that; plane accesses have no such constraint and stay conflict-free at pitch 33.
// (Stored as float2 halves: a plain 64-bit store of the packed register makes ptxas copy the pair into a
// scratch pair first -- two MOVs per STS.64 and a serialising dependency; float2 stores use it in place.)
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

__device__ __forceinline__ void unpack_words(uint32_t wa, uint32_t wb, uint32_t rand_mul, c2& v) {
    wa ^= (wa & 0x00010001u) * rand_mul;    // de-randomise both int16 halves (Core/fft_mt_r2iq.h:41-44)
    wb ^= (wb & 0x00010001u) * rand_mul;
    v.re = pk((float)(short)(wa & 0xFFFFu), (float)(short)(wb & 0xFFFFu));
    v.im = pk((float)((int)wa >> 16), (float)((int)wb >> 16));
}

struct FrameOut2 {
    long long base;
    int lo;
    bool live;
};
template <int D>
__device__ __forceinline__ FrameOut2 frame_out_dn(long long f, long long total) {
    constexpr int M = kHalfFft >> D;
    long long b = f / kFramesPerBlock;
    int k = (int)(f - b * kFramesPerBlock);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = b * (long long)((kTransferSamples / 2) >> D) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

}  // namespace

// forward and backward transforms both run packed; D = 2..5
// kVar (measurement variants, env R2IQ_VARIANT): bit 0 = frame pairs handed out by an atomic counter instead of a
// fixed stride (evens out SM-to-SM speed differences at the tail); bit 1 = no block barriers (warps run free).
template <int D, int kStage, int kVar = 0>
__global__ void __launch_bounds__(kDualThreads, kStage == 2 ? 3 : 2) r2iq_dual_dn_kernel(const FusedParams p) {
    constexpr bool kTma = kStage != 0;
    static_assert(kStage != 2 || 32 * ((32 >> D) + 1) * 8 <= kLandOff, "backward transposition overlaps the landing zone");
    constexpr bool kDyn = (kVar & 1) != 0;
    constexpr bool kFree = (kVar & 2) != 0;
    constexpr bool kMix = (kVar & kVarMix) != 0;    // fine-tune NCO in the store epilogue
    constexpr int M = kHalfFft >> D;
    constexpr int R = 32 >> D;            // band bins per lane = size of the register DFTs
    constexpr int G = 32 / R;             // lanes sharing one 32-point DFT in pass 2
    constexpr int RB = ilog2(R);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int kWarpBytes = warp_bytes<kStage>();
    unsigned char* wbase = smem_raw + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + (kStage == 2 ? kLandOff : kTile * 16));
    const uint32_t* inB = kStage == 2 ? reinterpret_cast<const uint32_t*>(wbase + kTile * 8 + kLandOff) : inA + kFrameBytes / 4;
    const uint32_t mbar = smem_u32(wbase + kTile * 16 + (kStage == 1 ? 2 * kFrameBytes : 0));
    const uint32_t dstA = smem_u32(inA), dstB = smem_u32(inB);

    const long long n_pairs = (p.total_frames + 1) / 2;

    if constexpr (kTma) {
        if (lane == 0) {
            mbar_init(mbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
    auto issue = [&](long long pair) {
        const long long fa = 2 * pair;
        const long long fb = (fa + 1 < p.total_frames) ? fa + 1 : fa;   // odd tail: B repeats A, never stored
        mbar_expect_tx(mbar, 2 * kFrameBytes);
        tma_load_1d(dstA, frame_src(p, fa), kFrameBytes, mbar);
        tma_load_1d(dstB, frame_src(p, fb), kFrameBytes, mbar);
    };
    // frame pairs from the device-wide counter (kDyn) or by fixed stride; see work_dispenser.cuh
    __shared__ long long s_base[3];
    PairDispenser<kDyn, kFree, kDualWarps> work;
    work.init(p.work_counter, s_base, lane, warp);
    long long q = work.cur(warp);
    if constexpr (kTma) {
        if (lane == 0 && q < n_pairs) issue(q);
    }
    uint32_t phase = 0;

    const float2* twf = p.tw_fwd + lane;
    const float2* twi = p.tw_aux + lane;            // [R][32]  exp(+2 pi i lane k2 / M)
    const float2* tw2 = p.tw_aux + R * 32 + lane;   // [32][32] exp(+2 pi i (lane % G) n1 / 32)

    // The loop bound is CTA-uniform and every trip starts with a block barrier: the four warps then
    // walk the (56 KB, fully unrolled) loop body roughly in step and share instruction-cache lines.
    while (work.more(n_pairs)) {
        work.top(p.work_counter, s_base);
        const long long qn = work.next(warp);
        const bool idle = q >= n_pairs;           // tail: keep barrier participation uniform, store nothing
        const long long qq = idle ? n_pairs - 1 : q;
        c2 v[32];
        if constexpr (kTma) {
            if (!idle) {                          // idle warps issued no copy; they chew on stale buffer contents
                mbar_wait(mbar, phase);
                phase ^= 1;
            }
#pragma unroll
            for (int j = 0; j < 32; j++) unpack_words(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
            __syncwarp();   // every lane has consumed its words: the buffers may be overwritten
            if constexpr (kStage == 1) {
                if (lane == 0 && qn < n_pairs) issue(qn);   // idle warps never issue: nobody would consume the copy
            }
        } else {
            const long long fa = 2 * qq;
            const long long fb = (fa + 1 < p.total_frames) ? fa + 1 : fa;   // odd tail: B repeats A, never stored
            const uint32_t* ga = frame_src(p, fa) + lane;
            const uint32_t* gb = frame_src(p, fb) + lane;
            uint32_t wa[32], wb[32];
#pragma unroll
            for (int j = 0; j < 32; j++) {
                wa[j] = __ldg(ga + 32 * j);
                wb[j] = __ldg(gb + 32 * j);
            }
#pragma unroll
            for (int j = 0; j < 32; j++) unpack_words(wa[j], wb[j], p.rand_mul, v[brev5(j)]);
        }

        // ---- forward 1024-point transform of z = x[2n] + i x[2n+1] ------------------------------
        dit2<32, false>(v);
#pragma unroll
        for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], __ldg(twf + 32 * k2));
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        PHASE_SYNC();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
        __syncwarp();
        dit2<32, false>(v);                       // v[k1] = Z[32*k1 + lane]

        // ---- spectrum to shared memory in natural order; only the rows the band cut will read -------
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++)
            if ((p.row_mask >> k1) & 1u) sts2(S, 32 * k1 + lane, v[k1]);
        PHASE_SYNC();
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
        PHASE_SYNC();   // every lane is done with the spectrum: the tile is free for the transposition
        if constexpr (kStage == 2) {
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
        PHASE_SYNC();
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

        // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
        const FrameOut2 oa = frame_out_dn<D>(2 * qq, p.total_frames);
        const FrameOut2 ob = frame_out_dn<D>(2 * qq + 1, p.total_frames);
        const int idx0 = k2p + R * h;
#pragma unroll
        for (int q4 = 0; q4 < R; q4++) {
            const int idx = idx0 + 32 * q4;
            float ar, br, ai, bi;
            upk(t[q4].re, ar, br);
            upk(t[q4].im, ai, bi);
            if (!idle && idx >= oa.lo && idx < 3 * M / 4) {
                float2 va = make_float2(ar, ai * p.imag_sign);
                if constexpr (kMix) va = nco_mix_sample(va, oa.base + idx, p.nco_starts, p.nco_tab);
                p.out[oa.base + idx] = va;
            }
            if (!idle && ob.live && idx >= ob.lo && idx < 3 * M / 4) {
                float2 vb = make_float2(br, bi * p.imag_sign);
                if constexpr (kMix) vb = nco_mix_sample(vb, ob.base + idx, p.nco_starts, p.nco_tab);
                p.out[ob.base + idx] = vb;
            }
        }
        work.advance();
        q = qn;
    }
}

template <int D, int kVar, int kStage = 1>
static cudaError_t setup_dual_one() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_dual_dn_kernel<D, kStage, kVar>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDualWarps * warp_bytes<kStage>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_dual_dn_kernel<D, kStage, kVar>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int D>
static cudaError_t setup_dual_all() {
    cudaError_t e;
    if ((e = setup_dual_one<D, 0>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<D, 1>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<D, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<D, 1, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<D, 3, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<D, 3 | kVarMix, 2>()) != cudaSuccess) return e;
    return setup_dual_one<D, 3>();
}

cudaError_t setup_dual_dn_kernels() {
    cudaError_t e;
    if ((e = setup_dual_all<1>()) != cudaSuccess) return e;
    if ((e = setup_dual_all<2>()) != cudaSuccess) return e;
    if ((e = setup_dual_all<3>()) != cudaSuccess) return e;
    if ((e = setup_dual_all<4>()) != cudaSuccess) return e;
    return setup_dual_all<5>();
}

template <int D, int kVar, int kStage = 1>
static cudaError_t launch_dual_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    constexpr int smem = kDualWarps * warp_bytes<kStage>();
    int ctas_per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, r2iq_dual_dn_kernel<D, kStage, kVar>, kDualThreads, smem);
    if (e != cudaSuccess) return e;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    if (const char* lim = getenv("R2IQ_CTAS_PER_SM")) {   // occupancy experiments only
        const int v = atoi(lim);
        if (v >= 1 && v < ctas_per_sm) ctas_per_sm = v;
    }
    const long long n_pairs = (p.total_frames + 1) / 2;
    long long want = (n_pairs + kDualWarps - 1) / kDualWarps;
    long long grid = (long long)sm_count * ctas_per_sm;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    if constexpr (kVar & 1) {
        e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    r2iq_dual_dn_kernel<D, kStage, kVar><<<(unsigned)grid, kDualThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

template <int D>
static cudaError_t launch_dual_var(const FusedParams& p, int sm_count, cudaStream_t stream) {
    // default 7 (measured on B200, 8192 blocks): 12 free-running warps per SM + dynamic distribution give
    // 760 / 869 / 929 / 968 Gsamples/s at d = 2..5, the 8-warp lock-step form (variant 1) 710 / 791 / 858 / 878
    if (p.nco_starts) return launch_dual_one<D, 3 | kVarMix, 2>(p, sm_count, stream);   // default variant + fused mixer
    switch (kernel_variant(7)) {
        case 0: return launch_dual_one<D, 0>(p, sm_count, stream);
        case 1: return launch_dual_one<D, 1>(p, sm_count, stream);
        case 2: return launch_dual_one<D, 2>(p, sm_count, stream);
        case 3: return launch_dual_one<D, 3>(p, sm_count, stream);
        case 5: return launch_dual_one<D, 1, 2>(p, sm_count, stream);
        default: return launch_dual_one<D, 3, 2>(p, sm_count, stream);
    }
}

cudaError_t launch_dual_dn(int d, const FusedParams& p, int sm_count, cudaStream_t stream) {
    switch (d) {
        case 1: return launch_dual_var<1>(p, sm_count, stream);
        case 2: return launch_dual_var<2>(p, sm_count, stream);
        case 3: return launch_dual_var<3>(p, sm_count, stream);
        case 4: return launch_dual_var<4>(p, sm_count, stream);
        case 5: return launch_dual_var<5>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
