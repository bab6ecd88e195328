// r2iq_dual_d0.cu -- the d = 0 (decimation 2) member of the dual-frame kernel family of r2iq_dual.cu,
// kept as its own specialised kernel in its own translation unit: forward AND backward transform
// run packed here, the loop body is ~59 KB of straight-line code, and its throughput follows the
// instruction-cache hit rate (ncu sm__icc_request_hit_rate 92 % for this layout vs 82 % when the
// same code is instantiated from the generic template -- 408 vs 369 Gsamples/s, profiles/README.md).
//
// TWO frames per warp on packed fp32x2 lanes, input frames prefetched into shared memory by TMA
// bulk copies (cp.async.bulk + mbarrier).
//
// Same mathematics and the same warp-per-frame decomposition as r2iq_kernels.cu (radix-32 register
// passes, one padded shared-memory transposition per 1024-point transform), but every register
// holds the value of frame A and frame B, so the butterflies issue as FADD2/FFMA2 -- half the issue
// slots per flop -- and the loads/stores/shuffles ride in the freed slots.  What changes because
// of the packing:
//   * the real-FFT mirror bin Z[1024-k] is fetched through shared memory (Z is written once in
//     natural order, 16 bytes per bin for the two frames) instead of by warp shuffle, which lets the
//     split/filter step  Y = P[k]*Z[k] + Q[k]*conj(Z[1024-k])  run IN PLACE in the registers;
//   * shared-memory exchanges move both frames of a bin (two 8-byte planes, re pairs and im pairs);
//   * the next pair of frames is in flight (TMA -> smem) while the current pair is transformed, so
//     8 resident warps per SM (255 registers each) are enough to cover DRAM latency.
// Reference lines replaced: Core/fft_mt_r2iq_impl.hpp:40-54,84-138 (see r2iq_kernels.cu).
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
//   kStage = 2 : TMA bulk copies land IN the transposition tile, which is idle between the last read of the
//                 backward transposition and the next forward one; 16.9 KB per warp, 12 warps per SM
//                 (3 CTAs, 168 registers) with TMA-staged input.
template <int kStage>
__host__ __device__ constexpr int warp_bytes() {
    return kTile * 16 + (kStage == 1 ? 2 * kFrameBytes : 0) + (kStage ? 16 : 0);
}

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
__device__ __forceinline__ float flip(float x, uint32_t sgn) { return __uint_as_float(__float_as_uint(x) ^ sgn); }

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
__device__ __forceinline__ FrameOut2 frame_out_d0(long long f, long long total) {
    constexpr int M = kHalfFft;
    long long b = f / kFramesPerBlock;
    int k = (int)(f - b * kFramesPerBlock);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = b * (long long)(kTransferSamples / 2) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

}  // namespace

// d = 0 (decimation 2, mfft = 1024): forward and backward transforms both run packed.
// kVar (measurement variants, env R2IQ_VARIANT): bit 0 = frame pairs handed out by an atomic counter instead of a
// fixed stride (evens out SM-to-SM speed differences at the tail); bit 1 = no block barriers (warps run free).
template <int kStage, int kVar = 0>
__global__ void __launch_bounds__(kDualThreads, kStage == 1 ? 2 : 3) r2iq_dual_d0_kernel(const FusedParams p) {
    constexpr bool kTma = kStage != 0;
    constexpr bool kDyn = (kVar & 1) != 0;
    constexpr bool kFree = (kVar & 2) != 0;
    constexpr bool kShfl = (kVar & 16) != 0;
    constexpr bool kMix = (kVar & kVarMix) != 0;    // fine-tune NCO in the store epilogue
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int kWarpBytes = warp_bytes<kStage>();
    unsigned char* wbase = smem_raw + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    // kStage 1: two frame buffers behind the tile; kStage 2: frame A at the head of the re plane, frame B at the
    // head of the im plane of the tile itself
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + (kStage == 2 ? 0 : kTile * 16));
    const uint32_t* inB = kStage == 2 ? reinterpret_cast<const uint32_t*>(wbase + kTile * 8) : inA + kFrameBytes / 4;
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
    const float2* twi = p.tw_inv + lane;
    const float2* gp = p.gp + lane;
    const float2* gq = p.gq + lane;
    const int n1_inv = (lane - p.tunebin) & 31;
    // sideband inversion (copy<true>, Core/fft_mt_r2iq.h:63-81) as a sign-bit flip: same bits as the multiply by
    // -1, on the integer pipe, and no constant-bank load in the store epilogue
    const uint32_t sgn = __float_as_uint(p.imag_sign) & 0x80000000u;

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
        if constexpr (kVar & 8) {
            twiddle_powers32(v, twf);
        } else {
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], __ldg(twf + 32 * k2));
        }
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        PHASE_SYNC();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
        __syncwarp();
        dit2<32, false>(v);                       // v[k1] = Z[32*k1 + lane]

        // ---- real-FFT split + tune rotation + filter, in place ---------------------------------
        c2 t[32];
        if constexpr (kShfl) {
            // Variant bit 4: the mirror bin Z[1024-k] comes by warp shuffle instead of a shared-memory round trip.
            // For k = 32*k1 + lane it sits in lane (32 - lane) & 31, register 31 - k1 (lane 0: register (32 - k1) & 31,
            // its own), so each lane SENDS v[31-k1] (lane 0: v[(32-k1)&31]): 4 SHFL.32 per bin pair (one wavefront each)
            // where the staged exchange costs 2 STS.64 + 2 LDS.64 (two wavefronts each).  Rows are walked from both
            // ends (k1 = 0, 31, 1, 30, ...) so that v[] dies as fast as t[] fills.
            const int src = (32 - lane) & 31;
            const bool l0 = lane == 0;
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const int k1 = (i & 1) ? 31 - (i >> 1) : (i >> 1);
                const c2 a = v[k1];
                c2 snd;
                snd.re = l0 ? v[(32 - k1) & 31].re : v[31 - k1].re;
                snd.im = l0 ? v[(32 - k1) & 31].im : v[31 - k1].im;
                c2 b;
                b.re = __shfl_sync(0xffffffffu, snd.re, src);
                b.im = __shfl_sync(0xffffffffu, snd.im, src);
                const float2 P = __ldg(gp + 32 * k1);
                const float2 Q = __ldg(gq + 32 * k1);
                c2 y;
                y.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
                y.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
                t[brev5(k1)] = y;
            }
            PHASE_SYNC();   // pacing only (instruction-cache sharing between the warps of the CTA)
        } else {
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++)
            sts2(S, 32 * k1 + lane, v[k1]);
        PHASE_SYNC();
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
        }

        // ---- backward 1024-point transform; lane l is residue (l - tunebin) mod 32 ---------------
        dit2<32, true>(t);
        if constexpr (kVar & 8) {
            twiddle_powers32(t, twi);
        } else {
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) t[k2] = cmul2s(t[k2], __ldg(twi + 32 * k2));
        }
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, n1_inv * 33 + k2, t[k2]);
        PHASE_SYNC();
#pragma unroll
        for (int n = 0; n < 32; n++) t[brev5(n)] = lds2(S, n * 33 + lane);
        __syncwarp();
        if constexpr (kStage == 2) {
            // the tile has been read for the last time in this trip: the next pair of frames lands in it while the
            // last radix-32 pass and the stores run (generic-proxy reads ordered before the async-proxy writes)
            if (lane == 0 && qn < n_pairs) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue(qn);
            }
        }
        dit2<32, true>(t);                        // t[k1] = y[lane + 32*k1] for both frames

        // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
        const FrameOut2 oa = frame_out_d0(2 * qq, p.total_frames);
        const FrameOut2 ob = frame_out_d0(2 * qq + 1, p.total_frames);
#pragma unroll
        for (int k1 = 0; k1 < 24; k1++) {         // idx >= 768 is always discarded
            const int idx = lane + 32 * k1;
            float ar, br, ai, bi;
            upk(t[k1].re, ar, br);
            upk(t[k1].im, ai, bi);
            if (!idle && idx >= oa.lo) {
                float2 va = make_float2(ar, flip(ai, sgn));
                if constexpr (kMix) va = nco_mix_sample(va, oa.base + idx, p.nco_starts, p.nco_tab);
                p.out[oa.base + idx] = va;
            }
            if (!idle && ob.live && idx >= ob.lo) {
                float2 vb = make_float2(br, flip(bi, sgn));
                if constexpr (kMix) vb = nco_mix_sample(vb, ob.base + idx, p.nco_starts, p.nco_tab);
                p.out[ob.base + idx] = vb;
            }
        }
        work.advance();
        q = qn;
    }
}

template <int kStage, int kVar>
static cudaError_t setup_dual_one() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_dual_d0_kernel<kStage, kVar>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDualWarps * warp_bytes<kStage>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_dual_d0_kernel<kStage, kVar>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

cudaError_t setup_dual_d0_kernels() {
    cudaError_t e;
    if ((e = setup_dual_one<1, 0>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 1>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 3>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 8>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 9>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 17>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 1 | kVarMix>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<1, 25>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<2, 0>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<2, 1>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<2, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<2, 3>()) != cudaSuccess) return e;
    return setup_dual_one<0, 0>();
}

template <int kStage, int kVar>
static cudaError_t launch_dual_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    constexpr int smem = kDualWarps * warp_bytes<kStage>();
    int ctas_per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, r2iq_dual_d0_kernel<kStage, kVar>, kDualThreads, smem);
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
    r2iq_dual_d0_kernel<kStage, kVar><<<(unsigned)grid, kDualThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

template <int kStage>
static cudaError_t launch_dual_var(const FusedParams& p, int sm_count, cudaStream_t stream) {
    switch (kernel_variant() & 3) {
        case 1: return launch_dual_one<kStage, 1>(p, sm_count, stream);
        case 2: return launch_dual_one<kStage, 2>(p, sm_count, stream);
        case 3: return launch_dual_one<kStage, 3>(p, sm_count, stream);
        default: return launch_dual_one<kStage, 0>(p, sm_count, stream);
    }
}

cudaError_t launch_dual_d0(const FusedParams& p, int sm_count, cudaStream_t stream, bool tma) {
    if (!tma) return launch_dual_one<0, 0>(p, sm_count, stream);
    if (p.nco_starts) return launch_dual_one<1, 1 | kVarMix>(p, sm_count, stream);   // default variant + fused mixer
    switch (kernel_variant()) {   // bit 3: inter-pass twiddles from the packed power recurrence
        case 8: return launch_dual_one<1, 8>(p, sm_count, stream);
        case 9: return launch_dual_one<1, 9>(p, sm_count, stream);
        case 17: return launch_dual_one<1, 17>(p, sm_count, stream);   // bit 4: mirror bins by warp shuffle
        case 25: return launch_dual_one<1, 25>(p, sm_count, stream);
        default: break;
    }
    if (kernel_variant() & 4) return launch_dual_var<2>(p, sm_count, stream);   // bit 2: input lands in the tile, 12 warps per SM
    return launch_dual_var<1>(p, sm_count, stream);
}

}  // namespace r2iq
