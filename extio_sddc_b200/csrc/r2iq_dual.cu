// r2iq_dual.cu -- FIRST dual-frame family (kept for A/B measurements, R2IQ_KERNEL_GEN=2/3; the default
// kernels are r2iq_dual_d0.cu, r2iq_dual_d1.cu and r2iq_dual_dn.cu, which also pack the backward transform):
// TWO frames per warp on packed fp32x2 lanes, input frames optionally prefetched into shared memory by TMA
// bulk copies (cp.async.bulk + mbarrier); for d >= 1 the bands are staged and transformed back by the
// scalar inverse_batch of the first-generation kernels.
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
#include "r2iq_common.cuh"

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
// per-warp shared memory: transposition tile (two 8-byte planes; doubles as the float2 tile of the
// scalar backward transform for d >= 1) [+ staged bands, d >= 1] [+ two input frames + one mbarrier]
template <int D, bool kTma>
__host__ __device__ constexpr int warp_bytes() {
    return kTile * 16 + (D > 0 ? kStage * 8 : 0) + (kTma ? 2 * kFrameBytes + 16 : 0);
}
// resident CTAs per SM the register budget is sized for: d = 0 TMA 2 (255 regs), d = 0 direct 3 (168);
// d >= 1 keep 2 (shared memory: 4 warps x 25.6 KB per CTA)
template <int D, bool kTma>
__host__ __device__ constexpr int min_ctas() {
    return kDualWarps == 8 ? 1 : ((D == 0 && !kTma) ? 3 : 2) * (4 / kDualWarps);
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
#define PHASE_SYNC() __syncthreads()

__device__ __forceinline__ int brev5(int v) { return (int)(__brev((unsigned)v) >> 27); }

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

// Packed split/filter of the bin held in register K1 (k = 32*K1 + lane), both frames at once:
//   Y = P[k]*Z[k] + Q[k]*conj(Z[1024-k]),  P = G + GC, Q = G - GC  (see r2iq_kernels.cu split_filter)
__device__ __forceinline__ c2 split_filter2(const c2 a, const c2 b, const float2 P, const float2 Q) {
    c2 y;
    y.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
    y.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
    return y;
}

template <int D, int K1>
__device__ __forceinline__ void band_row2(const FusedParams& p, const c2 (&v)[32], int lane, const u64* S,
                                          const float2* gp, const float2* gq, float2* stA, float2* stB) {
    constexpr int M = Geo<D>::M;
    if (32 * K1 + 31 >= p.band_lo && 32 * K1 < p.band_hi) {      // warp-uniform
        const int k = 32 * K1 + lane;
        const c2 b = lds2(S, (kHalfFft - k) & (kHalfFft - 1));
        const c2 y = split_filter2(v[K1], b, __ldg(gp + 32 * K1), __ldg(gq + 32 * K1));
        if (k >= p.band_lo && k < p.band_hi) {
            const int m = (k - p.tunebin) & (M - 1);
            float ar, br, ai, bi;
            upk(y.re, ar, br);
            upk(y.im, ai, bi);
            stA[m] = make_float2(ar, ai);
            stB[m] = make_float2(br, bi);
        }
    }
}
template <int D, int... K1>
__device__ __forceinline__ void band_rows2(const FusedParams& p, const c2 (&v)[32], int lane, const u64* S,
                                           const float2* gp, const float2* gq, float2* stA, float2* stB,
                                           std::integer_sequence<int, K1...>) {
    (band_row2<D, K1>(p, v, lane, S, gp, gq, stA, stB), ...);
}

}  // namespace

// One kernel per decimation index D.  A warp walks frame PAIRS (2i, 2i+1); for D >= 1 it stages the
// filtered bands of 2^D consecutive frames (2^(D-1) pairs) and then runs their backward transforms
// together (inverse_batch, one-frame lanes); for D = 0 forward and backward both run packed.
template <int D, bool kTma>
__global__ void __launch_bounds__(kDualThreads, min_ctas<D, kTma>()) r2iq_dual_kernel(const FusedParams p) {
    using G = Geo<D>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int kWarpBytes = warp_bytes<D, kTma>();
    constexpr int kPairs = (D == 0) ? 1 : (G::ROUNDS / 2);      // frame pairs per unit of work
    unsigned char* wbase = smem_raw + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    float2* stage = reinterpret_cast<float2*>(wbase + kTile * 16);
    unsigned char* in_base = wbase + kTile * 16 + (D > 0 ? kStage * 8 : 0);
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(in_base);
    const uint32_t* inB = inA + kFrameBytes / 4;
    const uint32_t mbar = smem_u32(in_base + 2 * kFrameBytes);
    const uint32_t dstA = smem_u32(inA), dstB = smem_u32(inB);

    const long long last_frame = p.total_frames - 1;
    const long long n_units = (p.total_frames + 2 * kPairs - 1) / (2 * kPairs);
    const long long stride = (long long)gridDim.x * kDualWarps;
    long long unit = (long long)blockIdx.x * kDualWarps + warp;

    if constexpr (D > 0) {
        // band bins outside [0, halfFft) are never written and must read as zero (impl.hpp:91-96)
        for (int i = lane; i < kStage; i += 32) stage[i] = make_float2(0.f, 0.f);
    }
    if constexpr (kTma) {
        if (lane == 0) {
            mbar_init(mbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncwarp();
    // frames past the end of the capture are clamped for loading and never stored
    auto issue = [&](long long fa) {
        const long long a = fa < last_frame ? fa : last_frame;
        const long long b = fa + 1 < last_frame ? fa + 1 : last_frame;
        mbar_expect_tx(mbar, 2 * kFrameBytes);
        tma_load_1d(dstA, frame_src(p, a), kFrameBytes, mbar);
        tma_load_1d(dstB, frame_src(p, b), kFrameBytes, mbar);
    };
    if constexpr (kTma) {
        if (lane == 0 && unit < n_units) issue(unit * (2 * kPairs));
    }
    uint32_t phase = 0;

    const float2* twf = p.tw_fwd + lane;
    const float2* twi = p.tw_inv + lane;
    const float2* gp = p.gp + lane;
    const float2* gq = p.gq + lane;
    const int n1_inv = (lane - p.tunebin) & 31;

    // The unit loop is CTA-uniform and phase boundaries are CTA barriers (see PHASE_SYNC); warps past
    // the end of the work list ("idle") keep participating on clamped frames and store nothing.
    for (long long u0 = (long long)blockIdx.x * kDualWarps; u0 < n_units; u0 += stride, unit += stride) {
        const bool idle = unit >= n_units;
        const long long f0 = (idle ? n_units - 1 : unit) * (2 * kPairs);
#pragma unroll 1
        for (int r = 0; r < kPairs; r++) {
            __syncthreads();
            const long long fa = f0 + 2 * r;
            c2 v[32];
            if constexpr (kTma) {
                if (!idle) {                      // idle warps issued no copy; they chew on stale buffer contents
                    mbar_wait(mbar, phase);
                    phase ^= 1;
                }
#pragma unroll
                for (int j = 0; j < 32; j++)
                    unpack_words(inA[lane + 32 * j], inB[lane + 32 * j], p.rand_mul, v[brev5(j)]);
                __syncwarp();   // every lane has consumed its words: the buffers may be overwritten
                if (lane == 0 && !idle) {
                    if (r + 1 < kPairs) issue(fa + 2);
                    else if (unit + stride < n_units) issue((unit + stride) * (2 * kPairs));
                }
            } else {
                const long long a = fa < last_frame ? fa : last_frame;
                const long long b = fa + 1 < last_frame ? fa + 1 : last_frame;
                const uint32_t* ga = frame_src(p, a) + lane;
                const uint32_t* gb = frame_src(p, b) + lane;
                uint32_t wa[32], wb[32];
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    wa[j] = __ldg(ga + 32 * j);
                    wb[j] = __ldg(gb + 32 * j);
                }
#pragma unroll
                for (int j = 0; j < 32; j++) unpack_words(wa[j], wb[j], p.rand_mul, v[brev5(j)]);
            }

            // ---- forward 1024-point transform of z = x[2n] + i x[2n+1] --------------------------
            dit2<32, false>(v);
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], __ldg(twf + 32 * k2));
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
            PHASE_SYNC();
#pragma unroll
            for (int n1 = 0; n1 < 32; n1++) v[brev5(n1)] = lds2(S, n1 * 33 + lane);
            __syncwarp();
            dit2<32, false>(v);                   // v[k1] = Z[32*k1 + lane]

            // ---- real-FFT split + tune rotation + filter ----------------------------------------
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
            PHASE_SYNC();
            if constexpr (D == 0) {
                c2 t[32];
#pragma unroll
                for (int k1 = 0; k1 < 32; k1++) {
                    const c2 b = lds2(S, (kHalfFft - (32 * k1 + lane)) & (kHalfFft - 1));
                    // in place, bit-reversed placement for the backward pass
                    t[brev5(k1)] = split_filter2(v[k1], b, __ldg(gp + 32 * k1), __ldg(gq + 32 * k1));
                }
                __syncwarp();
                // ---- backward 1024-point transform; lane l is residue (l - tunebin) mod 32 -------
                dit2<32, true>(t);
#pragma unroll
                for (int k2 = 1; k2 < 32; k2++) t[k2] = cmul2s(t[k2], __ldg(twi + 32 * k2));
#pragma unroll
                for (int k2 = 0; k2 < 32; k2++) sts2(S, n1_inv * 33 + k2, t[k2]);
                PHASE_SYNC();
#pragma unroll
                for (int n = 0; n < 32; n++) t[brev5(n)] = lds2(S, n * 33 + lane);
                __syncwarp();
                dit2<32, true>(t);                // t[k1] = y[lane + 32*k1] for both frames
                // ---- overlap-save discard + placement (impl.hpp:117-138) ------------------------
                const FrameOut oa = frame_out<0>(fa);
                const FrameOut ob = frame_out<0>(fa + 1);
                const bool liveA = !idle && fa <= last_frame;
                const bool liveB = !idle && fa + 1 <= last_frame;
#pragma unroll
                for (int k1 = 0; k1 < 24; k1++) {     // idx >= 768 is always discarded
                    const int idx = lane + 32 * k1;
                    float ar, br, ai, bi;
                    upk(t[k1].re, ar, br);
                    upk(t[k1].im, ai, bi);
                    if (liveA && idx >= oa.lo) p.out[oa.base + idx] = make_float2(ar, ai * p.imag_sign);
                    if (liveB && idx >= ob.lo) p.out[ob.base + idx] = make_float2(br, bi * p.imag_sign);
                }
            } else {
                band_rows2<D>(p, v, lane, S, gp, gq, stage + (2 * r) * G::GS, stage + (2 * r + 1) * G::GS,
                              std::make_integer_sequence<int, 32>{});
            }
        }
        if constexpr (D > 0) {
            __syncthreads();
            if (!idle) inverse_batch<D>(p, f0, lane, reinterpret_cast<float2*>(S), stage);
        }
    }
}

template <int D, bool kTma>
static cudaError_t setup_dual_one() {
    if (kDualWarps * warp_bytes<D, kTma>() > 227 * 1024) return cudaSuccess;   // variant not launchable at this CTA size
    cudaError_t e = cudaFuncSetAttribute(r2iq_dual_kernel<D, kTma>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDualWarps * warp_bytes<D, kTma>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_dual_kernel<D, kTma>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int D>
static cudaError_t setup_dual_d() {
    cudaError_t e = setup_dual_one<D, true>();
    if (e != cudaSuccess) return e;
    return setup_dual_one<D, false>();
}

cudaError_t setup_dual_kernels() {
    cudaError_t e;
    if ((e = setup_dual_d<0>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<1>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<2>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<3>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<4>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<5>()) != cudaSuccess) return e;
    if ((e = setup_dual_d<6>()) != cudaSuccess) return e;
    return cudaSuccess;
}

template <int D, bool kTma>
static cudaError_t launch_dual_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    constexpr int smem = kDualWarps * warp_bytes<D, kTma>();
    constexpr int kPairs = (D == 0) ? 1 : (Geo<D>::ROUNDS / 2);
    int ctas_per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, r2iq_dual_kernel<D, kTma>, kDualThreads, smem);
    if (e != cudaSuccess) return e;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    if (const char* lim = getenv("R2IQ_CTAS_PER_SM")) {   // occupancy experiments only
        const int v = atoi(lim);
        if (v >= 1 && v < ctas_per_sm) ctas_per_sm = v;
    }
    const long long n_units = (p.total_frames + 2 * kPairs - 1) / (2 * kPairs);
    long long want = (n_units + kDualWarps - 1) / kDualWarps;
    long long grid = (long long)sm_count * ctas_per_sm;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    r2iq_dual_kernel<D, kTma><<<(unsigned)grid, kDualThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

template <int D>
static cudaError_t launch_dual_d(const FusedParams& p, int sm_count, cudaStream_t stream, bool tma) {
    return tma ? launch_dual_one<D, true>(p, sm_count, stream) : launch_dual_one<D, false>(p, sm_count, stream);
}

cudaError_t launch_dual(int d, const FusedParams& p, int sm_count, cudaStream_t stream, bool tma) {
    switch (d) {
        case 0: return launch_dual_d<0>(p, sm_count, stream, tma);
        case 1: return launch_dual_d<1>(p, sm_count, stream, tma);
        case 2: return launch_dual_d<2>(p, sm_count, stream, tma);
        case 3: return launch_dual_d<3>(p, sm_count, stream, tma);
        case 4: return launch_dual_d<4>(p, sm_count, stream, tma);
        case 5: return launch_dual_d<5>(p, sm_count, stream, tma);
        case 6: return launch_dual_d<6>(p, sm_count, stream, tma);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
