// r2iq_dual_d1.cu -- the d = 1 (decimation 4, mfft = 512) member of the dual-frame kernel family.
//
// Forward transform exactly as r2iq_dual_d0.cu (two frames per warp on packed fp32x2 lanes, TMA-staged
// input).  What is specific to d = 1:
//   * the band is cut straight out of the natural-order spectrum in shared memory IN THE ORDER THE
//     BACKWARD TRANSFORM WANTS IT: lane l fetches bins k = tunebin + s(m), m = l + 32*j, j = 0..15, and
//     their mirrors, so the tune rotation is pure addressing, the P/Q tables (zero outside the band)
//     do the edge zero-fill, and the 16 filtered bins per lane never leave the registers -- no staging;
//   * the 512-point backward transform also runs packed on all 32 lanes: 512 = 16 (registers) x 32
//     (lanes).  Pass 1 is a radix-16 register DFT; pass 2 needs sixteen 32-point DFTs from 32 lanes,
//     so lane (k2, h) computes the outputs k1 = 2q + h of DFT number k2: one redundant radix-2 DIF
//     stage on all 32 inputs (sign and twiddle selected by h) and a radix-16 register DFT.
// The first-generation kernel does the same work with scalar lanes at 342 Gsamples/s.
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
template <bool kTma>
__host__ __device__ constexpr int warp_bytes() {
    return kTile * 16 + (kTma ? 2 * kFrameBytes + 16 : 0);
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
__device__ __forceinline__ FrameOut2 frame_out_d1(long long f, long long total) {
    constexpr int M = kHalfFft / 2;
    long long b = f / kFramesPerBlock;
    int k = (int)(f - b * kFramesPerBlock);
    FrameOut2 o;
    o.live = f < total;
    o.lo = k ? 0 : M / 4;
    o.base = b * (long long)(kTransferSamples / 4) + (k ? (M / 2 + (k - 1) * (3 * M / 4)) : 0) - o.lo;
    return o;
}

// One redundant radix-2 DIF stage of the 32-point DFT number k2p: u_J = (x_J + (-1)^h x_{J+16}) * w^(h*J),
// w = exp(+2 pi i / 32); deposited bit-reversed for the radix-16 DIT that follows.
template <int J>
__device__ __forceinline__ void dif_step(const u64* S, int k2p, int h, float sgn, c2 (&t)[16]) {
    const c2 lo = lds2(S, J * 17 + k2p);
    const c2 hi = lds2(S, (J + 16) * 17 + k2p);
    c2 u;
    u.re = fma2(bc(sgn), hi.re, lo.re);
    u.im = fma2(bc(sgn), hi.im, lo.im);
    if constexpr (J > 0) {
        constexpr float cr = (float)cx_cos2pi(J, 32);
        constexpr float ci = (float)cx_sin2pi(J, 32);
        u = cmul2s(u, make_float2(h ? cr : 1.0f, h ? ci : 0.0f));
    }
    t[bitrev(J, 4)] = u;
}
template <int... J>
__device__ __forceinline__ void dif_stage(const u64* S, int k2p, int h, float sgn, c2 (&t)[16],
                                          std::integer_sequence<int, J...>) {
    (dif_step<J>(S, k2p, h, sgn, t), ...);
}

}  // namespace

// d = 1 (decimation 4, mfft = 512): forward and backward transforms both run packed.
// kVar (env R2IQ_VARIANT): bit 0 = frame pairs handed out by an atomic counter (work_dispenser.cuh), bit 1 = no
// block barriers (warps run free).
template <bool kTma, int kVar = 0>
__global__ void __launch_bounds__(kDualThreads, kTma ? 2 : 3) r2iq_dual_d1_kernel(const FusedParams p) {
    constexpr bool kDyn = (kVar & 1) != 0;
    constexpr bool kFree = (kVar & 2) != 0;
    constexpr bool kMix = (kVar & kVarMix) != 0;    // fine-tune NCO in the store epilogue
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int kWarpBytes = warp_bytes<kTma>();
    unsigned char* wbase = smem_raw + (size_t)warp * kWarpBytes;
    u64* S = reinterpret_cast<u64*>(wbase);
    const uint32_t* inA = reinterpret_cast<const uint32_t*>(wbase + kTile * 16);
    const uint32_t* inB = inA + kFrameBytes / 4;
    const uint32_t mbar = smem_u32(wbase + kTile * 16 + 2 * kFrameBytes);
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
    const float2* twi = p.tw_aux + lane;      // [16][32] exp(+2 pi i lane k2 / 512)

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
            if (lane == 0 && qn < n_pairs) issue(qn);   // idle warps never issue: nobody would consume the copy
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

        // ---- spectrum to shared memory in natural order -------------------------------------------
#pragma unroll
        for (int k1 = 0; k1 < 32; k1++) sts2(S, 32 * k1 + lane, v[k1]);
        PHASE_SYNC();
        // ---- band cut in backward-transform order: m = lane + 32*j, bin k = tunebin + s(m) --------
        c2 y[16];
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const int m = lane + 32 * j;
            const int s = j < 8 ? m : m - 512;                       // upper half of the band is below the tune bin
            const int kk = (p.tunebin + s) & (kHalfFft - 1);         // wrapped bins meet P = Q = 0: zero fill
            const c2 a = lds2(S, kk);
            const c2 b = lds2(S, (kHalfFft - kk) & (kHalfFft - 1));
            const float2 P = __ldg(p.gp + kk);
            const float2 Q = __ldg(p.gq + kk);
            c2 r;
            r.re = fma2(bc(Q.y), b.im, fma2(bc(Q.x), b.re, fma2(bc(-P.y), a.im, mul2(bc(P.x), a.re))));
            r.im = fma2(bc(Q.y), b.re, fma2(bc(-Q.x), b.im, fma2(bc(P.y), a.re, mul2(bc(P.x), a.im))));
            y[(int)(__brev((unsigned)j) >> 28)] = r;                   // bit-reversed placement (4 bits)
        }
        PHASE_SYNC();   // every lane is done with the spectrum: the tile is free for the transposition

        // ---- backward 512-point transform: n = lane + 32*n2, k = k2 + 16*k1 ----------------------
        dit2<16, true>(y);                        // pass 1: y[k2] = sum over n2
#pragma unroll
        for (int k2 = 1; k2 < 16; k2++) y[k2] = cmul2s(y[k2], __ldg(twi + 32 * k2));   // exp(+2 pi i lane k2 / 512)
#pragma unroll
        for (int k2 = 0; k2 < 16; k2++) sts2(S, lane * 17 + k2, y[k2]);
        PHASE_SYNC();
        const int k2p = lane >> 1;                // which 32-point DFT this lane works on
        const int h = lane & 1;                   // outputs k1 = 2q + h
        const float sgn = h ? -1.0f : 1.0f;
        c2 t[16];
        dif_stage(S, k2p, h, sgn, t, std::make_integer_sequence<int, 16>{});
        __syncwarp();
        dit2<16, true>(t);                        // t[q] = y_out[k2p + 16*(2q + h)] for both frames

        // ---- overlap-save discard + placement (impl.hpp:117-138) --------------------------------
        const FrameOut2 oa = frame_out_d1(2 * qq, p.total_frames);
        const FrameOut2 ob = frame_out_d1(2 * qq + 1, p.total_frames);
#pragma unroll
        for (int q4 = 0; q4 < 12; q4++) {         // idx >= 384 is always discarded
            const int idx = k2p + 16 * h + 32 * q4;
            float ar, br, ai, bi;
            upk(t[q4].re, ar, br);
            upk(t[q4].im, ai, bi);
            if (!idle && idx >= oa.lo) {
                float2 va = make_float2(ar, ai * p.imag_sign);
                if constexpr (kMix) va = nco_mix_sample(va, oa.base + idx, p.nco_starts, p.nco_tab);
                p.out[oa.base + idx] = va;
            }
            if (!idle && ob.live && idx >= ob.lo) {
                float2 vb = make_float2(br, bi * p.imag_sign);
                if constexpr (kMix) vb = nco_mix_sample(vb, ob.base + idx, p.nco_starts, p.nco_tab);
                p.out[ob.base + idx] = vb;
            }
        }
        work.advance();
        q = qn;
    }
}

template <bool kTma, int kVar>
static cudaError_t setup_dual_one() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_dual_d1_kernel<kTma, kVar>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDualWarps * warp_bytes<kTma>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_dual_d1_kernel<kTma, kVar>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

cudaError_t setup_dual_d1_kernels() {
    cudaError_t e;
    if ((e = setup_dual_one<true, 0>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<true, 1>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<true, 2>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<true, 3>()) != cudaSuccess) return e;
    if ((e = setup_dual_one<true, 1 | kVarMix>()) != cudaSuccess) return e;
    return setup_dual_one<false, 0>();
}

template <bool kTma, int kVar>
static cudaError_t launch_dual_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    constexpr int smem = kDualWarps * warp_bytes<kTma>();
    int ctas_per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, r2iq_dual_d1_kernel<kTma, kVar>, kDualThreads, smem);
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
    r2iq_dual_d1_kernel<kTma, kVar><<<(unsigned)grid, kDualThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_dual_d1(const FusedParams& p, int sm_count, cudaStream_t stream, bool tma) {
    if (!tma) return launch_dual_one<false, 0>(p, sm_count, stream);
    if (p.nco_starts) return launch_dual_one<true, 1 | kVarMix>(p, sm_count, stream);   // default variant + fused mixer
    switch (kernel_variant() & 3) {
        case 1: return launch_dual_one<true, 1>(p, sm_count, stream);
        case 2: return launch_dual_one<true, 2>(p, sm_count, stream);
        case 3: return launch_dual_one<true, 3>(p, sm_count, stream);
        default: return launch_dual_one<true, 0>(p, sm_count, stream);
    }
}

}  // namespace r2iq
