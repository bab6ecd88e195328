// r2iq_gen1.cu -- FIRST-GENERATION fused kernels (one frame per warp, scalar FP32), FFTN = 2048.
// Built only with -DR2IQ_AB_VARIANTS (comparison runs: tools/variant_sweep.py, R2IQ_KERNEL_GEN=1); the shipped library
// runs r2iq_pair.cu.  Kept because its derivation comments are where the decomposition is explained.
//
// One kernel per decimation index d = 0..6 replaces the whole per-block worker loop of the
// reference (Core/fft_mt_r2iq_impl.hpp:15-151): de-randomise + int16->float (impl.hpp:40-54),
// 2048-point real forward FFT (impl.hpp:88), tune-bin select times filter response with edge
// zero-fill (impl.hpp:76-96), mfft-point backward FFT (impl.hpp:98), overlap-save discard,
// optional conjugate and placement in the output stream (impl.hpp:117-138).
//
// Mapping (see DESIGN.md for the derivation and the measured rooflines):
//   * a WARP owns a frame: lane n1 holds the 32 complex points z[n1 + 32*j] (z = packed int16
//     pairs x[2n] + i*x[2n+1]) and runs a radix-32 DFT entirely in registers (dft_regs.cuh);
//     one padded shared-memory transposition later the second radix-32 pass leaves
//     Z[32*k1 + lane] in register k1.  Only __syncwarp is ever needed -- no block barriers.
//   * the real-FFT split needs Z[k] and Z[1024-k]: the partner lives in lane (32-lane)&31,
//     register 31-k1, so it arrives by warp shuffle; split, tune rotation and filter multiply
//     collapse into  Y = GK[k]*(a+conj b) + GCK[k]*(a-conj b)  with two per-tune tables.
//   * d = 0: the tune rotation is absorbed into the inverse transform's inter-pass twiddles
//     (lane l simply plays residue (l - tunebin) mod 32), so Y never leaves registers between
//     the forward and the backward FFT.
//   * d >= 1: a warp runs 2^d forward frames, stages their mfft-bin bands in shared memory and
//     then transforms all of them at once, 32>>d lanes per frame, again 32 points per lane.
// No cuFFT, no tensor cores, no CPU fallback.
#include "r2iq_kernels.cuh"

#include <cstdio>
#include <cstdlib>

#include "r2iq_common.cuh"
#include "work_dispenser.cuh"

namespace r2iq {

// Small CTAs (4 warps) so that 3 of them fit an SM: 12 resident warps hide the DRAM latency of the
// 32 loads that open every frame.  A warp never talks to another warp, so CTA size is purely an
// occupancy knob (registers: 3*128*170 <= 64K; shared memory: 3*68.6 KB <= 227 KB for d >= 1).
constexpr int kWarpsPerCta = 4;
constexpr int kThreads = kWarpsPerCta * 32;
// d = 2, 5, 6: a unit of work is HALF the 2^d frames one backward batch could hold.  The backward
// transform then runs with half its lane groups idle, but staging shrinks to 4.3 KB per warp and the
// kernel fits 128 registers, so FOUR CTAs = 16 warps are resident per SM instead of 12.  Measured
// (Gsamples/s, 12 -> 16 warps): d=2 474 -> 501, d=5 683 -> 739, d=6 639 -> 726; d=3 591 -> 573 and
// d=4 656 -> 639 lose (their idle backward lanes cost more than the extra warps hide), d=1 is even:
// those keep full units and 3 CTAs.  d = 0 needs 165 registers: 3 CTAs.
template <int D>
__host__ __device__ constexpr bool half_units() { return D == 2 || D == 5 || D == 6; }
template <int D>
__host__ __device__ constexpr int unit_frames() { return D == 0 ? 1 : (half_units<D>() ? (1 << (D - 1)) : (1 << D)); }
template <int D>
__host__ __device__ constexpr int min_ctas_v1() { return half_units<D>() ? 4 : 3; }
template <int D>
__host__ __device__ constexpr int warp_smem_v1() { return kScratch + (D == 0 ? 0 : unit_frames<D>() * Geo<D>::GS); }   // float2
template <int D>
constexpr size_t fused_smem_bytes() { return (size_t)kWarpsPerCta * warp_smem_v1<D>() * sizeof(float2); }

// Forward transform of frame f: on return v[k1] = Z[32*k1 + lane], Z = FFT_1024 of the packed frame.
// `f_next` (>= 0) is the frame this warp will run after f: its 4 KB are requested into L2 now, one
// 128-byte line per lane, so that its loads find them there (L2 hit instead of a DRAM round trip).
__device__ __forceinline__ void forward_frame(const FusedParams& p, long long f, long long f_next, int lane, float2* S,
                                              float2 (&v)[32]) {
    const long long b = f / kFramesPerBlock;
    const int k = (int)(f - b * kFramesPerBlock);
    // frame k of block b starts k*1536 samples into [halo | block b]  (impl.hpp:88)
    const uint32_t* src = p.in_words + b * (kTransferSamples / 2) + k * (kHop / 2) + lane;
    uint32_t w[32];
#pragma unroll
    for (int j = 0; j < 32; j++) w[j] = __ldg(src + 32 * j);
    if (f_next >= 0 && f_next < p.total_frames) {
        const long long bn = f_next / kFramesPerBlock;
        const int kn = (int)(f_next - bn * kFramesPerBlock);
        const uint32_t* nxt = p.in_words + bn * (kTransferSamples / 2) + kn * (kHop / 2) + 32 * lane;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nxt));
    }
#pragma unroll
    for (int j = 0; j < 32; j++) v[brev(j, 5)] = unpack_pair(w[j], p.rand_mul);
    dit<32, false>(v);                                  // pass 1: DFT over n2, lane = n1
#pragma unroll
    for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul(v[k2], __ldg(&p.tw_fwd[k2 * 32 + lane]));
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) S[lane * 33 + k2] = v[k2];
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) v[brev(n1, 5)] = S[n1 * 33 + lane];
    __syncwarp();
    dit<32, false>(v);                                  // pass 2: DFT over n1, lane = k2
}

// Real-FFT split + tune rotation + filter for the bin k = 32*k1 + lane held in v[k1]:
//   X[k] = 0.5*(a + conj b) + 0.5*(-i W_2048^k)*(a - conj b),  a = Z[k], b = Z[1024-k]
//   Y    = X[k] * H[...]  ->  GK[k]*(a + conj b) + GCK[k]*(a - conj b)
template <int K1>
__device__ __forceinline__ float2 split_filter(const FusedParams& p, const float2 (&v)[32], int lane) {
    const float2 a = v[K1];
    const float2 bs = v[31 - K1];
    float bx = __shfl_sync(0xffffffffu, bs.x, (32 - lane) & 31);
    float by = __shfl_sync(0xffffffffu, bs.y, (32 - lane) & 31);
    if (lane == 0) {   // bins 32*k1: the mirror 1024-32*k1 is lane 0's own register (32-k1)&31
        bx = v[(32 - K1) & 31].x;
        by = v[(32 - K1) & 31].y;
    }
    const float ux = a.x + bx, uy = a.y - by;
    const float wx = a.x - bx, wy = a.y + by;
    const float2 g = __ldg(&p.gk[32 * K1 + lane]);
    const float2 gc = __ldg(&p.gck[32 * K1 + lane]);
    float2 y;
    y.x = fmaf(gc.x, wx, fmaf(-gc.y, wy, fmaf(g.x, ux, -g.y * uy)));
    y.y = fmaf(gc.x, wy, fmaf(gc.y, wx, fmaf(g.x, uy, g.y * ux)));
    return y;
}

template <int D, int... K1>
__device__ __forceinline__ void band_d0(const FusedParams& p, const float2 (&v)[32], int lane, float2 (&t)[32],
                                        std::integer_sequence<int, K1...>) {
    ((t[brev(K1, 5)] = split_filter<K1>(p, v, lane)), ...);
}

template <int D, int K1>
__device__ __forceinline__ void band_row(const FusedParams& p, const float2 (&v)[32], int lane, float2* stage) {
    constexpr int M = Geo<D>::M;
    if (32 * K1 + 31 >= p.band_lo && 32 * K1 < p.band_hi) {      // warp-uniform
        float2 y = split_filter<K1>(p, v, lane);
        const int k = 32 * K1 + lane;
        if (k >= p.band_lo && k < p.band_hi) stage[(k - p.tunebin) & (M - 1)] = y;
    }
}
template <int D, int... K1>
__device__ __forceinline__ void band_rows(const FusedParams& p, const float2 (&v)[32], int lane, float2* stage,
                                          std::integer_sequence<int, K1...>) {
    (band_row<D, K1>(p, v, lane, stage), ...);
}

// kDyn: units of work from the device-wide counter instead of a fixed stride (work_dispenser.cuh)
template <int D, bool kDyn = false>
__global__ void __launch_bounds__(kThreads, min_ctas_v1<D>()) r2iq_fused_kernel(const FusedParams p) {
    using G = Geo<D>;
    extern __shared__ float2 smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int UNIT = unit_frames<D>();
    float2* S = smem + warp * warp_smem_v1<D>();
    float2* stage = S + kScratch;

    if constexpr (D > 0) {
        // bins of the band that fall outside [0, halfFft) are never written: they must read as
        // zero (the memset calls of impl.hpp:91-96)
        for (int i = lane; i < UNIT * G::GS; i += 32) stage[i] = make_float2(0.f, 0.f);
        __syncwarp();
    }

    const long long n_units = (p.total_frames + UNIT - 1) / UNIT;
    PairDispenser<kDyn, true, kWarpsPerCta> work;          // warps run free: one unit per grab
    work.init(p.work_counter, nullptr, lane, warp);
    for (; work.more(n_units); work.advance()) {
        work.top(p.work_counter, nullptr);
        const long long unit = work.cur(warp);
        const long long next_unit = work.next(warp);       // its first frame is prefetched into L2
        const long long f0 = unit * UNIT;
        if constexpr (D == 0) {
            float2 v[32], t[32];
            forward_frame(p, f0, next_unit * UNIT, lane, S, v);
            band_d0<D>(p, v, lane, t, std::make_integer_sequence<int, 32>{});
            // backward pass 1 over the register index; lane l is residue (l - tunebin) mod 32
            dit<32, true>(t);
#pragma unroll
            for (int k2 = 1; k2 < 32; k2++) t[k2] = cmul(t[k2], __ldg(&p.tw_inv[k2 * 32 + lane]));
            const int n1 = (lane - p.tunebin) & 31;
#pragma unroll
            for (int k2 = 0; k2 < 32; k2++) S[n1 * 33 + k2] = t[k2];
            __syncwarp();
#pragma unroll
            for (int n = 0; n < 32; n++) t[brev(n, 5)] = S[n * 33 + lane];
            __syncwarp();
            dit<32, true>(t);                            // t[k1'] = y[lane + 32*k1']
            const FrameOut fo = frame_out<D>(f0);
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) store_iq(p.out, fo, lane + 32 * k1, t[k1], p.imag_sign);
        } else {
            for (int r = 0; r < UNIT; r++) {
                const long long f = f0 + r;
                if (f < p.total_frames) {                // warp-uniform
                    float2 v[32];
                    forward_frame(p, f, r + 1 < UNIT ? f + 1 : next_unit * UNIT,
                                  lane, S, v);
                    band_rows<D>(p, v, lane, stage + r * G::GS, std::make_integer_sequence<int, 32>{});
                }
            }
            inverse_bands<D, false, UNIT>(lane, S, stage, p.tw_inv, p.imag_sign, [&](int g) {
                Placement pl;
                pl.out = p.out;
                pl.fo = frame_out<D>(f0 + g);
                pl.live = f0 + g < p.total_frames;
                return pl;
            });
        }
    }
}

template <int D, bool kDyn>
static cudaError_t launch_one_v(const FusedParams& p, int sm_count, cudaStream_t stream) {
    int ctas_per_sm = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, r2iq_fused_kernel<D, kDyn>, kThreads,
                                                                 fused_smem_bytes<D>());
    if (e != cudaSuccess) return e;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    const long long n_units = (p.total_frames + unit_frames<D>() - 1) / unit_frames<D>();
    long long want = (n_units + kWarpsPerCta - 1) / kWarpsPerCta;
    long long grid = (long long)sm_count * ctas_per_sm;     // persistent: every CTA strides over units
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    if constexpr (kDyn) {
        e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    r2iq_fused_kernel<D, kDyn><<<(unsigned)grid, kThreads, fused_smem_bytes<D>(), stream>>>(p);
    return cudaGetLastError();
}
template <int D>
static cudaError_t launch_one(const FusedParams& p, int sm_count, cudaStream_t stream) {
    return (kernel_variant() & 1) ? launch_one_v<D, true>(p, sm_count, stream) : launch_one_v<D, false>(p, sm_count, stream);
}

template <int D, bool kDyn>
static cudaError_t setup_one_v() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_fused_kernel<D, kDyn>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)fused_smem_bytes<D>());
    if (e != cudaSuccess) return e;
    // ask for the largest shared-memory carve-out so that kMinCtasPerSm CTAs are resident
    return cudaFuncSetAttribute(r2iq_fused_kernel<D, kDyn>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}
template <int D>
static cudaError_t setup_one() {
    cudaError_t e = setup_one_v<D, false>();
    if (e != cudaSuccess) return e;
    return setup_one_v<D, true>();
}
cudaError_t setup_gen1_kernels() {
    cudaError_t e;
    if ((e = setup_one<0>()) != cudaSuccess) return e;
    if ((e = setup_one<1>()) != cudaSuccess) return e;
    if ((e = setup_one<2>()) != cudaSuccess) return e;
    if ((e = setup_one<3>()) != cudaSuccess) return e;
    if ((e = setup_one<4>()) != cudaSuccess) return e;
    if ((e = setup_one<5>()) != cudaSuccess) return e;
    return setup_one<6>();
}

cudaError_t launch_gen1(int d, const FusedParams& p, int sm_count, cudaStream_t stream) {
    switch (d) {
        case 0: return launch_one<0>(p, sm_count, stream);
        case 1: return launch_one<1>(p, sm_count, stream);
        case 2: return launch_one<2>(p, sm_count, stream);
        case 3: return launch_one<3>(p, sm_count, stream);
        case 4: return launch_one<4>(p, sm_count, stream);
        case 5: return launch_one<5>(p, sm_count, stream);
        case 6: return launch_one<6>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
