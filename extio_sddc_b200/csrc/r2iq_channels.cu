// r2iq_channels.cu -- wideband channelizer: many tuned DDC channels from ONE ADC stream.
//
// The reference runs one fft_mt_r2iq per tuned channel, so C channels cost C forward FFTs of every
// frame (Core/fft_mt_r2iq_impl.hpp:84-98).  Here a warp transforms a frame ONCE, leaves the 1024
// half-spectrum bins Z in shared memory, and then serves the channels in groups of 2^d: each
// channel's mfft-bin band is cut out at its own tune bin (same split / filter / edge zero-fill
// arithmetic as the single-channel kernels: impl.hpp:76-96), staged, and the whole group is
// transformed back at once by inverse_bands (32>>d lanes per channel).  Per channel the result is
// the stream a separate operator tuned to that bin would produce.
// Output is channel-major: channel c writes out[c*out_stride + ...].
#include "r2iq_common.cuh"
#include "work_dispenser.cuh"

namespace r2iq {

namespace {

constexpr int kChWarps = 4;
constexpr int kChThreads = kChWarps * 32;
// per warp: transposition tile (kScratch) + staged bands (kStage) + the frame's spectrum (1024 bins)
constexpr int kChWarpFloat2 = kScratch + kStage + kHalfFft;

// Forward transform of frame f (same two radix-32 register passes as r2iq_fused_kernel).
__device__ __forceinline__ void forward_frame_ch(const ChannelParams& p, long long f, int lane, float2* S,
                                                 float2 (&v)[32]) {
    const long long b = f / kFramesPerBlock;
    const int k = (int)(f - b * kFramesPerBlock);
    const uint32_t* src = p.in_words + b * (kTransferSamples / 2) + k * (kHop / 2) + lane;
    uint32_t w[32];
#pragma unroll
    for (int j = 0; j < 32; j++) w[j] = __ldg(src + 32 * j);
#pragma unroll
    for (int j = 0; j < 32; j++) v[brev(j, 5)] = unpack_pair(w[j], p.rand_mul);
    dit<32, false>(v);
#pragma unroll
    for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul(v[k2], __ldg(&p.tw_fwd[k2 * 32 + lane]));
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) S[lane * 33 + k2] = v[k2];
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) v[brev(n1, 5)] = S[n1 * 33 + lane];
    __syncwarp();
    dit<32, false>(v);
}

}  // namespace

template <int D>
__global__ void __launch_bounds__(kChThreads, 2) r2iq_channels_kernel(const ChannelParams p) {
    constexpr bool kDyn = true;   // frames from the device-wide counter (work_dispenser.cuh); warps run free
    using G = Geo<D>;
    constexpr int M = G::M;
    constexpr int GROUP = G::ROUNDS;     // channels transformed back together
    extern __shared__ float2 smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float2* S = smem + warp * kChWarpFloat2;
    float2* stage = S + kScratch;
    float2* Z = stage + kStage;

    PairDispenser<kDyn, true, kChWarps> work;
    work.init(p.work_counter, nullptr, lane, warp);
    for (; work.more(p.total_frames); work.advance()) {
        work.top(p.work_counter, nullptr);
        const long long f = work.cur(warp);
        {
            float2 v[32];
            forward_frame_ch(p, f, lane, S, v);
#pragma unroll
            for (int k1 = 0; k1 < 32; k1++) Z[32 * k1 + lane] = v[k1];
        }
        __syncwarp();
        const FrameOut fo = frame_out<D>(f);
        for (int c0 = 0; c0 < p.n_channels; c0 += GROUP) {
            // cut the bands of channels c0 .. c0+GROUP-1: element e = lane + 32*i is bin m of channel j
#pragma unroll 4
            for (int i = 0; i < 32; i++) {
                const int e = lane + 32 * i;
                const int j = e / M;              // GROUP * M = 1024 elements for every d
                const int m = e & (M - 1);
                const int ch = c0 + j;
                float2 y = make_float2(0.f, 0.f);
                if (ch < p.n_channels) {
                    const int s = m < M / 2 ? m : m - M;
                    const int k = __ldg(&p.tunebins[ch]) + s;
                    if (k >= 0 && k < kHalfFft) {      // outside: zero fill (impl.hpp:91-96)
                        const float2 a = Z[k];
                        const float2 b = Z[(kHalfFft - k) & (kHalfFft - 1)];
                        const float2 c = __ldg(&p.ctab[k]);       // -i * W_2048^k
                        const float2 h = __ldg(&p.gh[m]);         // 0.5 * H_d at inverse-FFT bin m
                        const float ux = a.x + b.x, uy = a.y - b.y;
                        const float wx = a.x - b.x, wy = a.y + b.y;
                        const float xx = fmaf(c.x, wx, fmaf(-c.y, wy, ux));   // 2*X[k] = u + c*w
                        const float xy = fmaf(c.x, wy, fmaf(c.y, wx, uy));
                        y = make_float2(fmaf(h.x, xx, -h.y * xy), fmaf(h.x, xy, h.y * xx));
                    }
                }
                stage[j * G::GS + m] = y;
            }
            inverse_bands<D, true>(lane, S, stage, p.tw_inv, p.imag_sign, [&](int g) {
                Placement pl;
                pl.out = p.out + (long long)(c0 + g) * p.out_stride;
                pl.fo = fo;
                pl.live = c0 + g < p.n_channels;
                return pl;
            });
        }
        __syncwarp();
    }
}

template <int D>
static cudaError_t setup_ch() {
    cudaError_t e = cudaFuncSetAttribute(r2iq_channels_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kChWarps * kChWarpFloat2 * sizeof(float2)));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(r2iq_channels_kernel<D>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

cudaError_t setup_channel_kernels() {
    cudaError_t e;
    if ((e = setup_ch<1>()) != cudaSuccess) return e;
    if ((e = setup_ch<2>()) != cudaSuccess) return e;
    if ((e = setup_ch<3>()) != cudaSuccess) return e;
    if ((e = setup_ch<4>()) != cudaSuccess) return e;
    if ((e = setup_ch<5>()) != cudaSuccess) return e;
    if ((e = setup_ch<6>()) != cudaSuccess) return e;
    return cudaSuccess;
}

template <int D>
static cudaError_t launch_ch(const ChannelParams& p, int sm_count, cudaStream_t stream) {
    const size_t smem = kChWarps * kChWarpFloat2 * sizeof(float2);
    int ctas = 1;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas, r2iq_channels_kernel<D>, kChThreads, smem);
    if (e != cudaSuccess) return e;
    if (ctas < 1) ctas = 1;
    long long want = (p.total_frames + kChWarps - 1) / kChWarps;
    long long grid = (long long)sm_count * ctas;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    r2iq_channels_kernel<D><<<(unsigned)grid, kChThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_channels(int d, const ChannelParams& p, int sm_count, cudaStream_t stream) {
    switch (d) {
        case 1: return launch_ch<1>(p, sm_count, stream);
        case 2: return launch_ch<2>(p, sm_count, stream);
        case 3: return launch_ch<3>(p, sm_count, stream);
        case 4: return launch_ch<4>(p, sm_count, stream);
        case 5: return launch_ch<5>(p, sm_count, stream);
        case 6: return launch_ch<6>(p, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace r2iq
