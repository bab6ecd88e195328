// Microbenchmark for the next kernel generation (DESIGN.md "where the kernels stand"): one radix-32 register pass +
// inter-pass twiddles + shared-memory transposition, repeated, in two register layouts
//   dual : TWO frames per warp in the packed fp32x2 lanes (today's kernels): 32 packed complex values = 128 data
//          registers, 8 warps per SM;
//   half : ONE frame per warp, even/odd-indexed points of the frame in the packed lanes: the two 16-point
//          sub-transforms of a decimation-in-time radix-32 pass share their twiddles and run packed (4 of 5 stages),
//          the last stage runs scalar; 16 packed complex values = 64 data registers, 16 warps per SM.
// Prints frame-passes per microsecond for the whole GPU; same arithmetic per frame in both.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I extio_sddc_b200/csrc tools/ubench/fft_pass_layouts.cu
#include <cstdio>
#include <utility>

#include "dft_regs2.cuh"

using namespace r2iq;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int kTile = 32 * 33;
__device__ __forceinline__ int brev5(int v) { return (int)(__brev((unsigned)v) >> 27); }
__device__ __forceinline__ int brev4(int v) { return (int)(__brev((unsigned)v) >> 28); }

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

__global__ void __launch_bounds__(128, 2) k_dual(float2* out, const float2* __restrict__ tw, int iters) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* S = reinterpret_cast<u64*>(sm + (size_t)warp * kTile * 16);
    c2 v[32];
#pragma unroll
    for (int j = 0; j < 32; j++) {
        v[j].re = pk(0.001f * (lane + j), 0.002f * (lane - j));
        v[j].im = pk(0.003f * j, 0.001f * lane);
    }
    for (int it = 0; it < iters; it++) {
        dit2<32, false>(v);
#pragma unroll
        for (int k2 = 1; k2 < 32; k2++) v[k2] = cmul2s(v[k2], __ldg(tw + 32 * k2 + lane));
#pragma unroll
        for (int k2 = 0; k2 < 32; k2++) sts2(S, lane * 33 + k2, v[k2]);
        __syncwarp();
#pragma unroll
        for (int n1 = 0; n1 < 32; n1++) {
            c2 t = lds2(S, n1 * 33 + lane);
            t.re = mul2(t.re, bc(0.03125f));      // keep magnitudes bounded over many passes
            t.im = mul2(t.im, bc(0.03125f));
            v[brev5(n1)] = t;
        }
        __syncwarp();
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 32; j++) {
        float a, b, c, d;
        upk(v[j].re, a, b);
        upk(v[j].im, c, d);
        s += a + b + c + d;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = make_float2(s, 0.f);
}

template <int... K>
__device__ __forceinline__ void last_stage(const c2 (&v)[16], float* re, float* im, std::integer_sequence<int, K...>) {
    (([&] {
         float er, orr, ei, oi;
         upk(v[K].re, er, orr);
         upk(v[K].im, ei, oi);
         float2 a = make_float2(er, ei), b = make_float2(orr, oi);
         bfly<K, 32, false>(a, b);               // X[k] = E[k] + w^k O[k], X[k+16] = E[k] - w^k O[k]
         re[2 * K] = a.x;                        // straight into the transposition tile: x[] never lives in registers
         im[2 * K] = a.y;
         re[2 * (K + 16)] = b.x;
         im[2 * (K + 16)] = b.y;
     }()),
     ...);
}

template <int MINB>
__global__ void __launch_bounds__(128, MINB) k_half(float2* out, const float4* __restrict__ tw4, int iters) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // tile: plane RE then plane IM, each [16][33][2] floats: point n1 of column k2 at ((n1 >> 1) * 33 + k2) * 2 + (n1 & 1),
    // so the reader finds points 2m and 2m+1 of its column adjacent -- one LDS.64 delivers a packed register pair
    float* RE = reinterpret_cast<float*>(sm + (size_t)warp * kTile * 8);
    float* IM = RE + kTile;
    c2 v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        v[j].re = pk(0.001f * (lane + j), 0.002f * (lane - j));
        v[j].im = pk(0.003f * j, 0.001f * lane);
    }
    for (int it = 0; it < iters; it++) {
        dit2<16, false>(v);                        // lo lanes: DFT16 of the even points, hi lanes: of the odd points
        last_stage(v, RE + (lane >> 1) * 66 + (lane & 1), IM + (lane >> 1) * 66 + (lane & 1), std::make_integer_sequence<int, 16>{});
        __syncwarp();
#pragma unroll
        for (int m = 0; m < 16; m++) {             // next pass input: points 2m, 2m+1 packed, inter-pass twiddles packed
            const float2 pr = reinterpret_cast<const float2*>(RE)[m * 33 + lane], pi = reinterpret_cast<const float2*>(IM)[m * 33 + lane];
            const float4 w = __ldg(tw4 + 32 * m + lane);            // (wr_2m, wr_2m+1, wi_2m, wi_2m+1)
            const u64 wr = pk(w.x, w.y), wi = pk(w.z, w.w);
            const u64 re = pk(pr.x, pr.y), im = pk(pi.x, pi.y);
            c2 t;
            t.re = fma2(wr, re, neg2(mul2(wi, im)));
            t.im = fma2(wr, im, mul2(wi, re));
            v[brev4(m)] = t;
        }
        __syncwarp();
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) {
        float a, b, c, d;
        upk(v[j].re, a, b);
        upk(v[j].im, c, d);
        s += a + b + c + d;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = make_float2(s, 0.f);
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    float2* out;
    float2* tw;
    CK(cudaMalloc(&out, sizeof(float2) * sms * 4 * 128));
    CK(cudaMalloc(&tw, sizeof(float4) * 32 * 32));
    float2 h[2 * 32 * 32];
    for (int i = 0; i < 2 * 32 * 32; i++) h[i] = make_float2(0.031f, 0.003f);   // |w| << 1: values stay bounded
    CK(cudaMemcpy(tw, h, sizeof(h), cudaMemcpyHostToDevice));
    const int iters = 2000;
    const int smem_dual = 4 * kTile * 16, smem_half = 4 * kTile * 8;
    CK(cudaFuncSetAttribute(k_dual, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dual));
    CK(cudaFuncSetAttribute(k_half<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_half));
    CK(cudaFuncSetAttribute(k_half<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_half));
    CK(cudaFuncSetAttribute(k_dual, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_half<4>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_half<3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int ctas = 1; ctas <= 4; ctas++) {
        for (int which = 0; which < 3; which++) {
            if (which == 0 && ctas > 2) continue;
            int occ = 0;
            if (which == 0) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_dual, 128, smem_dual));
            else if (which == 1) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_half<4>, 128, smem_half));
            else CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_half<3>, 128, smem_half));
            if (ctas > occ) continue;
            float best = 1e30f;
            for (int rep = 0; rep < 4; rep++) {
                CK(cudaEventRecord(e0));
                if (which == 0) k_dual<<<sms * ctas, 128, smem_dual>>>(out, tw, iters);
                else if (which == 1) k_half<4><<<sms * ctas, 128, smem_half>>>(out, (const float4*)tw, iters);
                else k_half<3><<<sms * ctas, 128, smem_half>>>(out, (const float4*)tw, iters);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep && ms < best) best = ms;
            }
            const double frames = (double)sms * ctas * 4 * iters * (which == 0 ? 2 : 1);
            printf("%-5s %2d warps/SM (occupancy limit %d CTAs): %8.3f ms  %8.1f frame-passes/us  (%.0f SM-cycles per frame-pass at %d MHz)\n",
                   which == 0 ? "dual" : which == 1 ? "half4" : "half3", ctas * 4, occ, best, frames / best / 1e3,
                   best * 1e-3 * prop.clockRate * 1e3 * sms / frames, prop.clockRate / 1000);
        }
    }
    CK(cudaGetLastError());
    printf("done\n");
    return 0;
}
