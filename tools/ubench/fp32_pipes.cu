// Microbenchmark: which FP32 instruction forms bound an FFT butterfly kernel on sm_100a?
// Measures per-SM throughput (thread-ops per clock per SM) of scalar FFMA/FADD, the packed
// f32x2 forms (PTX ISA 8.6, sm_100+), int16->float conversion variants, and LDS.64 mixed in.
// Used once to choose the register layout of the r2iq FFT kernels.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int NACC = 16;

__global__ void k_ffma(float* out, float a, float b) {
    float acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_fadd(float* out, float a, float b) {
    float acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = acc[i] + ((i & 1) ? a : b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// butterfly-like mix: c = a+b, d = a-b on independent registers (2 FADD, 3-reg form)
__global__ void k_bfly(float* out, float a0) {
    float x[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) x[i] = threadIdx.x * 0.001f + i * a0;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i += 2) {
            float s = x[i] + x[i + 1];
            float d = x[i] - x[i + 1];
            x[i] = s; x[i + 1] = d * 0.5f;   // keeps values bounded; FMUL counts too
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ unsigned long long pk(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

__global__ void k_ffma2(float* out, float a, float b) {
    unsigned long long acc[NACC];
    unsigned long long A = pk(a, a), B = pk(b, b);
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = pk(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(acc[i]) : "l"(acc[i]), "l"(A), "l"(B));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) { float lo, hi; upk(acc[i], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_fadd2(float* out, float a, float b) {
    unsigned long long acc[NACC];
    unsigned long long A = pk(a, b);
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = pk(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(acc[i]) : "l"(acc[i]), "l"(A));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) { float lo, hi; upk(acc[i], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// packed butterfly: s = x+y, d = x-y (sub.f32x2) on independent packed registers
__global__ void k_bfly2(float* out, float a0) {
    unsigned long long x[NACC];
    unsigned long long H = pk(0.5f, 0.5f);
#pragma unroll
    for (int i = 0; i < NACC; i++) x[i] = pk(threadIdx.x * 0.001f + i * a0, threadIdx.x * 0.003f + i);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i += 2) {
            unsigned long long s, d;
            asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(s) : "l"(x[i]), "l"(x[i + 1]));
            asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(x[i]), "l"(x[i + 1]));
            asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(d), "l"(H));
            x[i] = s; x[i + 1] = d;
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) { float lo, hi; upk(x[i], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// scalar FFMA interleaved 4:1 with LDS.64 (conflict-free) -- the FFT's transposition mix
__global__ void k_ffma_lds(float* out, float a, float b) {
    __shared__ float2 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_float2(i * 0.5f, i * 0.25f);
    __syncthreads();
    float acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 0.001f + i;
    int idx = threadIdx.x;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] = fmaf(acc[i], a, b);
#pragma unroll
        for (int j = 0; j < NACC / 4; j++) {
            float2 v = sm[(idx + 32 * j + it) & 1023];
            acc[j] += v.x; acc[j + 4] += v.y;
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// int16 pair -> two floats, plain C conversion (what does ptxas emit? I2F on which pipe?)
__global__ void k_cvt_plain(float* out, const uint32_t* in) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = in[threadIdx.x + 32 * i];
    float s = 0;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t v = w[i] + it;
            v ^= (v & 0x00010001u) * 0xFFFEu;
            s += (float)(short)(v & 0xFFFF);
            s += (float)((int)v >> 16);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// magic-number conversion: PRMT builds 0x4B40_0000 | (h ^ 0x8000), FADD removes the bias
__global__ void k_cvt_magic(float* out, const uint32_t* in) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = in[threadIdx.x + 32 * i];
    float s = 0;
    const float bias = 12582912.0f + 32768.0f;  // 0x4B400000 = 12582912.0
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t v = w[i] + it;
            v ^= (v & 0x00010001u) * 0xFFFEu;
            v ^= 0x80008000u;
            uint32_t lo = __byte_perm(v, 0x4B400000u, 0x7610);
            uint32_t hi = __byte_perm(v, 0x4B400000u, 0x7632);
            s += __uint_as_float(lo) - bias;
            s += __uint_as_float(hi) - bias;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_copy(const uint4* __restrict__ in, uint4* __restrict__ out, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = in[i];
}

template <typename F>
static float time_ms(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(e0);
        f();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("device %s  SMs %d  clock %d kHz  smem/SM %zu  regs/SM %d\n", p.name, p.multiProcessorCount, clk_khz,
           p.sharedMemPerMultiprocessor, p.regsPerMultiprocessor);
    const int sms = p.multiProcessorCount;
    float* out; uint32_t* in;
    CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 1024));
    CK(cudaMalloc(&in, sizeof(uint32_t) * 1024));
    CK(cudaMemset(in, 0x5a, sizeof(uint32_t) * 1024));
    for (int wpsm : {4, 8, 16, 32}) {   // warps per SM (one block per SM, wpsm*32 threads... up to 1024)
        int threads = wpsm * 32;
        dim3 g(sms), b(threads);
        double tot = (double)sms * threads * ITERS;
        auto rep = [&](const char* name, float ms, double ops_per_iter_thread, double flops_per_op) {
            double ops = tot * ops_per_iter_thread;
            printf("  warps/SM %2d  %-12s %8.3f ms  %7.2f Gop/s/SM-ish  %8.2f T-thread-ops/s  %8.2f TFLOP/s\n", wpsm, name, ms,
                   ops / ms / 1e6 / sms, ops / ms / 1e9, ops * flops_per_op / ms / 1e9);
        };
        rep("ffma", time_ms([&] { k_ffma<<<g, b>>>(out, 1.0001f, 0.5f); }), NACC, 2);
        rep("fadd", time_ms([&] { k_fadd<<<g, b>>>(out, 1.0001f, 0.5f); }), NACC, 1);
        rep("bfly(3/pair)", time_ms([&] { k_bfly<<<g, b>>>(out, 1.0001f); }), NACC / 2 * 3, 1);
        rep("ffma2", time_ms([&] { k_ffma2<<<g, b>>>(out, 1.0001f, 0.5f); }), NACC, 4);
        rep("fadd2", time_ms([&] { k_fadd2<<<g, b>>>(out, 1.0001f, 0.5f); }), NACC, 2);
        rep("bfly2", time_ms([&] { k_bfly2<<<g, b>>>(out, 1.0001f); }), NACC / 2 * 3, 2);
        rep("ffma+lds", time_ms([&] { k_ffma_lds<<<g, b>>>(out, 1.0001f, 0.5f); }), NACC + NACC / 4 + NACC / 2, 1);
        rep("cvt_plain", time_ms([&] { k_cvt_plain<<<g, b>>>(out, in); }), 16, 1);
        rep("cvt_magic", time_ms([&] { k_cvt_magic<<<g, b>>>(out, in); }), 16, 1);
    }
    // HBM copy for reference
    size_t nbytes = (size_t)2 << 30;
    uint4 *a, *b2;
    CK(cudaMalloc(&a, nbytes)); CK(cudaMalloc(&b2, nbytes));
    CK(cudaMemset(a, 1, nbytes));
    float ms = time_ms([&] { k_copy<<<sms * 16, 512>>>(a, b2, nbytes / 16); });
    printf("copy 2 GiB: %.3f ms  %.1f GB/s (read+write)\n", ms, 2.0 * nbytes / ms / 1e6);
    CK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
