// r2iq_fft.cu -- the register FFTs of the r2iq path as stand-alone batched transforms: the CUDA side of the
// reference's FFT backend abstraction (Core/fft_backend.h:22-50: plan_r2c / plan_c2c / execute_* / alloc / free),
// at the sizes its benchmark sweeps (Core/fft_benchmark.cpp:12-15: 4096 ... 128).  FFTW conventions throughout:
// unnormalised, forward = exp(-2 pi i nk/N), r2c of n reals yields n/2 + 1 bins.
//
// One plan executes `batch` transforms per launch.  A transform of N = A x B complex points is two passes of
// register DFTs around one shared-memory transposition (the scheme of the r2iq kernels, dft_regs.cuh):
//   pass 1   thread n2 < B loads x[n1 B + n2] (coalesced over n2), runs the A-point DFT over n1 in registers, applies
//            W_N^(n2 k1) and stores to S[k1][n2] (pitch B + 1: conflict-free both ways);
//   pass 2   thread k1 < A reads row k1, runs the B-point DFT over n2, and owns X[k1 + A k2] for every k2
//            (stores coalesced over k1).
// r2c of n = 2N reals is the N-point transform of z[m] = x[2m] + i x[2m+1] followed by the split
//   X[k] = E[k] + W_2N^k O[k],  X[N-k] = conj(E[k] - W_2N^k O[k]),  E = (Z[k] + conj Z[N-k])/2,  O = (Z[k] - conj Z[N-k])/(2i),
// done from shared memory by the same threads.  max(A, B) threads form the team of one transform; a CTA of 128
// threads holds 128 / max(A, B) teams.  HBM traffic is the algorithmic minimum: every point read once, written once.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>
#include <vector>

#include "../../include/r2iq_cuda.h"
#include "dft_regs.cuh"
#include "r2iq_kernels.cuh"

namespace r2iq {
void set_last_error(const char* msg);   // r2iq_capi.cu: the thread-local string behind r2iq_cuda_last_error()
}

using namespace r2iq;

namespace {

constexpr int kFftThreads = 128;
constexpr double kPiF64 = 3.14159265358979323846;

int fft_fail(int code, const char* fmt, ...) {
    char msg[384];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(msg, sizeof(msg), fmt, ap);
    va_end(ap);
    set_last_error(msg);
    return code;
}

template <int BITS>
__device__ __forceinline__ constexpr int rev_bits(int v) {
    int r = 0;
    for (int i = 0; i < BITS; i++) r |= ((v >> i) & 1) << (BITS - 1 - i);
    return r;
}

template <int A, int B>
struct FftGeo {
    static constexpr int N = A * B;
    static constexpr int kTeam = A > B ? A : B;
    static constexpr int kTeams = kFftThreads / kTeam;
    static constexpr int kPitch = B + 1;
    // the r2c split reads Z in natural order from a second region (N + 1 slots), so that no barrier is needed between
    // the last read of the transposition tile and the first write of Z
    static constexpr int kTile = A * kPitch;
    static constexpr int smem_float2(bool r2c) { return kTeams * (kTile + (r2c ? N : 0)); }
};

// in / out: `batch` transforms back to back; for R2C `in` is the float2 view of the 2N reals and `out` has pitch N + 1.
template <int A, int B, bool INV, bool R2C>
__global__ void __launch_bounds__(kFftThreads) fft_batch_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                                 const float2* __restrict__ tw, const float2* __restrict__ tw_split,
                                                                 long batch) {
    using G = FftGeo<A, B>;
    constexpr int N = G::N;
    extern __shared__ float2 fft_smem[];
    const int team = threadIdx.x / G::kTeam;
    const int t = threadIdx.x % G::kTeam;
    float2* S = fft_smem + team * (G::kTile + (R2C ? N : 0));
    float2* Z = S + G::kTile;
    const long per_cta = G::kTeams;
    for (long first = (long)blockIdx.x * per_cta; first < batch; first += (long)gridDim.x * per_cta) {
        const long x = first + team;
        const bool live = x < batch;
        const float2* src = in + x * N;
        if (live && t < B) {
            float2 v[A];
#pragma unroll
            for (int n1 = 0; n1 < A; n1++) v[rev_bits<ilog2(A)>(n1)] = src[n1 * B + t];
            dit<A, INV>(v);
#pragma unroll
            for (int k1 = 1; k1 < A; k1++) {
                const float2 w = __ldg(&tw[t * k1]);
                v[k1] = INV ? cmul_conj(v[k1], w) : cmul(v[k1], w);
            }
#pragma unroll
            for (int k1 = 0; k1 < A; k1++) S[k1 * G::kPitch + t] = v[k1];
        }
        __syncthreads();
        if (live && t < A) {
            float2 v[B];
#pragma unroll
            for (int n2 = 0; n2 < B; n2++) v[rev_bits<ilog2(B)>(n2)] = S[t * G::kPitch + n2];
            dit<B, INV>(v);
            if constexpr (R2C) {
#pragma unroll
                for (int k2 = 0; k2 < B; k2++) Z[t + A * k2] = v[k2];
            } else {
                float2* dst = out + x * N;
#pragma unroll
                for (int k2 = 0; k2 < B; k2++) dst[t + A * k2] = v[k2];
            }
        }
        __syncthreads();
        if constexpr (R2C) {
            if (live) {
                float2* dst = out + x * (N + 1);
                for (int k = t; k <= N / 2; k += G::kTeam) {
                    const float2 a = Z[k], b = Z[(N - k) & (N - 1)];
                    const float2 e = make_float2(0.5f * (a.x + b.x), 0.5f * (a.y - b.y));
                    const float2 o = make_float2(0.5f * (a.y + b.y), 0.5f * (b.x - a.x));   // (a - conj b) / (2i)
                    const float2 wo = cmul(o, __ldg(&tw_split[k]));
                    dst[k] = make_float2(e.x + wo.x, e.y + wo.y);
                    dst[N - k] = make_float2(e.x - wo.x, wo.y - e.y);
                }
            }
            __syncthreads();   // Z is rewritten by the next transform of this team
        }
    }
}

using FftKernel = void (*)(const float2*, float2*, const float2*, const float2*, long);

struct FftVariant {
    FftKernel fn;
    int smem_bytes;
    int teams;
};

template <int A, int B>
FftVariant pick_ab(int kind) {
    using G = FftGeo<A, B>;
    FftVariant v;
    v.teams = G::kTeams;
    if (kind == R2IQ_FFT_R2C) {
        v.fn = fft_batch_kernel<A, B, false, true>;
        v.smem_bytes = G::smem_float2(true) * (int)sizeof(float2);
    } else {
        v.fn = kind == R2IQ_FFT_C2C_FORWARD ? fft_batch_kernel<A, B, false, false> : fft_batch_kernel<A, B, true, false>;
        v.smem_bytes = G::smem_float2(false) * (int)sizeof(float2);
    }
    return v;
}

// complex length -> (A, B): squares where the length allows, the larger radix in pass 1 otherwise
bool pick(int n_complex, int kind, FftVariant* v) {
    switch (n_complex) {
        case 64: *v = pick_ab<8, 8>(kind); return true;
        case 128: *v = pick_ab<16, 8>(kind); return true;
        case 256: *v = pick_ab<16, 16>(kind); return true;
        case 512: *v = pick_ab<32, 16>(kind); return true;
        case 1024: *v = pick_ab<32, 32>(kind); return true;
        case 2048: *v = pick_ab<64, 32>(kind); return true;
        case 4096: *v = pick_ab<64, 64>(kind); return true;
    }
    return false;
}

}  // namespace

struct r2iq_cuda_fft_s {
    int device = 0;
    int kind = 0;
    int n = 0;            // transform length as the caller states it (reals for r2c)
    int n_complex = 0;    // points of the complex transform underneath
    long batch = 1;
    FftVariant var{};
    int grid_cap = 1;
    float2* d_tw = nullptr;
    float2* d_split = nullptr;
    void* d_in = nullptr;     // staging for pageable / large host buffers
    void* d_out = nullptr;
    size_t in_bytes = 0, out_bytes = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
};

namespace {

struct FftDeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit FftDeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != device) ok = cudaSetDevice(device) == cudaSuccess;
    }
    ~FftDeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int launch_fft(r2iq_cuda_fft_s* p, const void* d_in, void* d_out, long batch, cudaStream_t stream) {
    if (batch <= 0) return R2IQ_OK;
    long ctas = (batch + p->var.teams - 1) / p->var.teams;
    if (ctas > p->grid_cap) ctas = p->grid_cap;
    p->var.fn<<<(unsigned)ctas, kFftThreads, p->var.smem_bytes, stream>>>((const float2*)d_in, (float2*)d_out, p->d_tw, p->d_split, batch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fft_fail(R2IQ_E_CUDA, "fft kernel launch: %s", cudaGetErrorString(e));
    return R2IQ_OK;
}

// pinned (or registered) host memory is addressable from the device: small transforms then run straight on the
// caller's buffers, with no staging copies (what keeps a single 4096-point call near one launch latency)
bool device_addressable_host(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

constexpr size_t kZeroCopyLimit = 256 << 10;   // bytes per direction up to which host buffers are read in place

}  // namespace

extern "C" {

int r2iq_cuda_fft_plan(int device, int kind, int n, long batch, r2iq_cuda_fft_t* out) {
    if (!out) return fft_fail(R2IQ_E_INVALID, "fft_plan: out is null");
    *out = nullptr;
    if (kind < R2IQ_FFT_R2C || kind > R2IQ_FFT_C2C_BACKWARD) return fft_fail(R2IQ_E_INVALID, "fft_plan: kind %d", kind);
    if (batch < 1) return fft_fail(R2IQ_E_INVALID, "fft_plan: batch %ld", batch);
    const int nc = kind == R2IQ_FFT_R2C ? n / 2 : n;
    FftVariant var;
    if (n <= 0 || (kind == R2IQ_FFT_R2C && (n & 1)) || !pick(nc, kind, &var))
        return fft_fail(R2IQ_E_INVALID, "fft_plan: length %d unsupported (r2c 128..8192, c2c 64..4096, powers of two)", n);
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fft_fail(R2IQ_E_NODEVICE, "fft_plan: no CUDA device (there is no CPU path)");
    }
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return fft_fail(R2IQ_E_CUDA, "cudaGetDevice failed");
    if (device >= count) return fft_fail(R2IQ_E_INVALID, "fft_plan: device %d of %d", device, count);
    FftDeviceGuard guard(device);
    if (!guard.ok) return fft_fail(R2IQ_E_CUDA, "cudaSetDevice(%d) failed", device);
    auto* p = new r2iq_cuda_fft_s();
    p->device = device;
    p->kind = kind;
    p->n = n;
    p->n_complex = nc;
    p->batch = batch;
    p->var = var;
    cudaError_t e = cudaFuncSetAttribute(var.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, var.smem_bytes);
    int per_sm = 0, sms = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, var.fn, kFftThreads, var.smem_bytes);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    p->grid_cap = (per_sm > 0 ? per_sm : 1) * (sms > 0 ? sms : 1);
    // twiddles in float64, rounded once
    std::vector<float2> tw(nc), split(nc / 2 + 1);
    for (int j = 0; j < nc; j++) {
        const double a = -2.0 * kPiF64 * j / nc;
        tw[j] = make_float2((float)cos(a), (float)sin(a));
    }
    for (int k = 0; k <= nc / 2; k++) {
        const double a = -kPiF64 * k / nc;
        split[k] = make_float2((float)cos(a), (float)sin(a));
    }
    p->in_bytes = (size_t)batch * (kind == R2IQ_FFT_R2C ? (size_t)n * sizeof(float) : (size_t)n * sizeof(float2));
    p->out_bytes = (size_t)batch * (kind == R2IQ_FFT_R2C ? (size_t)(nc + 1) : (size_t)n) * sizeof(float2);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_tw, tw.size() * sizeof(float2));
    if (e == cudaSuccess) e = cudaMalloc(&p->d_split, split.size() * sizeof(float2));
    if (e == cudaSuccess) e = cudaMemcpy(p->d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_split, split.data(), split.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        const int code = e == cudaErrorMemoryAllocation ? R2IQ_E_NOMEM : R2IQ_E_CUDA;
        fft_fail(code, "fft_plan: %s", cudaGetErrorString(e));
        r2iq_cuda_fft_destroy(p);
        return code;
    }
    *out = p;
    return R2IQ_OK;
}

int r2iq_cuda_fft_destroy(r2iq_cuda_fft_t p) {
    if (!p) return R2IQ_OK;
    FftDeviceGuard guard(p->device);
    if (p->stream) {
        cudaStreamSynchronize(p->stream);
        cudaStreamDestroy(p->stream);
    }
    cudaFree(p->d_tw);
    cudaFree(p->d_split);
    cudaFree(p->d_in);
    cudaFree(p->d_out);
    delete p;
    return R2IQ_OK;
}

int r2iq_cuda_fft_execute_device(r2iq_cuda_fft_t p, const void* d_in, void* d_out, long batch, void* stream) {
    if (!p || !d_in || !d_out) return fft_fail(R2IQ_E_INVALID, "fft_execute_device: null argument");
    if (batch < 0) return fft_fail(R2IQ_E_INVALID, "fft_execute_device: batch %ld", batch);
    FftDeviceGuard guard(p->device);
    return launch_fft(p, d_in, d_out, batch, (cudaStream_t)stream);
}

int r2iq_cuda_fft_execute(r2iq_cuda_fft_t p, const void* in, void* out) {
    if (!p || !in || !out) return fft_fail(R2IQ_E_INVALID, "fft_execute: null argument");
    FftDeviceGuard guard(p->device);
    std::lock_guard<std::mutex> lk(p->mu);
    cudaError_t e = cudaSuccess;
    int rc;
    if (p->in_bytes <= kZeroCopyLimit && p->out_bytes <= kZeroCopyLimit && device_addressable_host(in) && device_addressable_host(out)) {
        rc = launch_fft(p, in, out, p->batch, p->stream);
        if (rc != R2IQ_OK) return rc;
    } else {
        if (!p->d_in) e = cudaMalloc(&p->d_in, p->in_bytes);
        if (e == cudaSuccess && !p->d_out) e = cudaMalloc(&p->d_out, p->out_bytes);
        if (e != cudaSuccess) return fft_fail(R2IQ_E_NOMEM, "fft_execute: %s", cudaGetErrorString(e));
        e = cudaMemcpyAsync(p->d_in, in, p->in_bytes, cudaMemcpyHostToDevice, p->stream);
        if (e != cudaSuccess) return fft_fail(R2IQ_E_CUDA, "fft_execute H2D: %s", cudaGetErrorString(e));
        rc = launch_fft(p, p->d_in, p->d_out, p->batch, p->stream);
        if (rc != R2IQ_OK) return rc;
        e = cudaMemcpyAsync(out, p->d_out, p->out_bytes, cudaMemcpyDeviceToHost, p->stream);
        if (e != cudaSuccess) return fft_fail(R2IQ_E_CUDA, "fft_execute D2H: %s", cudaGetErrorString(e));
    }
    e = cudaStreamSynchronize(p->stream);
    if (e != cudaSuccess) return fft_fail(R2IQ_E_CUDA, "fft_execute: %s", cudaGetErrorString(e));
    return R2IQ_OK;
}

long r2iq_cuda_fft_batch(r2iq_cuda_fft_t p) { return p ? p->batch : 0; }

}  // extern "C"
