// r2iq_capi.cu -- implementation of the C ABI declared in include/r2iq_cuda.h.
//
// Owns what the reference operator owns (filter spectra, per-worker scratch -- here device
// buffers, streams and lookup tables) and borrows what its caller owns (the sample buffers).
// All arithmetic on samples happens in the kernels of r2iq_kernels.cu; this file only designs
// filters, builds the small per-tune tables, and moves data.
#include <cuda_runtime.h>

#include <cmath>
#include <complex>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <atomic>
#include <sys/syscall.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <unistd.h>
#include <condition_variable>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/r2iq_cuda.h"
#include "fir_design.h"
#include "r2iq_kernels.cuh"

using namespace r2iq;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void poison(r2iq_cuda_t h, const char* why);   // defined once the handle type is complete
}  // namespace

namespace r2iq {
void set_last_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }   // for the other translation units
}

namespace {

constexpr double kPiD = 3.14159265358979323846;
constexpr int kDefaultChunkBlocks = 256;
constexpr int kSlots = 4;                 // chunks / submissions in flight on the host-buffer paths
constexpr size_t kMaxTableSets = 1024;    // cached (d, tunebin) table sets before the cache is recycled

struct TableSet {   // device tables for one (decimation, tune bin)
    float2* gk = nullptr;
    float2* gck = nullptr;
    float2* gp = nullptr;
    float2* gq = nullptr;
    float2* tw_inv = nullptr;
    float2* tw_aux = nullptr;
    float4* pq_band = nullptr;   // d >= 1, shipped kernels: (P, Q) of band position m (bin tunebin + s(m)), m < mfft
    int band_lo = 0, band_hi = 0;
    uint32_t row_mask = 0xFFFFFFFFu;
};

// One submission in flight on the host-buffer paths: its own stream, device buffers and pinned history.
//   kFree: nothing pending.
//   kAuto: the result is being copied to the caller's iq_out; whoever waits for `done` frees the slot
//          (r2iq_cuda_process, r2iq_cuda_process_rates, r2iq_cuda_process_async with an output pointer).
//   kHeld: the result stays in d_out until r2iq_cuda_release (r2iq_cuda_process_async without an output pointer:
//          the ring-buffer adapter fetches it slot by slot with r2iq_cuda_read).
struct Slot {
    enum State { kFree, kAuto, kHeld };
    cudaStream_t stream = nullptr;
    cudaEvent_t h2d_done = nullptr;   // the input of this submission has left the caller's buffer
    cudaEvent_t done = nullptr;       // everything enqueued for this submission has finished
    int16_t* d_in = nullptr;     // [halfFft | chunk_blocks * 65536]
    float* d_out = nullptr;      // chunk_blocks * 32768 complex (sized for d = 0)
    float* d_out_rates = nullptr;   // r2iq_cuda_process_rates: the chunk's output at up to 7 rates, allocated on first use
    float* d_starts = nullptr;   // fine-tune NCO: start phasors of the chunk's 128-sample blocks
    size_t starts_cap = 0;       // capacity of d_starts in 128-sample blocks
    int16_t* carry = nullptr;    // pinned: the halfFft samples in front of this submission
    State state = kFree;
    uint64_t ticket = 0;
    long n_blocks = 0;
    int d = 0;                   // decimation this submission ran at
};

// Host side of the fine-tune NCO: the state and small table of shift_limited_unroll_C_sse_init
// (Core/pffft/pf_mixer.cpp:750-787) and the once-per-128-samples recurrence of
// shift_limited_unroll_C_sse_inp_c (:840-853), in the same single-precision operations.
struct NcoState {
    bool on = false;
    float fc = 0.f;
    float tab[33][2];       // phasor of 4*(g+1) phase increments, g = 0..32 (dinterl_trig)
    float cs[4], sn[4];     // phase_state_i / phase_state_q
    // Direct-phase mode (R2IQ_NCO_DIRECT): no recurrence; phase of the next output sample in cycles, [0, 1)
    int mode = R2IQ_NCO_REFERENCE;
    double phase = 0.0;
    float* d_rot = nullptr;      // [128] exp(2 pi j fc 32 k): the 2048 family reads k < 32, the 8192 family k < 128
    float* d_tab = nullptr;
    float* d_starts = nullptr;   // device-path scratch (caller's stream keeps launches ordered)
    size_t starts_cap = 0;
    std::vector<float> h_starts;
};

}  // namespace

struct r2iq_cuda_s {
    int device = 0;
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
    float gain = 0.f;
    int chunk_blocks = kDefaultChunkBlocks;

    // control state (r2iqControlClass members, Core/r2iq.h:38-47; mtunebin, Core/fft_mt_r2iq.h:90)
    int fftn = kFftN;             // FFTN_R_ADC of this operator: 2048 (north star) or 8192 (reference as shipped)
    int half = kHalfFft;          // halfFft
    // decimate / tunebin / rand_on / sideband are written by the application thread while the worker streams
    // (Core/RadioHandler.cpp:289,335; unittest/core_test.cpp:176-197): atomics, and every launch reads the tune bin
    // ONCE -- tables, row mask and kernel parameter must describe the same bin (impl.hpp:20 snapshots it per block).
    std::atomic<int> decimate{0};
    std::atomic<int> tunebin{kHalfFft / 4};
    std::atomic<bool> rand_on{false};
    std::atomic<bool> sideband{false};
    bool on = false;
    int latched_decimate = 0;     // worker prologue latches these at TurnOn (impl.hpp:3-6)
    bool latched_sideband = false;

    bool poisoned = false;        // a CUDA error happened: refuse further work
    char poison_msg[512] = "";    // what poisoned it (g_err is thread-local: the application thread asks later)
    std::mutex mu;

    std::vector<std::complex<float>> hw[kNDec];   // filterHw[d], host copy
    uint64_t hw_version[kNDec] = {0};
    float2* d_tw_fwd = nullptr;
    // Work dispensers of the persistent kernels (work_dispenser.cuh).  Launches of one handle may be in flight on
    // several streams at once (the slots of the host pipeline), so every launch takes the next counter of a ring;
    // the ring is longer than the number of kernels a device can run concurrently.
    static constexpr unsigned kCounterRing = 256;
    unsigned long long* d_work_counter = nullptr;
    unsigned counter_next = 0;
    unsigned long long* next_counter() { return d_work_counter + (counter_next++ % kCounterRing); }
    std::map<long long, TableSet> tables;          // key = version<<20 | d<<12 | tunebin... see table_key
    std::map<long long, float2*> gh_tables;        // channelizer: 0.5*H_d in inverse-FFT bin order, key = version<<4 | d
    std::map<std::vector<int>, int*> chan_lists;   // channelizer: device copies of tune-bin lists
    float2* d_ctab = nullptr;                      // channelizer: -i * W_2048^k
    // FFTN 8192 family (r2iq_wide.cu)
    float2* d_twq = nullptr;                       // [3][1024] W_4096^(q n)
    float2* d_wctab = nullptr;                     // [2049] -i * W_8192^p
    float2* d_wctab_full = nullptr;                // [4096] the same for every bin (FFTN 8192 channelizer)
    struct WideSet { float2* g = nullptr; float2* tw_sub = nullptr; float4* pq = nullptr; float2* tw_inv = nullptr; uint32_t row_mask[4] = {0, 0, 0, 0}; };
    std::map<long long, WideSet> wide_tables;      // per (filter version, d, tunebin)
    Slot slots[kSlots];
    uint64_t next_ticket = 1;                      // submissions are numbered; ticket t uses slot t % kSlots
    std::condition_variable cv;                    // a held slot was released / the operator was switched off
    int16_t* carry = nullptr;                      // last halfFft samples of the stream (host copy, updated at submit)
    bool carry_valid = false;
    cudaStream_t util_stream = nullptr;

    NcoState nco;

    r2iq_cuda_stats stats{};
};

namespace {

void poison(r2iq_cuda_t h, const char* why) {
    if (!h || h->poisoned) return;
    h->poisoned = true;
    snprintf(h->poison_msg, sizeof(h->poison_msg), "%s", why);
}

#define CU(h, call)                                                                                       \
    do {                                                                                                  \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess) {                                                                         \
            const int rc__ = fail(R2IQ_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            poison(h, g_err);                                                                             \
            return rc__;                                                                                  \
        }                                                                                                 \
    } while (0)

int check_handle(r2iq_cuda_t h) {
    if (!h) return fail(R2IQ_E_INVALID, "null handle");
    if (h->poisoned) return fail(R2IQ_E_CUDA, "handle is unusable after an earlier CUDA error: %s", h->poison_msg);
    return R2IQ_OK;
}

// Frees device allocations when an entry point leaves early (the CU() macro returns from the middle of a function).
struct DeviceScratch {
    void* ptrs[4] = {nullptr, nullptr, nullptr, nullptr};
    int n = 0;
    template <typename T>
    cudaError_t alloc(T** p, size_t bytes) {
        cudaError_t e = cudaMalloc((void**)p, bytes);
        if (e == cudaSuccess) ptrs[n++] = (void*)*p;
        return e;
    }
    ~DeviceScratch() {
        for (int i = 0; i < n; i++) cudaFree(ptrs[i]);
    }
};

// Makes the handle's device current for the duration of an entry point and puts the caller's device back afterwards.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != device) {
            err = cudaSetDevice(device);
            switched = err == cudaSuccess;
        }
    }
    ~DeviceGuard() {
        if (switched && prev >= 0) cudaSetDevice(prev);
    }
};

// Twiddles of the R = M/32 register scheme for an M-point backward transform on 32 lanes (r2iq_dual_dn.cu,
// r2iq_wide.cu): [R][32] exp(+2 pi i lane k2 / M), then [32][32] exp(+2 pi i (lane % G) n1 / 32), G = 32/R.
// folded: the pass-2 weights carry the inter-pass twiddle exp(+2 pi i n1 (lane / G) / M) of the element they multiply
// (row n1, column lane / G of the transposition), so the kernels skip the twiddle multiplication before it
std::vector<float2> build_tw_aux(int M, bool folded = false) {
    const int R = M / 32, G = 32 / R;
    std::vector<float2> ta((size_t)(R + 32) * 32);
    for (int k2 = 0; k2 < R; k2++)
        for (int l = 0; l < 32; l++) {
            const double th = 2.0 * kPiD * (double)((l * k2) % M) / (double)M;
            ta[k2 * 32 + l] = make_float2((float)cos(th), (float)sin(th));
        }
    for (int n1 = 0; n1 < 32; n1++)
        for (int l = 0; l < 32; l++) {
            double th = 2.0 * kPiD * (double)(((l % G) * n1) % 32) / 32.0;
            if (folded) th += 2.0 * kPiD * (double)((n1 * (l / G)) % M) / (double)M;
            ta[(R + n1) * 32 + l] = make_float2((float)cos(th), (float)sin(th));
        }
    return ta;
}

void free_channel_tables(r2iq_cuda_t h) {
    for (auto& kv : h->gh_tables) cudaFree(kv.second);
    for (auto& kv : h->chan_lists) cudaFree(kv.second);
    h->gh_tables.clear();
    h->chan_lists.clear();
}

void free_tables(r2iq_cuda_t h) {
    for (auto& kv : h->tables) {
        cudaFree(kv.second.gk);
        cudaFree(kv.second.gck);
        cudaFree(kv.second.gp);
        cudaFree(kv.second.gq);
        cudaFree(kv.second.pq_band);
        cudaFree(kv.second.tw_inv);
        cudaFree(kv.second.tw_aux);
    }
    h->tables.clear();
}

long long table_key(r2iq_cuda_t h, int d, int tunebin) {
    return ((long long)h->hw_version[d] << 16) | ((long long)d << 12) | (long long)tunebin;
}

// Build (or fetch) the per-(d, tunebin) tables:
//   gk[k]   = 0.5 * H_d[(k - tunebin) mod halfFft]  for bins k inside the band, else 0
//             (band = tunebin + [-mfft/2, mfft/2) clipped to [0, halfFft): impl.hpp:76-96)
//   gck[k]  = (-i * exp(-2 pi i k / FFTN)) * gk[k]   (the odd-part twiddle of the real-FFT split)
//   tw_inv  = inter-pass twiddles of the backward transform (see r2iq_kernels.cu)
int get_tables(r2iq_cuda_t h, int d, int tunebin, TableSet* out) {
    const long long key = table_key(h, d, tunebin);
    auto it = h->tables.find(key);
    if (it != h->tables.end()) {
        *out = it->second;
        return R2IQ_OK;
    }
    if (h->tables.size() >= kMaxTableSets) {
        CU(h, cudaDeviceSynchronize());   // nobody may still be reading the tables we are about to free
        free_tables(h);
    }
    const int M = kHalfFft >> d;
    std::vector<float2> gk(kHalfFft), gck(kHalfFft), gp(kHalfFft), gq(kHalfFft);
    const std::vector<std::complex<float>>& H = h->hw[d];
    TableSet ts;
    ts.band_lo = tunebin - M / 2 < 0 ? 0 : tunebin - M / 2;
    ts.band_hi = tunebin + M / 2 > kHalfFft ? kHalfFft : tunebin + M / 2;
    ts.row_mask = 0;   // rows the band cut of the dual-frame kernels reads (wrapped bins and mirrors included)
    for (int s = -M / 2; s < M / 2; s++) {
        const int kk = (tunebin + s) & (kHalfFft - 1);
        ts.row_mask |= 1u << (kk >> 5);
        ts.row_mask |= 1u << (((kHalfFft - kk) & (kHalfFft - 1)) >> 5);
    }
    for (int k = 0; k < kHalfFft; k++) {
        const int s = k - tunebin;
        if (s >= -M / 2 && s < M / 2) {
            const std::complex<double> hh = std::complex<double>(H[(s + kHalfFft) % kHalfFft]) * 0.5;
            const double th = 2.0 * kPiD * (double)k / (double)kFftN;
            const std::complex<double> c(-sin(th), -cos(th));   // -i * exp(-i th)
            const std::complex<double> gc = c * hh;
            gk[k] = make_float2((float)hh.real(), (float)hh.imag());
            gck[k] = make_float2((float)gc.real(), (float)gc.imag());
            gp[k] = make_float2((float)(hh.real() + gc.real()), (float)(hh.imag() + gc.imag()));
            gq[k] = make_float2((float)(hh.real() - gc.real()), (float)(hh.imag() - gc.imag()));
        } else {
            gk[k] = gck[k] = gp[k] = gq[k] = make_float2(0.f, 0.f);
        }
    }
    std::vector<float2> tw(32 * 32);
    const bool pair_tables = kernel_generation() == 0;   // the shipped kernels (r2iq_pair.cu); comparison builds keep the older forms
    if (d == 0 && pair_tables) {
        // The twiddles are applied on the READING side of the backward transposition (dit2_tw): lane k2 reads row n, which
        // the lane l = (n + tunebin) mod 32 wrote -- the lane that plays residue n1' = (l - tunebin) mod 1024.
        for (int n = 0; n < 32; n++)
            for (int k2 = 0; k2 < 32; k2++) {
                const int l = (n + tunebin) & 31;
                const long long n1 = ((l - tunebin) % kHalfFft + kHalfFft) % kHalfFft;
                const long long e = (n1 * k2) % kHalfFft;
                const double th = 2.0 * kPiD * (double)e / (double)kHalfFft;
                tw[n * 32 + k2] = make_float2((float)cos(th), (float)sin(th));
            }
    } else if (d == 0) {
        // lane l plays residue n1' = (l - tunebin) mod 1024: twiddle exp(+2 pi i n1' k2 / 1024)
        for (int k2 = 0; k2 < 32; k2++)
            for (int l = 0; l < 32; l++) {
                const long long n1 = ((l - tunebin) % kHalfFft + kHalfFft) % kHalfFft;
                const long long e = (n1 * k2) % kHalfFft;
                const double th = 2.0 * kPiD * (double)e / (double)kHalfFft;
                tw[k2 * 32 + l] = make_float2((float)cos(th), (float)sin(th));
            }
    } else {
        const int L = d <= 5 ? (32 >> d) : 1;
        for (int k2 = 0; k2 < 32; k2++)
            for (int n1 = 0; n1 < L; n1++) {
                const double th = 2.0 * kPiD * (double)((n1 * k2) % M) / (double)M;
                tw[k2 * L + n1] = make_float2((float)cos(th), (float)sin(th));
            }
    }
    if (d >= 5 && pair_tables) {
        // r2iq_pair.cu, d = 5: lane c evaluates the band bin and the mirror bin of spectrum column c directly from the
        // first forward pass: coefficients exp(-2 pi i n1 k / 1024), [c][n1] as (band.re, band.im, mirror.re, mirror.im)
        std::vector<float4> zc(32 * 32);
        for (int c = 0; c < 32; c++) {
            // the band offset s in [-M/2, M/2) with tunebin + s = c (band bin) resp. = -c (mirror bin) modulo 32; at
            // d = 6 (M = 16) half of the columns hold none: zero coefficients, the value is never fetched
            auto offset = [&](int u, int* s_out) {
                u &= 31;
                if (u < M / 2) { *s_out = u; return true; }
                if (u >= 32 - M / 2) { *s_out = u - 32; return true; }
                return false;
            };
            int sb = 0, sm = 0;
            const bool has_b = offset(c - tunebin, &sb), has_m = offset(-c - tunebin, &sm);
            const int kb = (tunebin + sb) & (kHalfFft - 1);
            const int km = (kHalfFft - ((tunebin + sm) & (kHalfFft - 1))) & (kHalfFft - 1);
            for (int n1 = 0; n1 < 32; n1++) {
                const double tb_ = -2.0 * kPiD * (double)((n1 * kb) % kHalfFft) / (double)kHalfFft;
                const double tm_ = -2.0 * kPiD * (double)((n1 * km) % kHalfFft) / (double)kHalfFft;
                zc[c * 32 + n1] = make_float4(has_b ? (float)cos(tb_) : 0.f, has_b ? (float)sin(tb_) : 0.f,
                                              has_m ? (float)cos(tm_) : 0.f, has_m ? (float)sin(tm_) : 0.f);
            }
        }
        CU(h, cudaMalloc(&ts.tw_aux, sizeof(float4) * zc.size()));
        CU(h, cudaMemcpy(ts.tw_aux, zc.data(), sizeof(float4) * zc.size(), cudaMemcpyHostToDevice));
    } else if (d >= 1 && d <= 5) {
        // dual-frame kernels, backward transform mfft = 32*R on 32 lanes: pass-1 twiddles exp(+2 pi i lane k2 / mfft)
        // (k2 < R), then the pass-2 weights exp(+2 pi i (lane % G) n1 / 32), G = 32/R
        const std::vector<float2> ta = build_tw_aux(M, pair_tables);
        CU(h, cudaMalloc(&ts.tw_aux, sizeof(float2) * ta.size()));
        CU(h, cudaMemcpy(ts.tw_aux, ta.data(), sizeof(float2) * ta.size(), cudaMemcpyHostToDevice));
    }
    if (d >= 1 && pair_tables) {
        // the band cut reads (P, Q) in the order of the backward transform: position m = lane + 32 j holds bin
        // tunebin + (m < mfft/2 ? m : m - mfft) -- one coalesced 16-byte load per bin and no index arithmetic
        std::vector<float4> pqb(M);
        for (int m = 0; m < M; m++) {
            const int kk = (tunebin + (m < M / 2 ? m : m - M)) & (kHalfFft - 1);
            pqb[m] = make_float4(gp[kk].x, gp[kk].y, gq[kk].x, gq[kk].y);
        }
        CU(h, cudaMalloc(&ts.pq_band, sizeof(float4) * M));
        CU(h, cudaMemcpy(ts.pq_band, pqb.data(), sizeof(float4) * M, cudaMemcpyHostToDevice));
    }
    CU(h, cudaMalloc(&ts.gk, sizeof(float2) * kHalfFft));
    CU(h, cudaMalloc(&ts.gck, sizeof(float2) * kHalfFft));
    CU(h, cudaMalloc(&ts.gp, sizeof(float2) * kHalfFft));
    CU(h, cudaMalloc(&ts.gq, sizeof(float2) * kHalfFft));
    CU(h, cudaMalloc(&ts.tw_inv, sizeof(float2) * 32 * 32));
    CU(h, cudaMemcpy(ts.gk, gk.data(), sizeof(float2) * kHalfFft, cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(ts.gck, gck.data(), sizeof(float2) * kHalfFft, cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(ts.gp, gp.data(), sizeof(float2) * kHalfFft, cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(ts.gq, gq.data(), sizeof(float2) * kHalfFft, cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(ts.tw_inv, tw.data(), sizeof(float2) * 32 * 32, cudaMemcpyHostToDevice));
    h->tables[key] = ts;
    *out = ts;
    return R2IQ_OK;
}

void nco_init(NcoState& s, float relative_freq) {
    const float PI_F = 3.14159265358979323846f;             // PI as the reference's float arithmetic sees it
    const float inc = 2 * relative_freq * PI_F;
    float myphase = 0.0F;
    for (int g = 0; g < 33; g++) {
        for (int k = 0; k < 4; k++) {
            myphase += inc;
            while (myphase > PI_F) myphase -= 2 * PI_F;
            while (myphase < -PI_F) myphase += 2 * PI_F;
        }
        s.tab[g][0] = cosf(myphase);
        s.tab[g][1] = sinf(myphase);
    }
    myphase = 0.0F;                                          // phase_start_rad = 0 (RadioHandler.cpp:295)
    for (int i = 0; i < 4; i++) {
        s.cs[i] = cosf(myphase);
        s.sn[i] = sinf(myphase);
        myphase += inc;
        while (myphase > PI_F) myphase -= 2 * PI_F;
        while (myphase < -PI_F) myphase += 2 * PI_F;
    }
    s.fc = relative_freq;
}

// Start phasors of the next n128 blocks of 128 samples into s.h_starts ([block][lane] (cos, sin)); advances the state.
// This recurrence is inherently sequential (every step renormalises the previous one) and is what bounds the mixer
// in bulk use: ~2.1 M steps per 8192-block launch at d = 0.  The four lanes are the reference's own SSE lanes
// (Core/pffft/pf_mixer.cpp:791-856), so the x86 path runs them as one __m128 chain -- mulps/subps/addps/sqrtps/divps
// are correctly rounded and never contracted, i.e. the same bits as the scalar form below.
void nco_advance(NcoState& s, size_t n128) {
    s.h_starts.resize(n128 * 8);
#if defined(__SSE2__)
    const __m128 tc = _mm_set1_ps(s.tab[31][0]), ts = _mm_set1_ps(s.tab[31][1]);   // after 32 groups vals = table[31] * starts
    __m128 cs = _mm_loadu_ps(s.cs), sn = _mm_loadu_ps(s.sn);
    float* dst = s.h_starts.data();
    for (size_t b = 0; b < n128; b++, dst += 8) {
        _mm_storeu_ps(dst, _mm_unpacklo_ps(cs, sn));        // (c0, s0, c1, s1)
        _mm_storeu_ps(dst + 4, _mm_unpackhi_ps(cs, sn));    // (c2, s2, c3, s3)
        const __m128 cv = _mm_sub_ps(_mm_mul_ps(tc, cs), _mm_mul_ps(ts, sn));
        const __m128 sv = _mm_add_ps(_mm_mul_ps(ts, cs), _mm_mul_ps(tc, sn));
        const __m128 mag = _mm_sqrt_ps(_mm_add_ps(_mm_mul_ps(cv, cv), _mm_mul_ps(sv, sv)));
        cs = _mm_div_ps(cv, mag);
        sn = _mm_div_ps(sv, mag);
    }
    _mm_storeu_ps(s.cs, cs);
    _mm_storeu_ps(s.sn, sn);
#else
    const volatile float tc = s.tab[31][0], ts = s.tab[31][1];   // after 32 groups vals = table[31] * starts
    for (size_t b = 0; b < n128; b++) {
        for (int j = 0; j < 4; j++) {
            s.h_starts[b * 8 + 2 * j] = s.cs[j];
            s.h_starts[b * 8 + 2 * j + 1] = s.sn[j];
            // volatile temporaries keep every product separately rounded (no contraction, as the SSE code)
            volatile float a = tc * s.cs[j], bb = ts * s.sn[j], cc = ts * s.cs[j], dd = tc * s.sn[j];
            volatile float cv = a - bb, sv = cc + dd;
            volatile float m1 = cv * cv, m2 = sv * sv;
            volatile float mag = sqrtf(m1 + m2);
            s.cs[j] = cv / mag;
            s.sn[j] = sv / mag;
        }
    }
#endif
}

// Fine-tune NCO for the n_out complex samples of one launch: nco_prepare advances the phasor recurrence on the host
// and uploads the start phasors of every 128-sample block (on the launch stream, BEFORE the kernel, so that a kernel
// with the mixer in its store epilogue can read them); nco_stage is prepare + the stand-alone in-place pass for the
// kernels that do not mix themselves.
int nco_prepare(r2iq_cuda_t h, size_t n_out, float** d_starts, size_t* cap, cudaStream_t stream) {
    NcoState& s = h->nco;
    if (n_out % 128) return fail(R2IQ_E_INVALID, "fine-tune mixer needs a multiple of 128 output samples per call");
    const size_t n128 = n_out / 128;
    if (*cap < n128) {
        if (*d_starts) {
            CU(h, cudaStreamSynchronize(stream));
            cudaFree(*d_starts);
        }
        CU(h, cudaMalloc(d_starts, sizeof(float) * 8 * n128));
        *cap = n128;
    }
    nco_advance(s, n128);
    // pageable source: the call returns once the data sits in the driver's staging buffer, so h_starts may be reused
    CU(h, cudaMemcpyAsync(*d_starts, s.h_starts.data(), sizeof(float) * 8 * n128, cudaMemcpyHostToDevice, stream));
    return R2IQ_OK;
}
int nco_stage(r2iq_cuda_t h, float* d_out, size_t n_out, float** d_starts, size_t* cap, cudaStream_t stream) {
    NcoState& s = h->nco;
    if (!s.on || n_out == 0) return R2IQ_OK;
    int rc = nco_prepare(h, n_out, d_starts, cap, stream);
    if (rc != R2IQ_OK) return rc;
    CU(h, launch_nco_mix(d_out, (long long)n_out, *d_starts, s.d_tab, stream));
    return R2IQ_OK;
}
// Direct-phase mode: the phase moves on by fc * n cycles per launch (float64; fc is a float32 value, n < 2^31).
void nco_direct_advance(NcoState& s, size_t n_out) {
    double ph = s.phase + (double)s.fc * (double)n_out;
    s.phase = ph - floor(ph);
}

// FFTN 8192: the tables that depend on nothing but the size, once per handle
int ensure_wide_static_tables(r2iq_cuda_t h) {
    if (h->d_twq) return R2IQ_OK;
    std::vector<float2> tq(3 * 1024), ct(2049), cf(4096);
    for (int q = 1; q <= 3; q++)
        for (int n = 0; n < 1024; n++) {
            const double th = -2.0 * kPiD * (double)((q * n) % 4096) / 4096.0;
            tq[(q - 1) * 1024 + n] = make_float2((float)cos(th), (float)sin(th));
        }
    for (int k = 0; k < 4096; k++) {
        const double th = 2.0 * kPiD * (double)k / 8192.0;
        cf[k] = make_float2((float)-sin(th), (float)-cos(th));          // -i * exp(-i th)
        if (k <= 2048) ct[k] = cf[k];
    }
    CU(h, cudaMalloc(&h->d_twq, sizeof(float2) * tq.size()));
    CU(h, cudaMalloc(&h->d_wctab, sizeof(float2) * ct.size()));
    CU(h, cudaMalloc(&h->d_wctab_full, sizeof(float2) * cf.size()));
    CU(h, cudaMemcpy(h->d_twq, tq.data(), sizeof(float2) * tq.size(), cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(h->d_wctab, ct.data(), sizeof(float2) * ct.size(), cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(h->d_wctab_full, cf.data(), sizeof(float2) * cf.size(), cudaMemcpyHostToDevice));
    return R2IQ_OK;
}

// FFTN 8192: static twiddles once per handle, filter gains (and small-transform twiddles) per (d, tune bin).
// The kernels index frames with 32 bits (dual_frame.cuh: frame_src): 2^31 frames is 6.5 TB of samples at FFTN 2048.
constexpr long kMaxBlocksPerLaunch = 0x7fffffffL / 43 - 1;

int launch_wide_impl(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_out, cudaStream_t stream, int tunebin,
                     bool direct_mixer = false) {
    if (n_blocks > kMaxBlocksPerLaunch) return fail(R2IQ_E_INVALID, "%ld blocks in one launch (at most %ld)", n_blocks, kMaxBlocksPerLaunch);
    const int d = h->latched_decimate;
    const int HF = 4096, M = HF >> d;
    int rc0 = ensure_wide_static_tables(h);
    if (rc0 != R2IQ_OK) return rc0;
    const long long key = ((long long)h->hw_version[d] << 20) | ((long long)d << 16) | (long long)tunebin;
    r2iq_cuda_s::WideSet ws;
    auto it = h->wide_tables.find(key);
    if (it != h->wide_tables.end()) ws = it->second;
    else {
        if (h->wide_tables.size() >= kMaxTableSets) {
            CU(h, cudaDeviceSynchronize());
            for (auto& kv : h->wide_tables) {
                cudaFree(kv.second.g);
                cudaFree(kv.second.tw_sub);
                cudaFree(kv.second.pq);
                cudaFree(kv.second.tw_inv);
            }
            h->wide_tables.clear();
        }
        std::vector<float2> g(HF);
        for (int k = 0; k < HF; k++) {
            const int s = k - tunebin;
            if (s >= -M / 2 && s < M / 2) {
                const std::complex<float> hh = h->hw[d][(s + HF) % HF];
                g[k] = make_float2(0.5f * hh.real(), 0.5f * hh.imag());
            } else {
                g[k] = make_float2(0.f, 0.f);
            }
        }
        CU(h, cudaMalloc(&ws.g, sizeof(float2) * HF));
        CU(h, cudaMemcpy(ws.g, g.data(), sizeof(float2) * HF, cudaMemcpyHostToDevice));
        // (P, Q) of the packed kernels, plane-major (bin 4 m + r at [r][m]), and the rows of each plane the band cut reads
        std::vector<float4> pq(HF);
        for (int k = 0; k < HF; k++) {
            const std::complex<double> hh(g[k].x, g[k].y);
            const double th = 2.0 * kPiD * (double)k / 8192.0;
            const std::complex<double> gc = std::complex<double>(-sin(th), -cos(th)) * hh;   // (-i exp(-i th)) g
            pq[(size_t)(k & 3) * 1024 + (k >> 2)] = make_float4((float)(hh.real() + gc.real()), (float)(hh.imag() + gc.imag()),
                                                                (float)(hh.real() - gc.real()), (float)(hh.imag() - gc.imag()));
        }
        for (int s = -M / 2; s < M / 2; s++) {
            const int kk = (tunebin + s) & (HF - 1), km = (HF - kk) & (HF - 1);
            ws.row_mask[kk & 3] |= 1u << (kk >> 7);
            ws.row_mask[km & 3] |= 1u << (km >> 7);
        }
        CU(h, cudaMalloc(&ws.pq, sizeof(float4) * HF));
        CU(h, cudaMemcpy(ws.pq, pq.data(), sizeof(float4) * HF, cudaMemcpyHostToDevice));
        if (d == 0) {
            // lane l of the backward 1024-point transforms plays residue (l - tunebin/4) mod 32 (r2iq_wide_pair.cu, D1)
            std::vector<float2> ti(1024);
            const bool reader_side = kernel_generation() == 0;   // the packed kernels apply them after the transposition
            for (int k2 = 0; k2 < 32; k2++)
                for (int l = 0; l < 32; l++) {
                    const long long n1 = ((l - tunebin / 4) % 1024 + 1024) % 1024;
                    const double th = 2.0 * kPiD * (double)((n1 * k2) % 1024) / 1024.0;
                    // writer lane l stores row (l - tunebin/4) mod 32; reader lane k2 finds it there
                    if (reader_side) ti[(int)(n1 & 31) * 32 + k2] = make_float2((float)cos(th), (float)sin(th));
                    else ti[k2 * 32 + l] = make_float2((float)cos(th), (float)sin(th));
                }
            CU(h, cudaMalloc(&ws.tw_inv, sizeof(float2) * ti.size()));
            CU(h, cudaMemcpy(ws.tw_inv, ti.data(), sizeof(float2) * ti.size(), cudaMemcpyHostToDevice));
        }
        if (d >= 1) {
            const std::vector<float2> ta = build_tw_aux(d == 6 ? M : M / 4, kernel_generation() == 0);
            CU(h, cudaMalloc(&ws.tw_sub, sizeof(float2) * ta.size()));
            CU(h, cudaMemcpy(ws.tw_sub, ta.data(), sizeof(float2) * ta.size(), cudaMemcpyHostToDevice));
        }
        h->wide_tables[key] = ws;
    }
    WideParams p;
    p.in_words = (const uint32_t*)d_in;
    p.out = (float2*)d_out;
    p.total_frames = (long long)n_blocks * 11;
    p.tw1024 = h->d_tw_fwd;
    p.twq = h->d_twq;
    p.ctab = h->d_wctab;
    p.g = ws.g;
    p.tw_sub = ws.tw_sub;
    p.pq = ws.pq;
    p.tw_inv = ws.tw_inv;
    for (int r = 0; r < 4; r++) p.row_mask[r] = ws.row_mask[r];
    p.tunebin = tunebin;
    p.rand_mul = h->rand_on ? 0xFFFEu : 0u;
    p.imag_sign = h->latched_sideband ? -1.0f : 1.0f;
    p.work_counter = h->next_counter();
    WideNco nco{nullptr, 0.0, 0.0};
    if (direct_mixer) {                   // mixed in the store epilogue, as the 2048 family does
        nco.nco_rot = (const float2*)h->nco.d_rot;
        nco.nco_phase0 = h->nco.phase;
        nco.nco_fc = (double)h->nco.fc;
    }
    CU(h, launch_wide(d, p, nco, h->sm_count, stream));
    return R2IQ_OK;
}

// Parameter block of one single-channel launch at `tunebin` (tables from the per-(d, tune bin) cache), mixer off.
int fused_params(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_out, int tunebin, FusedParams* out) {
    if (n_blocks > kMaxBlocksPerLaunch) return fail(R2IQ_E_INVALID, "%ld blocks in one launch (at most %ld)", n_blocks, kMaxBlocksPerLaunch);
    TableSet ts;
    const int d = h->latched_decimate;
    int rc = get_tables(h, d, tunebin, &ts);
    if (rc != R2IQ_OK) return rc;
    FusedParams& p = *out;
    p.in_words = (const uint32_t*)d_in;
    p.out = (float2*)d_out;
    p.total_frames = (long long)n_blocks * kFramesPerBlock;
    p.tw_fwd = h->d_tw_fwd;
    p.gk = ts.gk;
    p.gck = ts.gck;
    p.gp = ts.gp;
    p.gq = ts.gq;
    if (ts.pq_band) {                     // r2iq_pair.cu, d >= 1: `gp` carries the band-ordered (P, Q) table, `gq` is unused
        p.gp = reinterpret_cast<const float2*>(ts.pq_band);
        p.gq = nullptr;
    }
    p.tw_inv = ts.tw_inv;
    p.tw_aux = ts.tw_aux;
    p.tunebin = tunebin;
    p.band_lo = ts.band_lo;
    p.band_hi = ts.band_hi;
    p.row_mask = ts.row_mask;
    p.rand_mul = h->rand_on ? 0xFFFEu : 0u;
    p.imag_sign = h->latched_sideband ? -1.0f : 1.0f;
    p.work_counter = h->next_counter();
    p.nco_starts = nullptr;
    p.nco_tab = nullptr;
    p.nco_rot = nullptr;
    p.nco_phase0 = 0.0;
    p.nco_fc = 0.0;
    return R2IQ_OK;
}

int launch(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_out, cudaStream_t stream, float** nco_starts,
           size_t* nco_cap) {
    if (h->fftn == 8192) {
        const size_t n_out8 = (size_t)n_blocks * (size_t)((kTransferSamples / 2) >> h->latched_decimate);
        const bool direct8 = h->nco.on && n_out8 > 0 && h->nco.mode == R2IQ_NCO_DIRECT;
        int rc = launch_wide_impl(h, d_in, n_blocks, d_out, stream, h->tunebin.load(), direct8);   // one snapshot per launch
        if (rc != R2IQ_OK) return rc;
        h->stats.kernel_launches++;
        h->stats.blocks_processed += (uint64_t)n_blocks;
        if (direct8) {
            nco_direct_advance(h->nco, n_out8);
            return R2IQ_OK;
        }
        return nco_stage(h, d_out, n_out8, nco_starts, nco_cap, stream);
    }
    const int d = h->latched_decimate;
    const int tunebin = h->tunebin.load();   // one snapshot per launch: a retune from the application thread between the
                                             // table lookup and the parameter block would pair tables with another bin
    FusedParams p;
    int rc = fused_params(h, d_in, n_blocks, d_out, tunebin, &p);
    if (rc != R2IQ_OK) return rc;
    const size_t n_out = (size_t)n_blocks * (size_t)((kTransferSamples / 2) >> d);
    const bool direct = h->nco.on && n_out > 0 && h->nco.mode == R2IQ_NCO_DIRECT;
    if (direct) {                         // the default kernels mix in their store epilogue: no second pass, no host work
        p.nco_rot = (const float2*)h->nco.d_rot;
        p.nco_phase0 = h->nco.phase;
        p.nco_fc = (double)h->nco.fc;
        nco_direct_advance(h->nco, n_out);
    }
    CU(h, launch_fused(d, p, h->sm_count, stream));
    h->stats.kernel_launches++;
    h->stats.blocks_processed += (uint64_t)n_blocks;
    if (direct) return R2IQ_OK;
    return nco_stage(h, d_out, n_out, nco_starts, nco_cap, stream);
}

// Channelizer launch: per-d filter table in inverse-FFT bin order + the tune-bin list on the device.
int launch_channels_impl(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, const int* tunebins, int n_channels,
                         float* d_out, cudaStream_t stream) {
    if (n_blocks > kMaxBlocksPerLaunch) return fail(R2IQ_E_INVALID, "%ld blocks in one launch (at most %ld)", n_blocks, kMaxBlocksPerLaunch);
    const int d = h->latched_decimate;
    const int half = h->half;
    for (int c = 0; c < n_channels; c++)
        if (tunebins[c] < 0 || tunebins[c] >= half || (tunebins[c] & 3))
            return fail(R2IQ_E_INVALID, "channel %d: tune bin %d must be a multiple of 4 in [0,%d)", c, tunebins[c], half);
    if (h->fftn == 8192 && (d == 4 || d == 5) && n_channels >= 2) {
        // FFTN 8192, d = 4, 5: one forward transform and split per frame pair for all channels (r2iq_wide_channels.cu)
        int rc = ensure_wide_static_tables(h);
        if (rc != R2IQ_OK) return rc;
        const int HF = 4096, M = HF >> d;
        const long long gkey = ((long long)h->hw_version[d] << 4) | d | (1LL << 40);
        float2* gh = nullptr;
        auto git = h->gh_tables.find(gkey);
        if (git != h->gh_tables.end()) gh = git->second;
        else {
            std::vector<float2> g(M);
            for (int m = 0; m < M; m++) {
                const int s = m < M / 2 ? m : m - M;
                const std::complex<float> hh = h->hw[d][(s + HF) % HF];
                g[m] = make_float2(0.5f * hh.real(), 0.5f * hh.imag());
            }
            CU(h, cudaMalloc(&gh, sizeof(float2) * M));
            CU(h, cudaMemcpy(gh, g.data(), sizeof(float2) * M, cudaMemcpyHostToDevice));
            h->gh_tables[gkey] = gh;
        }
        std::vector<int> key(tunebins, tunebins + n_channels);
        int* d_list = nullptr;
        auto lit = h->chan_lists.find(key);
        if (lit != h->chan_lists.end()) d_list = lit->second;
        else {
            if (h->chan_lists.size() >= 256) {
                CU(h, cudaDeviceSynchronize());
                for (auto& kv : h->chan_lists) cudaFree(kv.second);
                h->chan_lists.clear();
            }
            CU(h, cudaMalloc(&d_list, sizeof(int) * n_channels));
            CU(h, cudaMemcpy(d_list, tunebins, sizeof(int) * n_channels, cudaMemcpyHostToDevice));
            h->chan_lists[key] = d_list;
        }
        WideChannelParams p;
        p.in_words = (const uint32_t*)d_in;
        p.out = (float2*)d_out;
        p.out_stride = (long long)n_blocks * ((kTransferSamples / 2) >> d);
        p.total_frames = (long long)n_blocks * 11;
        p.tw1024 = h->d_tw_fwd;
        p.twq = h->d_twq;
        p.ctab = h->d_wctab_full;
        p.gh = gh;
        p.tunebins = d_list;
        p.n_channels = n_channels;
        p.rand_mul = h->rand_on ? 0xFFFEu : 0u;
        p.imag_sign = h->latched_sideband ? -1.0f : 1.0f;
        p.work_counter = h->next_counter();
        CU(h, launch_wide_channels(d, p, h->sm_count, stream));
        h->stats.kernel_launches++;
        h->stats.blocks_processed += (uint64_t)n_blocks;
        return R2IQ_OK;
    }
    if (h->fftn != kFftN || d <= 1) {
        // FFTN 8192, and d = 0, 1 (every channel needs a backward transform of the size of the forward one or half of it:
        // sharing the forward transform saves little, and the single-channel kernels are the faster code -- 64 channels,
        // 2048 blocks at d = 1: 12.9 ms against 17.8 ms of the shared-forward kernel): one launch of the single-channel
        // kernel per channel, the way the reference runs one operator per tuned instance
        // (Core/fft_mt_r2iq_impl.hpp:76-98); same result, no shared forward transform
        const size_t per_ch = (size_t)n_blocks * (size_t)((kTransferSamples / 2) >> d);
        for (int c = 0; c < n_channels; c++) {
            float* dst = d_out + 2 * per_ch * (size_t)c;
            int rc;
            if (h->fftn == 8192) rc = launch_wide_impl(h, d_in, n_blocks, dst, stream, tunebins[c]);
            else {
                FusedParams p;
                rc = fused_params(h, d_in, n_blocks, dst, tunebins[c], &p);
                if (rc == R2IQ_OK) CU(h, launch_fused(d, p, h->sm_count, stream));
            }
            if (rc != R2IQ_OK) return rc;
            h->stats.kernel_launches++;
        }
        h->stats.blocks_processed += (uint64_t)n_blocks;
        return R2IQ_OK;
    }
    const int M = kHalfFft >> d;
    TableSet ts;   // for tw_inv (independent of the tune bin when d >= 1)
    int rc = get_tables(h, d, 0, &ts);
    if (rc != R2IQ_OK) return rc;
    if (!h->d_ctab) {
        std::vector<float2> c(kHalfFft);
        for (int k = 0; k < kHalfFft; k++) {
            const double th = 2.0 * kPiD * (double)k / (double)kFftN;
            c[k] = make_float2((float)-sin(th), (float)-cos(th));
        }
        CU(h, cudaMalloc(&h->d_ctab, sizeof(float2) * kHalfFft));
        CU(h, cudaMemcpy(h->d_ctab, c.data(), sizeof(float2) * kHalfFft, cudaMemcpyHostToDevice));
    }
    const long long gkey = ((long long)h->hw_version[d] << 4) | d;
    float2* gh = nullptr;
    auto git = h->gh_tables.find(gkey);
    if (git != h->gh_tables.end()) gh = git->second;
    else {
        std::vector<float2> g(M);
        for (int m = 0; m < M; m++) {
            const int s = m < M / 2 ? m : m - M;
            const std::complex<float> hh = h->hw[d][(s + kHalfFft) % kHalfFft];
            g[m] = make_float2(0.5f * hh.real(), 0.5f * hh.imag());
        }
        CU(h, cudaMalloc(&gh, sizeof(float2) * M));
        CU(h, cudaMemcpy(gh, g.data(), sizeof(float2) * M, cudaMemcpyHostToDevice));
        h->gh_tables[gkey] = gh;
    }
    std::vector<int> key(tunebins, tunebins + n_channels);
    int* d_list = nullptr;
    auto lit = h->chan_lists.find(key);
    if (lit != h->chan_lists.end()) d_list = lit->second;
    else {
        if (h->chan_lists.size() >= 256) {
            CU(h, cudaDeviceSynchronize());
            for (auto& kv : h->chan_lists) cudaFree(kv.second);
            h->chan_lists.clear();
        }
        CU(h, cudaMalloc(&d_list, sizeof(int) * n_channels));
        CU(h, cudaMemcpy(d_list, tunebins, sizeof(int) * n_channels, cudaMemcpyHostToDevice));
        h->chan_lists[key] = d_list;
    }
    ChannelParams p;
    p.in_words = (const uint32_t*)d_in;
    p.out = (float2*)d_out;
    p.out_stride = (long long)n_blocks * ((kTransferSamples / 2) >> d);
    p.total_frames = (long long)n_blocks * kFramesPerBlock;
    p.tw_fwd = h->d_tw_fwd;
    p.ctab = h->d_ctab;
    p.gh = gh;
    p.tw_inv = ts.tw_inv;
    p.tunebins = d_list;
    p.n_channels = n_channels;
    p.rand_mul = h->rand_on ? 0xFFFEu : 0u;
    p.imag_sign = h->latched_sideband ? -1.0f : 1.0f;
    p.work_counter = h->next_counter();
    CU(h, launch_channels(d, p, h->sm_count, stream));
    h->stats.kernel_launches++;
    h->stats.blocks_processed += (uint64_t)n_blocks;
    return R2IQ_OK;
}

int ensure_slots(r2iq_cuda_t h) {
    for (int i = 0; i < kSlots; i++) {
        Slot& s = h->slots[i];
        if (s.d_in) continue;
        CU(h, cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CU(h, cudaEventCreateWithFlags(&s.h2d_done, cudaEventDisableTiming));
        CU(h, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        CU(h, cudaMallocHost(&s.carry, sizeof(int16_t) * (size_t)h->half));
        CU(h, cudaMalloc(&s.d_in, sizeof(int16_t) * ((size_t)h->half + (size_t)h->chunk_blocks * kTransferSamples)));
        CU(h, cudaMalloc(&s.d_out, sizeof(float) * 2 * (size_t)h->chunk_blocks * (kTransferSamples / 2)));
    }
    return R2IQ_OK;
}

}  // namespace

extern "C" {

const char* r2iq_cuda_last_error(void) { return g_err; }
#ifdef R2IQ_AB_VARIANTS
const char* r2iq_cuda_version(void) { return "r2iq-b200 0.2 sm_100a fftn=2048,8192 ab-variants"; }
#else
const char* r2iq_cuda_version(void) { return "r2iq-b200 0.2 sm_100a fftn=2048,8192"; }
#endif

int r2iq_cuda_create(const r2iq_cuda_config* cfg, r2iq_cuda_t* out) {
    if (!cfg || !out) return fail(R2IQ_E_INVALID, "null argument");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(R2IQ_E_NODEVICE, "no CUDA device (%s); this backend has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    int dev = cfg->device;
    if (dev < 0) {
        if (cudaGetDevice(&dev) != cudaSuccess) return fail(R2IQ_E_NODEVICE, "cudaGetDevice failed");
    }
    if (dev >= ndev) return fail(R2IQ_E_INVALID, "device %d out of range (%d devices)", dev, ndev);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess)
        return fail(R2IQ_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return fail(R2IQ_E_NODEVICE, "device %d is sm_%d%d; kernels are built for sm_100a only", dev, prop.major,
                    prop.minor);
    if (cfg->fftn != 0 && cfg->fftn != 2048 && cfg->fftn != 8192)
        return fail(R2IQ_E_INVALID, "fftn %d: kernels exist for FFTN_R_ADC 2048 and 8192", cfg->fftn);
    r2iq_cuda_s* h = new (std::nothrow) r2iq_cuda_s();
    if (!h) return fail(R2IQ_E_NOMEM, "out of host memory");
    h->device = dev;
    h->sm_count = prop.multiProcessorCount;
    h->cc_major = prop.major;
    h->cc_minor = prop.minor;
    h->gain = cfg->gain;
    h->fftn = cfg->fftn == 0 ? kFftN : cfg->fftn;
    h->half = h->fftn / 2;
    h->tunebin = h->half / 4;     // constructor default mtunebin = halfFft / 4 (Core/fft_mt_r2iq.cpp:43)
    h->chunk_blocks = cfg->chunk_blocks > 0 ? cfg->chunk_blocks : kDefaultChunkBlocks;
    h->stats.sm_count = h->sm_count;
    h->stats.cc_major = prop.major;
    h->stats.cc_minor = prop.minor;
    auto bail = [&](int rc) {
        r2iq_cuda_destroy(h);
        return rc;
    };
    DeviceGuard dg(dev);                  // the caller's current device is put back when create returns
    if ((e = dg.err) != cudaSuccess) return bail(fail(R2IQ_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e)));
#ifdef R2IQ_AB_VARIANTS
    if (const char* gen = getenv("R2IQ_KERNEL_GEN")) set_kernel_generation(atoi(gen));   // comparison builds only
#endif
    if ((e = setup_channel_kernels()) != cudaSuccess)
        return bail(fail(R2IQ_E_CUDA, "channelizer kernel setup: %s", cudaGetErrorString(e)));
    if ((e = setup_fused_kernels()) != cudaSuccess)
        return bail(fail(R2IQ_E_CUDA, "kernel setup: %s", cudaGetErrorString(e)));
    if ((e = setup_wide_kernels()) != cudaSuccess || (e = setup_wide_channel_kernels()) != cudaSuccess)
        return bail(fail(R2IQ_E_CUDA, "FFTN 8192 kernel setup: %s", cudaGetErrorString(e)));
    // forward inter-pass twiddles exp(-2 pi i k2 l / 1024)
    std::vector<float2> tw(32 * 32);
    for (int k2 = 0; k2 < 32; k2++)
        for (int l = 0; l < 32; l++) {
            const double th = -2.0 * kPiD * (double)(k2 * l) / (double)kHalfFft;
            tw[k2 * 32 + l] = make_float2((float)cos(th), (float)sin(th));
        }
    if ((e = cudaMalloc(&h->d_work_counter, sizeof(unsigned long long) * r2iq_cuda_s::kCounterRing)) != cudaSuccess ||
        (e = cudaMemset(h->d_work_counter, 0, sizeof(unsigned long long) * r2iq_cuda_s::kCounterRing)) != cudaSuccess ||
        (e = cudaMalloc(&h->d_tw_fwd, sizeof(float2) * 32 * 32)) != cudaSuccess ||
        (e = cudaMemcpy(h->d_tw_fwd, tw.data(), sizeof(float2) * 32 * 32, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (h->carry = (int16_t*)malloc(sizeof(int16_t) * h->half)) == nullptr ||
        (e = cudaStreamCreateWithFlags(&h->util_stream, cudaStreamNonBlocking)) != cudaSuccess)
        return bail(fail(R2IQ_E_CUDA, "device setup: %s", cudaGetErrorString(e)));
    memset(h->carry, 0, sizeof(int16_t) * h->half);
    for (int d = 0; d < kNDec; d++) {
        h->hw[d] = design_filter_response(h->gain, d, h->fftn);
        h->hw_version[d] = 1;
    }
    *out = h;
    return R2IQ_OK;
}

int r2iq_cuda_destroy(r2iq_cuda_t h) {
    if (!h) return R2IQ_OK;
    DeviceGuard dg(h->device);
    cudaDeviceSynchronize();
    free_tables(h);
    free_channel_tables(h);
    cudaFree(h->d_ctab);
    for (auto& kv : h->wide_tables) {
        cudaFree(kv.second.g);
        cudaFree(kv.second.tw_sub);
        cudaFree(kv.second.pq);
        cudaFree(kv.second.tw_inv);
    }
    cudaFree(h->d_twq);
    cudaFree(h->d_wctab);
    cudaFree(h->d_wctab_full);
    for (int i = 0; i < kSlots; i++) {
        Slot& s = h->slots[i];
        cudaFree(s.d_out_rates);
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.h2d_done) cudaEventDestroy(s.h2d_done);
        if (s.done) cudaEventDestroy(s.done);
        if (s.carry) cudaFreeHost(s.carry);
        cudaFree(s.d_in);
        cudaFree(s.d_out);
        cudaFree(s.d_starts);
    }
    cudaFree(h->nco.d_tab);
    cudaFree(h->nco.d_rot);
    cudaFree(h->nco.d_starts);
    if (h->util_stream) cudaStreamDestroy(h->util_stream);
    cudaFree(h->d_tw_fwd);
    cudaFree(h->d_work_counter);
    free(h->carry);
    delete h;
    return R2IQ_OK;
}

int r2iq_cuda_set_filter(r2iq_cuda_t h, int d, const float* hw, int n_complex) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (d < 0 || d >= kNDec || !hw || n_complex != h->half)
        return fail(R2IQ_E_INVALID, "set_filter: d=%d n_complex=%d (want 0..6, %d)", d, n_complex, h->half);
    std::lock_guard<std::mutex> lk(h->mu);
    for (int k = 0; k < h->half; k++) h->hw[d][k] = std::complex<float>(hw[2 * k], hw[2 * k + 1]);
    h->hw_version[d]++;   // cached tables of older versions are simply never looked up again
    return R2IQ_OK;
}

int r2iq_cuda_get_filter(r2iq_cuda_t h, int d, float* hw, int n_complex) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (d < 0 || d >= kNDec || !hw || n_complex != h->half)
        return fail(R2IQ_E_INVALID, "get_filter: d=%d n_complex=%d (want 0..6, %d)", d, n_complex, h->half);
    std::lock_guard<std::mutex> lk(h->mu);
    for (int k = 0; k < h->half; k++) {
        hw[2 * k] = h->hw[d][k].real();
        hw[2 * k + 1] = h->hw[d][k].imag();
    }
    return R2IQ_OK;
}

int r2iq_cuda_set_decimate(r2iq_cuda_t h, int d) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (d < 0 || d >= kNDec) return fail(R2IQ_E_INVALID, "decimation index %d out of 0..%d", d, kNDec - 1);
    h->decimate = d;
    return R2IQ_OK;
}
int r2iq_cuda_get_decimate(r2iq_cuda_t h) { return h ? h->decimate.load() : R2IQ_E_INVALID; }
int r2iq_cuda_get_ratio(r2iq_cuda_t h) { return h ? (1 << h->decimate.load()) : R2IQ_E_INVALID; }

float r2iq_cuda_set_freq_offset(r2iq_cuda_t h, float offset) {
    if (!h) return 0.f;
    // float arithmetic and truncation exactly as the reference (Core/fft_mt_r2iq.cpp:104-106)
    // float arithmetic and truncation of Core/fft_mt_r2iq.cpp:104-106 with this operator's halfFft
    int tb = int(offset * h->half / 4) * 4;
    const float ret = (((float)tb / h->half) - offset) * (float)(1 << h->decimate.load());
    // The reference stores whatever comes out; bins outside [0, halfFft) would index out of
    // bounds there (offset is documented as [0,1)), so clamp instead of reading garbage.
    if (tb < 0) tb = 0;
    if (tb > h->half - 4) tb = h->half - 4;
    h->tunebin = tb;
    return ret;
}

int r2iq_cuda_set_tunebin(r2iq_cuda_t h, int tunebin) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (tunebin < 0 || tunebin >= h->half || (tunebin & 3))
        return fail(R2IQ_E_INVALID, "tune bin %d must be a multiple of 4 in [0,%d)", tunebin, h->half);
    h->tunebin = tunebin;
    return R2IQ_OK;
}
int r2iq_cuda_get_tunebin(r2iq_cuda_t h) { return h ? h->tunebin.load() : R2IQ_E_INVALID; }

int r2iq_cuda_set_rand(r2iq_cuda_t h, int on) {
    if (!h) return fail(R2IQ_E_INVALID, "null handle");
    h->rand_on = on != 0;
    return R2IQ_OK;
}
int r2iq_cuda_get_rand(r2iq_cuda_t h) { return h ? (int)h->rand_on.load() : R2IQ_E_INVALID; }
int r2iq_cuda_set_sideband(r2iq_cuda_t h, int lsb) {
    if (!h) return fail(R2IQ_E_INVALID, "null handle");
    h->sideband = lsb != 0;
    return R2IQ_OK;
}
int r2iq_cuda_get_sideband(r2iq_cuda_t h) { return h ? (int)h->sideband.load() : R2IQ_E_INVALID; }

float r2iq_cuda_set_fine_tune(r2iq_cuda_t h, float fc) {
    if (!h || h->poisoned) return 0.f;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);            // the table lives on the operator's device, whatever the caller's current one is
    if (dg.err != cudaSuccess) return 0.f;
    NcoState& s = h->nco;
    if (fc == 0.0f) {                 // RadioHandler.cpp:33: the mixer is skipped when fc == 0
        s.on = false;
        s.fc = 0.f;
        return 0.f;
    }
    if (s.on && s.fc == fc) return fc;   // RadioHandler.cpp:292: the state restarts only when fc changes
    nco_init(s, fc);
    s.phase = 0.0;                        // phase_start_rad = 0 (RadioHandler.cpp:295)
    float rot[256];
    for (int k = 0; k < 128; k++) {
        double ph = (double)fc * 32.0 * (double)k;
        ph = 2.0 * kPiD * (ph - floor(ph));
        rot[2 * k] = (float)cos(ph);
        rot[2 * k + 1] = (float)sin(ph);
    }
    if (!s.d_tab && cudaMalloc(&s.d_tab, sizeof(float) * 66) != cudaSuccess) return 0.f;
    if (!s.d_rot && cudaMalloc(&s.d_rot, sizeof(rot)) != cudaSuccess) return 0.f;
    // ordered after everything already queued: the tables may still be read by launches in flight
    if (cudaDeviceSynchronize() != cudaSuccess ||
        cudaMemcpy(s.d_rot, rot, sizeof(rot), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(s.d_tab, &s.tab[0][0], sizeof(float) * 66, cudaMemcpyHostToDevice) != cudaSuccess) {
        fail(R2IQ_E_CUDA, "fine-tune table upload failed");
        poison(h, g_err);
        return 0.f;
    }
    s.on = true;
    return fc;
}

int r2iq_cuda_set_fine_tune_mode(r2iq_cuda_t h, int mode) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (mode != R2IQ_NCO_REFERENCE && mode != R2IQ_NCO_DIRECT) return fail(R2IQ_E_INVALID, "fine-tune mode %d", mode);
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->nco.on && h->nco.mode != mode)
        return fail(R2IQ_E_STATE, "switch the fine-tune mixer off (fc = 0) before changing its mode: the two forms keep separate phase state");
    h->nco.mode = mode;
    return R2IQ_OK;
}
int r2iq_cuda_get_fine_tune_mode(r2iq_cuda_t h) { return h ? h->nco.mode : R2IQ_E_INVALID; }

int r2iq_cuda_turn_on(r2iq_cuda_t h) {
    int rc = check_handle(h);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    h->latched_decimate = h->decimate.load();
    h->latched_sideband = h->sideband.load();
    memset(h->carry, 0, sizeof(int16_t) * h->half);
    h->carry_valid = false;
    for (int i = 0; i < kSlots; i++) {      // results nobody fetched before the last TurnOff are gone
        if (h->slots[i].state != Slot::kFree && h->slots[i].done) cudaEventSynchronize(h->slots[i].done);
        h->slots[i].state = Slot::kFree;
    }
    h->on = true;
    return R2IQ_OK;
}

int r2iq_cuda_turn_off(r2iq_cuda_t h) {
    if (!h) return fail(R2IQ_E_INVALID, "null handle");
    {
        std::lock_guard<std::mutex> lk(h->mu);
        h->on = false;
    }
    h->cv.notify_all();                     // a submitter waiting for a held slot gives up
    if (h->poisoned) return R2IQ_OK;
    return r2iq_cuda_sync(h);
}
int r2iq_cuda_is_on(r2iq_cuda_t h) { return h ? (int)h->on : 0; }

long r2iq_cuda_out_per_block(r2iq_cuda_t h) { return h ? (long)((kTransferSamples / 2) >> h->latched_decimate) : 0; }

// ---- host-buffer paths: a ring of kSlots submissions -----------------------------------------------------------
}  // extern "C"

namespace {

// Next slot of the ring, free for a new submission.  An automatic slot is freed here by waiting for its work; a held
// slot belongs to the caller until r2iq_cuda_release -- wait for that (another thread of the caller does it: the
// ring-buffer adapter's publisher) unless the operator is switched off meanwhile.
int acquire_slot(r2iq_cuda_t h, std::unique_lock<std::mutex>& lk, Slot** out) {
    Slot& s = h->slots[h->next_ticket % kSlots];
    while (s.state != Slot::kFree) {
        if (s.state == Slot::kAuto) {
            CU(h, cudaEventSynchronize(s.done));
            s.state = Slot::kFree;
        } else {
            h->cv.wait(lk, [&] { return s.state != Slot::kHeld || !h->on || h->poisoned; });
            if (s.state == Slot::kHeld)
                return fail(R2IQ_E_STATE, "submission %llu still holds its result (r2iq_cuda_release) and the operator was switched off",
                            (unsigned long long)s.ticket);
        }
    }
    s.ticket = h->next_ticket++;
    *out = &s;
    return R2IQ_OK;
}

// [history | blocks] -> s.d_in on the slot's stream.  The history comes from the handle's carry (copied into the
// slot's own pinned buffer: several submissions are in flight), which then moves on to the tail of these blocks.
int upload_input(r2iq_cuda_t h, Slot& s, const int16_t* blocks, long nb) {
    memcpy(s.carry, h->carry, sizeof(int16_t) * h->half);
    CU(h, cudaMemcpyAsync(s.d_in, s.carry, sizeof(int16_t) * h->half, cudaMemcpyHostToDevice, s.stream));
    CU(h, cudaMemcpyAsync(s.d_in + h->half, blocks, sizeof(int16_t) * (size_t)nb * kTransferSamples, cudaMemcpyHostToDevice,
                          s.stream));
    CU(h, cudaEventRecord(s.h2d_done, s.stream));
    memcpy(h->carry, blocks + (size_t)nb * kTransferSamples - h->half, sizeof(int16_t) * h->half);
    h->carry_valid = true;
    h->stats.h2d_bytes += sizeof(int16_t) * ((size_t)h->half + (size_t)nb * kTransferSamples);
    s.n_blocks = nb;
    return R2IQ_OK;
}

// One submission of nb <= chunk_blocks blocks at the latched decimation; iq_out == nullptr keeps the result on the device.
int submit(r2iq_cuda_t h, std::unique_lock<std::mutex>& lk, const int16_t* blocks, long nb, float* iq_out, uint64_t* ticket) {
    Slot* sp = nullptr;
    int rc = acquire_slot(h, lk, &sp);
    if (rc != R2IQ_OK) return rc;
    Slot& s = *sp;
    s.d = h->latched_decimate;
    if ((rc = upload_input(h, s, blocks, nb)) != R2IQ_OK) return rc;
    if ((rc = launch(h, s.d_in, nb, s.d_out, s.stream, &s.d_starts, &s.starts_cap)) != R2IQ_OK) return rc;
    if (iq_out) {
        const size_t out_bytes = sizeof(float) * 2 * (size_t)nb * (size_t)((kTransferSamples / 2) >> s.d);
        CU(h, cudaMemcpyAsync(iq_out, s.d_out, out_bytes, cudaMemcpyDeviceToHost, s.stream));
        h->stats.d2h_bytes += out_bytes;
    }
    CU(h, cudaEventRecord(s.done, s.stream));
    s.state = iq_out ? Slot::kAuto : Slot::kHeld;
    if (ticket) *ticket = s.ticket;
    return R2IQ_OK;
}

int drain_auto_slots(r2iq_cuda_t h) {
    for (int i = 0; i < kSlots; i++) {
        Slot& s = h->slots[i];
        if (s.state == Slot::kAuto) {
            CU(h, cudaEventSynchronize(s.done));
            s.state = Slot::kFree;
        }
    }
    return R2IQ_OK;
}

Slot* find_ticket(r2iq_cuda_t h, uint64_t ticket) {
    Slot& s = h->slots[ticket % kSlots];
    return (s.state != Slot::kFree && s.ticket == ticket) ? &s : nullptr;
}

}  // namespace

extern "C" {

int r2iq_cuda_process(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history, float* iq_out) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_blocks < 0 || (n_blocks > 0 && (!blocks || !iq_out))) return fail(R2IQ_E_INVALID, "process: bad buffers");
    if (n_blocks == 0) return R2IQ_OK;
    std::unique_lock<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    if ((rc = ensure_slots(h)) != R2IQ_OK) return rc;
    const size_t out_per_block = (size_t)((kTransferSamples / 2) >> h->latched_decimate);
    if (history) memcpy(h->carry, history, sizeof(int16_t) * h->half);
    for (long done = 0; done < n_blocks;) {
        const long nb = (n_blocks - done) < h->chunk_blocks ? (n_blocks - done) : h->chunk_blocks;
        rc = submit(h, lk, blocks + (size_t)done * kTransferSamples, nb, iq_out + 2 * (size_t)done * out_per_block, nullptr);
        if (rc != R2IQ_OK) return rc;
        done += nb;
    }
    return drain_auto_slots(h);   // returns when iq_out is complete
}

int r2iq_cuda_process_async(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history, float* iq_out,
                            uint64_t* ticket) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process_async called while the operator is off (call r2iq_cuda_turn_on)");
    if (!ticket || n_blocks < 1 || !blocks) return fail(R2IQ_E_INVALID, "process_async: bad arguments");
    if (n_blocks > h->chunk_blocks)
        return fail(R2IQ_E_INVALID, "process_async: %ld blocks exceed one submission (chunk_blocks = %d)", n_blocks, h->chunk_blocks);
    std::unique_lock<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    if ((rc = ensure_slots(h)) != R2IQ_OK) return rc;
    if (history) memcpy(h->carry, history, sizeof(int16_t) * h->half);
    return submit(h, lk, blocks, n_blocks, iq_out, ticket);
}

int r2iq_cuda_wait_input(r2iq_cuda_t h, uint64_t ticket) {
    int rc = check_handle(h);
    if (rc) return rc;
    cudaEvent_t ev;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        Slot* s = find_ticket(h, ticket);
        if (!s) return R2IQ_OK;            // already waited for and recycled: its input left the host long ago
        ev = s->h2d_done;
    }
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    CU(h, cudaEventSynchronize(ev));
    return R2IQ_OK;
}

int r2iq_cuda_wait(r2iq_cuda_t h, uint64_t ticket) {
    int rc = check_handle(h);
    if (rc) return rc;
    cudaEvent_t ev;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        Slot* s = find_ticket(h, ticket);
        if (!s) return R2IQ_OK;
        ev = s->done;
    }
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    CU(h, cudaEventSynchronize(ev));
    std::lock_guard<std::mutex> lk(h->mu);
    Slot* s = find_ticket(h, ticket);
    if (s && s->state == Slot::kAuto) s->state = Slot::kFree;
    return R2IQ_OK;
}

int r2iq_cuda_read(r2iq_cuda_t h, uint64_t ticket, long first_complex, long n_complex, float* dst) {
    int rc = check_handle(h);
    if (rc) return rc;
    cudaStream_t stream;
    const float* src;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        Slot* s = find_ticket(h, ticket);
        if (!s || s->state != Slot::kHeld) return fail(R2IQ_E_STATE, "read: ticket %llu holds no result", (unsigned long long)ticket);
        const long have = s->n_blocks * (long)((kTransferSamples / 2) >> s->d);
        if (first_complex < 0 || n_complex < 0 || first_complex + n_complex > have || (n_complex > 0 && !dst))
            return fail(R2IQ_E_INVALID, "read: [%ld, %ld) outside the %ld complex samples of ticket %llu", first_complex,
                        first_complex + n_complex, have, (unsigned long long)ticket);
        stream = s->stream;
        src = s->d_out + 2 * (size_t)first_complex;
        h->stats.d2h_bytes += sizeof(float) * 2 * (size_t)n_complex;
    }
    if (n_complex == 0) return R2IQ_OK;
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    // same stream as the kernel: ordered behind it without an event
    CU(h, cudaMemcpyAsync(dst, src, sizeof(float) * 2 * (size_t)n_complex, cudaMemcpyDeviceToHost, stream));
    CU(h, cudaStreamSynchronize(stream));
    return R2IQ_OK;
}

int r2iq_cuda_release(r2iq_cuda_t h, uint64_t ticket) {
    if (!h) return fail(R2IQ_E_INVALID, "null handle");
    {
        std::lock_guard<std::mutex> lk(h->mu);
        Slot* s = find_ticket(h, ticket);
        if (s && s->state == Slot::kHeld) s->state = Slot::kFree;   // its stream is idle: every read was synchronous
    }
    h->cv.notify_all();
    return R2IQ_OK;
}

int r2iq_cuda_host_register(void* p, size_t bytes) {
    if (!p || !bytes) return fail(R2IQ_E_INVALID, "host_register: null range");
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return R2IQ_OK;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_NOMEM, "cudaHostRegister(%zu bytes): %s", bytes, cudaGetErrorString(e));
    }
    return R2IQ_OK;
}
int r2iq_cuda_host_unregister(void* p) {
    if (!p) return R2IQ_OK;
    cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_INVALID, "cudaHostUnregister: %s", cudaGetErrorString(e));
    }
    return R2IQ_OK;
}

int r2iq_cuda_process_rates(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history,
                            const int* decimations, int n_rates, float* const* iq_outs) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process_rates called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_rates < 1 || n_rates > kNDec || !decimations || !iq_outs) return fail(R2IQ_E_INVALID, "process_rates: 1..%d rates", kNDec);
    for (int r = 0; r < n_rates; r++)
        if (decimations[r] < 0 || decimations[r] >= kNDec || (n_blocks > 0 && !iq_outs[r]))
            return fail(R2IQ_E_INVALID, "process_rates: bad decimation index or output buffer at position %d", r);
    // every rate at most once: the per-chunk device buffer holds ONE copy of each decimation's output
    // (sum over d of 32768 >> d < 2 * 32768 complex per block)
    for (int r = 0; r < n_rates; r++)
        for (int q = 0; q < r; q++)
            if (decimations[q] == decimations[r])
                return fail(R2IQ_E_INVALID, "process_rates: decimation index %d listed twice (positions %d and %d)",
                            decimations[r], q, r);
    if (h->nco.on) return fail(R2IQ_E_STATE, "process_rates: the fine-tune mixer belongs to one rate; switch it off first");
    if (n_blocks < 0 || (n_blocks > 0 && !blocks)) return fail(R2IQ_E_INVALID, "process_rates: bad buffers");
    if (n_blocks == 0) return R2IQ_OK;
    std::unique_lock<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    if ((rc = ensure_slots(h)) != R2IQ_OK) return rc;
    const size_t rates_floats = 2 * 2 * (size_t)h->chunk_blocks * (kTransferSamples / 2);
    for (int i = 0; i < kSlots; i++)
        if (!h->slots[i].d_out_rates) CU(h, cudaMalloc(&h->slots[i].d_out_rates, sizeof(float) * rates_floats));
    if (history) memcpy(h->carry, history, sizeof(int16_t) * h->half);
    const int latched = h->latched_decimate;
    for (long done = 0; done < n_blocks && rc == R2IQ_OK;) {
        const long nb = (n_blocks - done) < h->chunk_blocks ? (n_blocks - done) : h->chunk_blocks;
        Slot* sp = nullptr;
        if ((rc = acquire_slot(h, lk, &sp)) != R2IQ_OK) break;
        Slot& s = *sp;
        if ((rc = upload_input(h, s, blocks + (size_t)done * kTransferSamples, nb)) != R2IQ_OK) break;
        size_t off = 0;   // floats into d_out_rates
        for (int r = 0; r < n_rates && rc == R2IQ_OK; r++) {
            const int d = decimations[r];
            const size_t opb = (size_t)((kTransferSamples / 2) >> d);
            if (off + 2 * (size_t)nb * opb > rates_floats) {
                rc = fail(R2IQ_E_INVALID, "process_rates: outputs exceed the per-chunk device buffer");
                break;
            }
            h->latched_decimate = d;      // the kernels and tables are chosen per launch from the latched index
            rc = launch(h, s.d_in, nb, s.d_out_rates + off, s.stream, &s.d_starts, &s.starts_cap);
            if (rc != R2IQ_OK) break;
            const size_t out_bytes = sizeof(float) * 2 * (size_t)nb * opb;
            cudaError_t ce = cudaMemcpyAsync(iq_outs[r] + 2 * (size_t)done * opb, s.d_out_rates + off, out_bytes,
                                             cudaMemcpyDeviceToHost, s.stream);
            if (ce != cudaSuccess) rc = fail(R2IQ_E_CUDA, "process_rates: %s", cudaGetErrorString(ce));
            h->stats.d2h_bytes += out_bytes;
            off += 2 * (size_t)h->chunk_blocks * opb;
        }
        h->latched_decimate = latched;
        if (rc != R2IQ_OK) break;
        CU(h, cudaEventRecord(s.done, s.stream));
        s.state = Slot::kAuto;
        done += nb;
    }
    const int rc2 = drain_auto_slots(h);
    return rc != R2IQ_OK ? rc : rc2;
}

int r2iq_cuda_process_device(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_iq_out, void* stream) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_blocks < 0 || (n_blocks > 0 && (!d_in || !d_iq_out))) return fail(R2IQ_E_INVALID, "process_device: bad buffers");
    if (((uintptr_t)d_in & 15) || ((uintptr_t)d_iq_out & 7))
        return fail(R2IQ_E_INVALID, "process_device: input must be 16-byte and output 8-byte aligned");
    if (n_blocks == 0) return R2IQ_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);            // tables are built (cudaMalloc/cudaMemcpy) on first use of a (d, tune bin)
    CU(h, dg.err);
    return launch(h, d_in, n_blocks, d_iq_out, (cudaStream_t)stream, &h->nco.d_starts, &h->nco.starts_cap);
}

int r2iq_cuda_process_channels_device(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, const int* tunebins,
                                      int n_channels, float* d_iq_out, void* stream) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_blocks < 0 || n_channels < 0 || ((n_blocks > 0 && n_channels > 0) && (!d_in || !d_iq_out || !tunebins)))
        return fail(R2IQ_E_INVALID, "process_channels_device: bad arguments");
    if (((uintptr_t)d_in & 15) || ((uintptr_t)d_iq_out & 15))
        return fail(R2IQ_E_INVALID, "process_channels_device: input and output must be 16-byte aligned");
    if (n_blocks == 0 || n_channels == 0) return R2IQ_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    return launch_channels_impl(h, d_in, n_blocks, tunebins, n_channels, d_iq_out, (cudaStream_t)stream);
}

int r2iq_cuda_process_channels(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history,
                               const int* tunebins, int n_channels, float* iq_out) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_blocks < 0 || n_channels < 0 || ((n_blocks > 0 && n_channels > 0) && (!blocks || !iq_out || !tunebins)))
        return fail(R2IQ_E_INVALID, "process_channels: bad arguments");
    if (n_blocks == 0 || n_channels == 0) return R2IQ_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    CU(h, cudaSetDevice(h->device));
    const int d = h->latched_decimate;
    const size_t per_ch = (size_t)n_blocks * ((kTransferSamples / 2) >> d);
    int16_t* d_in = nullptr;
    float* d_out = nullptr;
    DeviceScratch scratch;                // freed on every way out, the early returns of CU() included
    const size_t hist = (size_t)h->half;
    cudaError_t e = scratch.alloc(&d_in, sizeof(int16_t) * (hist + (size_t)n_blocks * kTransferSamples));
    if (e == cudaSuccess) e = scratch.alloc(&d_out, sizeof(float) * 2 * per_ch * (size_t)n_channels);
    if (e != cudaSuccess) {
        cudaGetLastError();               // out of memory is the caller's problem, not a poisoned handle
        return fail(R2IQ_E_NOMEM, "cudaMalloc(channelizer buffers): %s", cudaGetErrorString(e));
    }
    if (history) CU(h, cudaMemcpyAsync(d_in, history, sizeof(int16_t) * hist, cudaMemcpyHostToDevice, h->util_stream));
    else CU(h, cudaMemsetAsync(d_in, 0, sizeof(int16_t) * hist, h->util_stream));
    CU(h, cudaMemcpyAsync(d_in + hist, blocks, sizeof(int16_t) * (size_t)n_blocks * kTransferSamples,
                          cudaMemcpyHostToDevice, h->util_stream));
    rc = launch_channels_impl(h, d_in, n_blocks, tunebins, n_channels, d_out, h->util_stream);
    if (rc == R2IQ_OK) {
        CU(h, cudaMemcpyAsync(iq_out, d_out, sizeof(float) * 2 * per_ch * (size_t)n_channels, cudaMemcpyDeviceToHost,
                              h->util_stream));
        h->stats.h2d_bytes += sizeof(int16_t) * (hist + (size_t)n_blocks * kTransferSamples);
        h->stats.d2h_bytes += sizeof(float) * 2 * per_ch * (size_t)n_channels;
    }
    CU(h, cudaStreamSynchronize(h->util_stream));
    return rc;
}

int r2iq_cuda_sync(r2iq_cuda_t h) {
    int rc = check_handle(h);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaDeviceSynchronize());
    return R2IQ_OK;
}

// ---- NUMA placement of page-locked buffers ---------------------------------------------------------------------
// A pinned buffer is what the GPU's copy engines read and write over PCIe.  On a two-socket 8-GPU box a buffer that
// sits on the other socket makes every copy cross the socket interconnect, and with all ranks allocating from the
// socket their launcher happened to start on, the copies of all GPUs share one memory controller set.  The pages
// are therefore placed on the NUMA node the GPU's PCIe root port belongs to (sysfs numa_node of its bus id) by
// setting the calling thread's memory policy to "preferred: that node" around cudaMallocHost (raw set_mempolicy
// syscalls: libnuma is not assumed).  Every failure (no sysfs entry, node -1, syscall filtered) degrades to the
// default first-touch placement.  R2IQ_NUMA=0 switches it off (A/B measurement).
namespace {
int device_numa_node_impl(int device) {
    if (const char* v = getenv("R2IQ_NUMA")) {
        if (atoi(v) == 0) return -1;
    }
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return -1;
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, (int)sizeof(bus), device) != cudaSuccess) return -1;
    for (char* c = bus; *c; ++c)
        if (*c >= 'A' && *c <= 'Z') *c = (char)(*c - 'A' + 'a');
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return (node >= 0 && node < 1024) ? node : -1;
}
constexpr int kMpolDefault = 0, kMpolPreferred = 1;
constexpr unsigned long kMaskWords = 1024 / (8 * sizeof(unsigned long));
}  // namespace

int r2iq_cuda_device_numa_node(int device) { return device_numa_node_impl(device); }

void* r2iq_cuda_host_alloc_on(size_t bytes, int device) {
    const int node = device_numa_node_impl(device);
    int old_mode = kMpolDefault;
    unsigned long old_mask[kMaskWords] = {0}, mask[kMaskWords] = {0};
    bool policy_set = false;
    if (node >= 0 && syscall(SYS_get_mempolicy, &old_mode, old_mask, 1024UL, nullptr, 0UL) == 0) {
        mask[node / (8 * sizeof(unsigned long))] = 1UL << (node % (8 * sizeof(unsigned long)));
        policy_set = syscall(SYS_set_mempolicy, kMpolPreferred, mask, 1024UL) == 0;
    }
    void* p = nullptr;
    cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
    if (policy_set) syscall(SYS_set_mempolicy, old_mode, old_mode == kMpolDefault ? nullptr : old_mask, old_mode == kMpolDefault ? 0UL : 1024UL);
    if (e != cudaSuccess) {
        fail(R2IQ_E_NOMEM, "cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}
void* r2iq_cuda_host_alloc(size_t bytes) { return r2iq_cuda_host_alloc_on(bytes, -1); }
int r2iq_cuda_host_free(void* p) {
    if (!p) return R2IQ_OK;
    cudaError_t e = cudaFreeHost(p);
    if (e != cudaSuccess) return fail(R2IQ_E_CUDA, "cudaFreeHost: %s", cudaGetErrorString(e));
    return R2IQ_OK;
}

// ---- device memory shared between the ranks of one box (time-slab / channel sharding) ---------------
// One process per GPU: the rank that collects the IQ allocates the destination, the others map it (CUDA IPC, peer
// access over NVLink) and hand the mapped pointer to r2iq_cuda_process_device as d_iq_out -- the fused kernel's own
// stores then cross NVLink while it computes, and no gather follows the kernel.
void* r2iq_cuda_device_alloc(int device, size_t bytes) {
    DeviceGuard dg(device);
    if (dg.err != cudaSuccess) {
        fail(R2IQ_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(dg.err));
        return nullptr;
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        fail(R2IQ_E_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}
int r2iq_cuda_device_free(void* p) {
    if (!p) return R2IQ_OK;
    cudaError_t e = cudaFree(p);
    if (e != cudaSuccess) return fail(R2IQ_E_CUDA, "cudaFree: %s", cudaGetErrorString(e));
    return R2IQ_OK;
}
int r2iq_cuda_ipc_export(void* device_ptr, unsigned char handle[R2IQ_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == R2IQ_IPC_HANDLE_BYTES, "handle size");
    if (!device_ptr || !handle) return fail(R2IQ_E_INVALID, "ipc_export: null argument");
    cudaIpcMemHandle_t hnd;
    cudaError_t e = cudaIpcGetMemHandle(&hnd, device_ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    }
    memcpy(handle, &hnd, sizeof(hnd));
    return R2IQ_OK;
}
int r2iq_cuda_ipc_open(int device, const unsigned char handle[R2IQ_IPC_HANDLE_BYTES], void** device_ptr) {
    if (!handle || !device_ptr) return fail(R2IQ_E_INVALID, "ipc_open: null argument");
    *device_ptr = nullptr;
    DeviceGuard dg(device);
    if (dg.err != cudaSuccess) return fail(R2IQ_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(dg.err));
    cudaIpcMemHandle_t hnd;
    memcpy(&hnd, handle, sizeof(hnd));
    cudaError_t e = cudaIpcOpenMemHandle(device_ptr, hnd, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
    }
    return R2IQ_OK;
}
int r2iq_cuda_ipc_close(void* device_ptr) {
    if (!device_ptr) return R2IQ_OK;
    cudaError_t e = cudaIpcCloseMemHandle(device_ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_CUDA, "cudaIpcCloseMemHandle: %s", cudaGetErrorString(e));
    }
    return R2IQ_OK;
}

int r2iq_cuda_convert(r2iq_cuda_t h, const int16_t* in, long n, int rand, float* out) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (n < 0 || (n & 1) || (n > 0 && (!in || !out))) return fail(R2IQ_E_INVALID, "convert: n must be even and buffers non-null");
    if (n == 0) return R2IQ_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    CU(h, cudaSetDevice(h->device));
    int16_t* d_in = nullptr;
    float* d_out = nullptr;
    DeviceScratch scratch;
    cudaError_t e = scratch.alloc(&d_in, sizeof(int16_t) * (size_t)n);
    if (e == cudaSuccess) e = scratch.alloc(&d_out, sizeof(float) * (size_t)n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(R2IQ_E_NOMEM, "cudaMalloc(convert buffers): %s", cudaGetErrorString(e));
    }
    CU(h, cudaMemcpy(d_in, in, sizeof(int16_t) * (size_t)n, cudaMemcpyHostToDevice));
    CU(h, launch_convert(d_in, d_out, n, rand, h->util_stream));
    CU(h, cudaStreamSynchronize(h->util_stream));
    CU(h, cudaMemcpy(out, d_out, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    return R2IQ_OK;
}

int r2iq_cuda_get_fftn(r2iq_cuda_t h) { return h ? h->fftn : R2IQ_E_INVALID; }

int r2iq_cuda_kernel_name(r2iq_cuda_t h, int d, char* buf, int n) {
    if (!h || !buf || n < 1 || d < 0 || d >= kNDec) return fail(R2IQ_E_INVALID, "kernel_name: bad arguments");
    if (h->fftn == 8192) snprintf(buf, (size_t)n, "%s", wide_kernel_name(d, h->nco.on && h->nco.mode == R2IQ_NCO_DIRECT));
    else snprintf(buf, (size_t)n, "%s", fused_kernel_name(d, h->nco.on && h->nco.mode == R2IQ_NCO_DIRECT));
    return R2IQ_OK;
}

int r2iq_cuda_get_stats(r2iq_cuda_t h, r2iq_cuda_stats* out) {
    if (!h || !out) return fail(R2IQ_E_INVALID, "null argument");
    *out = h->stats;
    return R2IQ_OK;
}

int r2iq_cuda_design_filter(float gain, int d, float* hw, int n_complex) {
    if (d < 0 || d >= kNDec || !hw || n_complex != kHalfFft)
        return fail(R2IQ_E_INVALID, "design_filter: d=%d n_complex=%d (want 0..6, %d)", d, n_complex, kHalfFft);
    const std::vector<std::complex<float>> H = design_filter_response(gain, d, kFftN);
    for (int k = 0; k < kHalfFft; k++) {
        hw[2 * k] = H[k].real();
        hw[2 * k + 1] = H[k].imag();
    }
    return R2IQ_OK;
}

int r2iq_cuda_kaiser(int num_taps, float astop_db, float f_pass, float f_stop, float* taps) {
    if (num_taps <= 0 || !taps) return fail(R2IQ_E_INVALID, "kaiser: num_taps must be positive");
    const std::vector<float> t = kaiser_lowpass(num_taps, astop_db, f_pass, f_stop);
    memcpy(taps, t.data(), sizeof(float) * (size_t)num_taps);
    return R2IQ_OK;
}

float r2iq_cuda_freq_to_tunebin(float offset, int d, int* tunebin) {
    const int tb = int(offset * kHalfFft / 4) * 4;
    const float delta = ((float)tb / kHalfFft) - offset;
    if (tunebin) *tunebin = tb;
    return delta * (float)(1 << (d < 0 ? 0 : d));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// r2iq_cuda_process_device_static: the device path without any allocation (include/r2iq_cuda.h).  No counterpart in
// the reference.  The fused FFTN 2048 kernel of r2iq_pair.cu is launched
// with its whole working set -- work counter and per-(filter, d, tune bin) tables -- inside g_static_scratch below
// instead of the handle's cudaMalloc'ed buffers.  Nothing above this line is touched by it.
//
// Layout of the array (bytes):      0  work counter (8 used; cleared before every launch)
//                                 256  d = 0: gp[1024] float2      d >= 1: (P, Q) in band order, float4[mfft] (<= 8 KB)
//                                8448  d = 0: gq[1024] float2
//                               16640  tw_inv[32][32] float2
//                               24832  tw_aux (d = 1..4: (R + 32) * 32 float2 <= 12 KB; d = 5, 6: float4[1024] = 16 KB)
//                               41216  end
// ---------------------------------------------------------------------------------------------
namespace {

constexpr size_t kStCounter = 0;
constexpr size_t kStP = 256;
constexpr size_t kStQ = kStP + sizeof(float2) * kHalfFft;
constexpr size_t kStTwInv = kStQ + sizeof(float2) * kHalfFft;
constexpr size_t kStTwAux = kStTwInv + sizeof(float2) * 32 * 32;
constexpr size_t kStEnd = kStTwAux + sizeof(float4) * 32 * 32;
static_assert(kStEnd <= R2IQ_STATIC_SCRATCH_BYTES, "table set larger than the static scratch array");

// In host code this name is only a stand-in for the device symbol: never pass it to a kernel -- static_scratch_address()
// resolves the device address (cudaGetSymbolAddress) once per device.
static __device__ __align__(256) unsigned char g_static_scratch[R2IQ_STATIC_SCRATCH_BYTES];

constexpr int kStaticMaxDevices = 64;
struct StaticDevice {
    unsigned char* base = nullptr;        // device address of g_static_scratch on this device
    bool tables_valid = false;            // the array holds the tables of (filter, d, tunebin)
    int d = -1, tunebin = -1;
    uint32_t row_mask = 0;
    std::complex<float> filter[kHalfFft];
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mu;                                    // the staging image and the per-device records
alignas(16) unsigned char g_static_img[kStEnd];            // host image of the tables on their way to the device

// Set-up, once per device, with that device current (the caller holds a DeviceGuard).
int static_scratch_address(r2iq_cuda_t h, StaticDevice& sd) {
    if (sd.base) return R2IQ_OK;
    unsigned char* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
        cudaGetLastError();
        return fail(R2IQ_E_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", h->device,
                    e == cudaSuccess ? "null address" : cudaGetErrorString(e));
    }
    sd.base = ptr;
    return R2IQ_OK;
}

// The tables get_tables() builds for the shipped kernels, written into the host image (same float64 expressions, so a
// launch from the array delivers the bits of a launch from the handle's tables).  Returns the row mask.
uint32_t static_build_tables(const std::complex<float>* H, int d, int tunebin) {
    const int M = kHalfFft >> d;
    memset(g_static_img, 0, sizeof(g_static_img));
    float2* gp = reinterpret_cast<float2*>(g_static_img + kStP);
    float2* gq = reinterpret_cast<float2*>(g_static_img + kStQ);
    float4* pqb = reinterpret_cast<float4*>(g_static_img + kStP);
    float2* tw = reinterpret_cast<float2*>(g_static_img + kStTwInv);
    uint32_t row_mask = 0;
    // (P, Q) = gk +- (-i W_2048^k) gk of band offset s; outside [0, halfFft) the band is cut (impl.hpp:91-96)
    auto pq_at = [&](int s, float2* P, float2* Q) {
        const int k = tunebin + s;
        *P = *Q = make_float2(0.f, 0.f);
        if (k < 0 || k >= kHalfFft) return;
        const std::complex<double> hh = std::complex<double>(H[(s + kHalfFft) % kHalfFft]) * 0.5;
        const double th = 2.0 * kPiD * (double)k / (double)kFftN;
        const std::complex<double> c(-sin(th), -cos(th));
        const std::complex<double> gc = c * hh;
        *P = make_float2((float)(hh.real() + gc.real()), (float)(hh.imag() + gc.imag()));
        *Q = make_float2((float)(hh.real() - gc.real()), (float)(hh.imag() - gc.imag()));
    };
    for (int m = 0; m < M; m++) {
        const int s = m < M / 2 ? m : m - M;
        const int kk = (tunebin + s) & (kHalfFft - 1);
        row_mask |= 1u << (kk >> 5);
        row_mask |= 1u << (((kHalfFft - kk) & (kHalfFft - 1)) >> 5);
        float2 P, Q;
        pq_at(s, &P, &Q);
        if (d == 0) {
            if (tunebin + s >= 0 && tunebin + s < kHalfFft) {
                gp[tunebin + s] = P;
                gq[tunebin + s] = Q;
            }
        } else {
            pqb[m] = make_float4(P.x, P.y, Q.x, Q.y);
        }
    }
    if (d == 0) {
        for (int n = 0; n < 32; n++)
            for (int k2 = 0; k2 < 32; k2++) {
                const int l = (n + tunebin) & 31;
                const long long n1 = ((l - tunebin) % kHalfFft + kHalfFft) % kHalfFft;
                const double th = 2.0 * kPiD * (double)((n1 * k2) % kHalfFft) / (double)kHalfFft;
                tw[n * 32 + k2] = make_float2((float)cos(th), (float)sin(th));
            }
    } else {
        const int L = d <= 5 ? (32 >> d) : 1;
        for (int k2 = 0; k2 < 32; k2++)
            for (int n1 = 0; n1 < L; n1++) {
                const double th = 2.0 * kPiD * (double)((n1 * k2) % M) / (double)M;
                tw[k2 * L + n1] = make_float2((float)cos(th), (float)sin(th));
            }
    }
    if (d >= 5) {
        float4* zc = reinterpret_cast<float4*>(g_static_img + kStTwAux);
        for (int c = 0; c < 32; c++) {
            auto offset = [&](int u, int* s_out) {
                u &= 31;
                if (u < M / 2) { *s_out = u; return true; }
                if (u >= 32 - M / 2) { *s_out = u - 32; return true; }
                return false;
            };
            int sb = 0, sm = 0;
            const bool has_b = offset(c - tunebin, &sb), has_m = offset(-c - tunebin, &sm);
            const int kb = (tunebin + sb) & (kHalfFft - 1);
            const int km = (kHalfFft - ((tunebin + sm) & (kHalfFft - 1))) & (kHalfFft - 1);
            for (int n1 = 0; n1 < 32; n1++) {
                const double tb_ = -2.0 * kPiD * (double)((n1 * kb) % kHalfFft) / (double)kHalfFft;
                const double tm_ = -2.0 * kPiD * (double)((n1 * km) % kHalfFft) / (double)kHalfFft;
                zc[c * 32 + n1] = make_float4(has_b ? (float)cos(tb_) : 0.f, has_b ? (float)sin(tb_) : 0.f,
                                              has_m ? (float)cos(tm_) : 0.f, has_m ? (float)sin(tm_) : 0.f);
            }
        }
    } else if (d >= 1) {
        const std::vector<float2> ta = build_tw_aux(M, true);
        memcpy(g_static_img + kStTwAux, ta.data(), sizeof(float2) * ta.size());
    }
    return row_mask;
}

}  // namespace

extern "C" int r2iq_cuda_process_device_static(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_iq_out, void* stream) {
    int rc = check_handle(h);
    if (rc) return rc;
    if (!h->on) return fail(R2IQ_E_STATE, "process called while the operator is off (call r2iq_cuda_turn_on)");
    if (n_blocks < 0 || (n_blocks > 0 && (!d_in || !d_iq_out))) return fail(R2IQ_E_INVALID, "process_device_static: bad buffers");
    if (((uintptr_t)d_in & 15) || ((uintptr_t)d_iq_out & 7))
        return fail(R2IQ_E_INVALID, "process_device_static: input must be 16-byte and output 8-byte aligned");
    if (n_blocks > R2IQ_STATIC_MAX_BLOCKS)
        return fail(R2IQ_E_INVALID, "process_device_static: %ld blocks in one call, the static entry point takes at most %d "
                    "(use r2iq_cuda_process_device for larger launches)", n_blocks, R2IQ_STATIC_MAX_BLOCKS);
    if (h->fftn != kFftN)
        return fail(R2IQ_E_INVALID, "process_device_static: the FFTN %d tables do not fit the %d-byte static scratch array "
                    "(FFTN 2048 operators only)", h->fftn, R2IQ_STATIC_SCRATCH_BYTES);
    if (h->device < 0 || h->device >= kStaticMaxDevices)
        return fail(R2IQ_E_INVALID, "process_device_static: device %d (at most %d devices)", h->device, kStaticMaxDevices);
    if (kernel_generation() != 0)
        return fail(R2IQ_E_STATE, "process_device_static: comparison kernel generations use other table layouts");
    if (n_blocks == 0) return R2IQ_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->nco.on && h->nco.mode != R2IQ_NCO_DIRECT)
        return fail(R2IQ_E_STATE, "process_device_static: the R2IQ_NCO_REFERENCE fine-tune mixer needs start phasors that grow "
                    "with the launch; switch to R2IQ_NCO_DIRECT or use r2iq_cuda_process_device");
    DeviceGuard dg(h->device);
    CU(h, dg.err);
    std::lock_guard<std::mutex> slk(g_static_mu);
    StaticDevice& sd = g_static_dev[h->device];
    if ((rc = static_scratch_address(h, sd)) != R2IQ_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int d = h->latched_decimate;
    const int tunebin = h->tunebin.load();
    const std::complex<float>* H = h->hw[d].data();
    if (!sd.tables_valid || sd.d != d || sd.tunebin != tunebin || memcmp(sd.filter, H, sizeof(sd.filter)) != 0) {
        sd.tables_valid = false;
        sd.row_mask = static_build_tables(H, d, tunebin);
        // pageable source: the call returns once the image sits in the driver's staging buffer; in stream order before the kernel
        CU(h, cudaMemcpyAsync(sd.base + kStP, g_static_img + kStP, kStEnd - kStP, cudaMemcpyHostToDevice, st));
        memcpy(sd.filter, H, sizeof(sd.filter));
        sd.d = d;
        sd.tunebin = tunebin;
        sd.tables_valid = true;
    }
    // the array keeps the last call's contents: the counter must start at zero
    CU(h, cudaMemsetAsync(sd.base + kStCounter, 0, 256, st));
    const int M = kHalfFft >> d;
    FusedParams p;
    p.in_words = (const uint32_t*)d_in;
    p.out = (float2*)d_iq_out;
    p.total_frames = (long long)n_blocks * kFramesPerBlock;
    p.tw_fwd = h->d_tw_fwd;               // a constant of the handle, uploaded by r2iq_cuda_create
    p.gk = nullptr;                       // read by the comparison generations only
    p.gck = nullptr;
    p.gp = reinterpret_cast<const float2*>(sd.base + kStP);
    p.gq = d == 0 ? reinterpret_cast<const float2*>(sd.base + kStQ) : nullptr;
    p.tw_inv = reinterpret_cast<const float2*>(sd.base + kStTwInv);
    p.tw_aux = d == 0 ? nullptr : reinterpret_cast<const float2*>(sd.base + kStTwAux);
    p.tunebin = tunebin;
    p.band_lo = tunebin - M / 2 < 0 ? 0 : tunebin - M / 2;
    p.band_hi = tunebin + M / 2 > kHalfFft ? kHalfFft : tunebin + M / 2;
    p.row_mask = sd.row_mask;
    p.rand_mul = h->rand_on ? 0xFFFEu : 0u;
    p.imag_sign = h->latched_sideband ? -1.0f : 1.0f;
    p.work_counter = reinterpret_cast<unsigned long long*>(sd.base + kStCounter);
    p.nco_starts = nullptr;
    p.nco_tab = nullptr;
    p.nco_rot = nullptr;
    p.nco_phase0 = 0.0;
    p.nco_fc = 0.0;
    if (h->nco.on) {                      // direct-phase mixer in the store epilogue: d_rot was uploaded by set_fine_tune
        p.nco_rot = (const float2*)h->nco.d_rot;
        p.nco_phase0 = h->nco.phase;
        p.nco_fc = (double)h->nco.fc;
        nco_direct_advance(h->nco, (size_t)n_blocks * (size_t)((kTransferSamples / 2) >> d));
    }
    CU(h, launch_fused(d, p, h->sm_count, st));
    h->stats.kernel_launches++;
    h->stats.blocks_processed += (uint64_t)n_blocks;
    return R2IQ_OK;
}
