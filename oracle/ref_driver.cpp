// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// C-ABI harness around the UNMODIFIED reference operator (class fft_mt_r2iq, compiled in place
// from /root/reference/Core by oracle/Makefile against oracle/shim/fftw3.h).  It plays the two
// roles RadioHandlerClass plays around the operator -- owner of the input/output ring buffers and
// producer/consumer of their slots (Core/RadioHandler.cpp:142,166-171; mock producers in
// unittest/core_test.cpp:71-85 and unittest/benchmark_test.cpp:78-107) -- and nothing else.
// Outputs go to oracle/_ref/libref_r2iq_<FFTN>.so, which is git-ignored but travels to the GPU
// box; nothing here is linked into, imported by or called from the product path.
//
// Used for: (1) pinning the numpy restatement (oracle/r2iq_oracle.py) and generating the golden
// fixtures under tests/golden/; (2) the parity checks in tests/; (3) the timed CPU baseline
// ("kind": "reference") in bench.py.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

// The harness needs two private members of the reference class (filterHw to read back the
// designed filter spectra, mtunebin/threadArgs to pin a tune bin and to force an ISA variant).
// The reference offers no accessor, so widen access for THIS translation unit only; the class
// layout is unchanged, so it stays link-compatible with the unmodified reference objects.
#define private public
#define protected public
#include "config.h"
#undef FFTN_R_ADC
#define FFTN_R_ADC (ORACLE_FFTN)
#include "fft_mt_r2iq.h"
#include "fir.h"
#undef private
#undef protected

#include "fftw3.h"

namespace {

struct Session {
    ringbuffer<int16_t> in;
    ringbuffer<float> out;
    fft_mt_r2iq* op = nullptr;
    std::thread worker;          // only when an ISA variant is forced
    bool forced = false;
};

// Equivalent of fft_mt_r2iq::TurnOn (Core/fft_mt_r2iq.cpp:111-129) that runs a chosen ISA
// variant of the worker instead of the cpuid dispatcher's pick (Core/fft_mt_r2iq.cpp:263-308).
void turn_on(Session& s, int variant) {
    if (variant == 0) {
        s.op->TurnOn();
        s.forced = false;
        return;
    }
    fft_mt_r2iq* op = s.op;
    op->r2iqOn = true;
    op->bufIdx = 0;
    op->lastThread = op->threadArgs[0];
    op->inputSeq.store(0);
    op->outputWriteTurn.store(0);
    op->sharedPout = nullptr;
    s.in.Start();
    s.out.Start();
    s.forced = true;
    s.worker = std::thread([op, variant] {
        r2iqThreadArg* th = op->threadArgs[0];
        switch (variant) {
            case 1: op->r2iqThreadf_def(th); break;
            case 2: op->r2iqThreadf_avx(th); break;
            case 3: op->r2iqThreadf_avx2(th); break;
            default: op->r2iqThreadf_avx512(th); break;
        }
    });
}

void turn_off(Session& s) {
    if (!s.forced) {
        s.op->TurnOff();
        return;
    }
    s.op->r2iqOn = false;
    s.in.Stop();
    s.out.Stop();
    s.op->outputCV.notify_all();
    s.worker.join();
}

void setup(Session& s, float gain, int d, int tunebin, int rnd, int sideband) {
    s.in.setBlockSize(transferSamples);                      // unittest/core_test.cpp:73
    s.out.setBlockSize(EXT_BLOCKLEN * 2 * sizeof(float));    // Core/RadioHandler.cpp:166
    s.op = new fft_mt_r2iq();
    s.op->Init(gain, &s.in, &s.out);
    s.op->setDecimate(d);
    s.op->updateRand(rnd != 0);
    s.op->setSideband(sideband != 0);
    if (tunebin >= 0) s.op->mtunebin = tunebin;
}

}  // namespace

extern "C" {

int refr2iq_fftn(void) { return FFTN_R_ADC; }
int refr2iq_half_fft(void) { return halfFft; }
int refr2iq_fft_per_buf(void) { return fftPerBuf; }
int refr2iq_max_threads(void) { return N_MAX_R2IQ_THREADS; }

int refr2iq_set_fft_backend(int backend, const char* libpath) { return fftwshim_set_backend(backend, libpath); }
int refr2iq_get_fft_backend(void) { return fftwshim_get_backend(); }

// Core/fir.cpp:48 KaiserWindow, straight through.
int refr2iq_kaiser(int num_taps, float astop, float fpass, float fstop, float* coef) {
    return KaiserWindow(num_taps, astop, fpass, fstop, coef);
}

// filterHw[d] as built by fft_mt_r2iq::Init (Core/fft_mt_r2iq.cpp:163-209); hw = halfFft complex floats.
int refr2iq_filter(float gain, int d, float* hw) {
    if (d < 0 || d >= NDECIDX) return -1;
    ringbuffer<int16_t> in;
    ringbuffer<float> out;
    fft_mt_r2iq* op = new fft_mt_r2iq();
    op->Init(gain, &in, &out);
    memcpy(hw, op->filterHw[d], sizeof(float) * 2 * halfFft);
    delete op;
    return halfFft;
}

// fft_mt_r2iq::setFreqOffset (Core/fft_mt_r2iq.cpp:101-109): returns the residual fine-tune value
// and reports the tune bin it selected.
float refr2iq_set_freq_offset(float offset, int d, int* tunebin) {
    fft_mt_r2iq op;
    op.setDecimate(d);
    float r = op.setFreqOffset(offset);
    if (tunebin) *tunebin = op.mtunebin;
    return r;
}

// fft_mt_r2iq::convert_float<rand> (Core/fft_mt_r2iq.h:36-51), scalar template.
void refr2iq_convert(const int16_t* in, float* out, int n, int rnd) {
    fft_mt_r2iq op;
    if (rnd) op.convert_float<true>(in, out, n);
    else op.convert_float<false>(in, out, n);
}

// Run the operator over `n_blocks` input blocks of transferSamples int16 each and collect the IQ
// it publishes: n_blocks * (transferSamples/2 >> d) complex floats, interleaved re,im.
// `history` = the halfFft samples that precede block 0 (NULL = zeros).  The reference reads them
// from the tail of the previous ring slot (Core/fft_mt_r2iq_impl.hpp:32), which is uninitialised
// heap for the very first block; the harness writes them there so the result is deterministic.
// tunebin < 0 keeps the constructor default halfFft/4.  variant: 0 dispatcher, 1 def, 2 avx,
// 3 avx2 intrinsics, 4 avx512.  Returns the number of complex samples written, <0 on error.
long refr2iq_run(const int16_t* blocks, long n_blocks, const int16_t* history, float gain, int d, int tunebin,
                 int rnd, int sideband, int variant, float* iq_out) {
    if (d < 0 || d >= NDECIDX || n_blocks <= 0) return -1;
    Session s;
    setup(s, gain, d, tunebin, rnd, sideband);
    const long ratio = 1L << d;
    const long n_padded = (n_blocks + ratio - 1) / ratio * ratio;   // complete the last output slot
    const long n_slots = n_padded / ratio;
    const long per_block = (transferSamples / 2) >> d;              // complex samples per input block

    turn_on(s, variant);   // Start() resets both rings: read = write = 0
    // history lives in the tail of slot (read_index - 1) = the last slot of the ring
    {
        int16_t* prev = s.in.peekWritePtr(-1);
        if (history) memcpy(prev + transferSamples - halfFft, history, sizeof(int16_t) * halfFft);
        else memset(prev + transferSamples - halfFft, 0, sizeof(int16_t) * halfFft);
    }
    std::thread producer([&] {
        for (long b = 0; b < n_padded; b++) {
            int16_t* p = s.in.getWritePtr();
            if (b < n_blocks) memcpy(p, blocks + b * (long)transferSamples, sizeof(int16_t) * transferSamples);
            else memset(p, 0, sizeof(int16_t) * transferSamples);
            s.in.WriteDone();
        }
    });
    long written = 0;
    const long want = n_blocks * per_block;
    for (long sl = 0; sl < n_slots; sl++) {
        const float* p = s.out.getReadPtr();
        long n = (transferSamples / 2);
        if (written + n > want) n = want - written;
        memcpy(iq_out + 2 * written, p, sizeof(float) * 2 * n);
        written += n;
        s.out.ReadDone();
    }
    producer.join();
    turn_off(s);
    delete s.op;
    return written;
}

// Timed throughput of the reference loop, modelled on fx3handler_benchmark
// (unittest/benchmark_test.cpp:78-142): `n_instances` independent operator instances, each with its
// own rings, a producer that memcpy's pre-built blocks as fast as the ring accepts them and a
// consumer that drains output slots.  `warm_blocks` blocks per instance are processed before the
// clock starts, then `timed_blocks` per instance are timed.  `blocks` holds `n_distinct` blocks that
// are cycled.  Returns aggregate input samples per second.
double refr2iq_bench(const int16_t* blocks, long n_distinct, long warm_blocks, long timed_blocks, float gain, int d,
                     int tunebin, int rnd, int sideband, int variant, int n_instances, double* seconds_out) {
    if (d < 0 || d >= NDECIDX || n_instances <= 0 || n_distinct <= 0) return -1.0;
    const long ratio = 1L << d;
    warm_blocks = (warm_blocks + ratio - 1) / ratio * ratio;
    timed_blocks = (timed_blocks + ratio - 1) / ratio * ratio;
    std::vector<Session*> ss(n_instances);
    for (auto& p : ss) {
        p = new Session();
        setup(*p, gain, d, tunebin, rnd, sideband);
    }
    std::mutex mu;
    std::condition_variable cv;
    int warmed = 0;
    bool go = false;
    std::vector<std::thread> producers, consumers;
    std::vector<std::chrono::steady_clock::time_point> t_done(n_instances);
    std::chrono::steady_clock::time_point t_start;
    for (int i = 0; i < n_instances; i++) {
        Session& s = *ss[i];
        turn_on(s, variant);
        memset(s.in.peekWritePtr(-1), 0, sizeof(int16_t) * transferSamples);
        producers.emplace_back([&s, blocks, n_distinct, warm_blocks, timed_blocks, i] {
            long total = warm_blocks + timed_blocks;
            for (long b = 0; b < total; b++) {
                int16_t* p = s.in.getWritePtr();
                memcpy(p, blocks + ((b + i) % n_distinct) * (long)transferSamples, sizeof(int16_t) * transferSamples);
                s.in.WriteDone();
            }
        });
        consumers.emplace_back([&, i] {
            Session& s = *ss[i];
            volatile float sink = 0;
            for (long sl = 0; sl < warm_blocks / ratio; sl++) {
                const float* p = s.out.getReadPtr();
                sink = sink + p[0];
                s.out.ReadDone();
            }
            {
                std::unique_lock<std::mutex> lk(mu);
                if (++warmed == n_instances) {
                    t_start = std::chrono::steady_clock::now();
                    go = true;
                    cv.notify_all();
                }
            }
            for (long sl = 0; sl < timed_blocks / ratio; sl++) {
                const float* p = s.out.getReadPtr();
                sink = sink + p[0];
                s.out.ReadDone();
            }
            t_done[i] = std::chrono::steady_clock::now();
        });
    }
    for (auto& t : consumers) t.join();
    for (auto& t : producers) t.join();
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return go; });
    }
    auto t_end = t_done[0];
    for (auto& t : t_done) if (t > t_end) t_end = t;
    for (auto& p : ss) turn_off(*p);
    double sec = std::chrono::duration<double>(t_end - t_start).count();
    if (seconds_out) *seconds_out = sec;
    for (auto& p : ss) {
        delete p->op;
        delete p;
    }
    return (double)n_instances * (double)timed_blocks * (double)transferSamples / sec;
}

}  // extern "C"
