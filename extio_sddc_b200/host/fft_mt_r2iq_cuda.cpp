// fft_mt_r2iq_cuda.cpp -- the reference's OWN class `fft_mt_r2iq` (Core/fft_mt_r2iq.h:22-113), implemented over the
// B200 backend: a build that lists this file INSTEAD of Core/fft_mt_r2iq.cpp and its per-ISA workers
// (fft_mt_r2iq_def/avx/avx2/avx512/neon.cpp) runs every application of the reference on the GPU with no source edit --
// libsddc/libsddc.cpp:262 and SoapySDDC/Settings.cpp pass a null operator to RadioHandlerClass::Init, which then does
// `new fft_mt_r2iq()` (Core/RadioHandler.cpp:94-95) and gets this implementation.  Compiled inside the reference tree
// only (it needs Core/fft_mt_r2iq.h and the fftw3.h that header includes for its member types).
//
// Each member forwards to a `cuda_r2iq` kept beside the object (the class layout is the reference's, so there is no
// room inside it).  Control state set through the non-virtual setters of r2iqControlClass (setDecimate, updateRand,
// setSideband: Core/r2iq.h:22-28) lives in THIS object; the backend follows it: decimation is forwarded where the
// reference latches or uses it (TurnOn, setFreqOffset), rand and sideband are read through cuda_r2iq::follow().
//
// R2IQ_CUDA_FUSE_FINE_TUNE=0|1|2 (environment, read once per operator) selects where RadioHandlerClass's residual
// mixer runs (see cuda_r2iq.h); default 0 = on the CPU as the reference does.
#ifndef R2IQ_USE_REFERENCE_HEADERS
#error "fft_mt_r2iq_cuda.cpp is built inside the reference tree: compile with -DR2IQ_USE_REFERENCE_HEADERS"
#endif
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "cuda_r2iq.h"
#include "fft_mt_r2iq.h"

namespace {
std::mutex g_mu;
std::unordered_map<const fft_mt_r2iq*, cuda_r2iq*> g_backend;

cuda_r2iq* backend_of(const fft_mt_r2iq* self) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_backend.find(self);
    return it == g_backend.end() ? nullptr : it->second;
}
}  // namespace

// The base-class constructor lives in the file this one replaces (Core/fft_mt_r2iq.cpp:26-37): decimation 0, ratio table
// 1, 2, 4 ... 64, operator off, randomizer and sideband flags clear.
r2iqControlClass::r2iqControlClass() {
    r2iqOn = false;
    mdecimation = 0;
    for (int d = 0; d < NDECIDX; d++) mratio[d] = 1 << d;
    updateRand(false);
    setSideband(false);
}

fft_mt_r2iq::fft_mt_r2iq() : r2iqControlClass() {
    // constructor defaults of the reference (Core/fft_mt_r2iq.cpp:39-49); the members behind them are unused here
    inputbuffer = nullptr;
    outputbuffer = nullptr;
    bufIdx = 0;
    lastThread = nullptr;
    GainScale = 0.0f;
    mtunebin = halfFft / 4;
    for (int d = 0; d < NDECIDX; d++) mfftdim[d] = halfFft >> d;
    filterHw = nullptr;
    plan_t2f_r2c = nullptr;
    plan_f2t_c2c = nullptr;
    for (int d = 0; d < NDECIDX; d++) plans_f2t_c2c[d] = nullptr;
    processor_count = 1;
    for (int t = 0; t < N_MAX_R2IQ_THREADS; t++) threadArgs[t] = nullptr;
    const char* fuse = getenv("R2IQ_CUDA_FUSE_FINE_TUNE");
    cuda_r2iq* be = new cuda_r2iq(-1, 16, fuse ? atoi(fuse) : 0);
    be->follow(this);
    std::lock_guard<std::mutex> lk(g_mu);
    g_backend[this] = be;
}

fft_mt_r2iq::~fft_mt_r2iq() {
    cuda_r2iq* be = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_backend.find(this);
        if (it != g_backend.end()) {
            be = it->second;
            g_backend.erase(it);
        }
    }
    delete be;
}

void fft_mt_r2iq::Init(float gain, ringbuffer<int16_t>* buffers, ringbuffer<float>* obuffers) {
    inputbuffer = buffers;
    outputbuffer = obuffers;
    GainScale = gain;
    cuda_r2iq* be = backend_of(this);
    be->Init(gain, buffers, obuffers);
    if (be->lastError()[0]) fprintf(stderr, "fft_mt_r2iq (CUDA): %s\n", be->lastError());   // there is no CPU path to fall back to
}

float fft_mt_r2iq::setFreqOffset(float offset) {
    cuda_r2iq* be = backend_of(this);
    be->setDecimate(mdecimation);          // the residual scales with getRatio() (Core/fft_mt_r2iq.cpp:101-109)
    return be->setFreqOffset(offset);
}

void fft_mt_r2iq::TurnOn() {
    cuda_r2iq* be = backend_of(this);
    be->setDecimate(mdecimation);
    be->TurnOn();
    this->r2iqOn = be->IsOn();
}

void fft_mt_r2iq::TurnOff(void) {
    this->r2iqOn = false;
    backend_of(this)->TurnOff();
}

bool fft_mt_r2iq::IsOn(void) { return backend_of(this)->IsOn(); }
