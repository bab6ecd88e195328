// cuda_r2iq.h -- `cuda_r2iq : r2iqControlClass`, the B200 backend as a drop-in operator.
//
// Pass an instance as the third argument of RadioHandlerClass::Init (Core/RadioHandler.h:34); when that
// argument is null the reference creates `new fft_mt_r2iq()` (Core/RadioHandler.cpp:94-95).  The class
// holds no DSP: it moves ring-buffer slots to and from the C ABI of include/r2iq_cuda.h.
//
// Streaming contract kept from fft_mt_r2iq (Core/fft_mt_r2iq_impl.hpp:15-151):
//  * one worker thread blocks in inputbuffer->getReadPtr(), consumes 65536-sample slots in order and
//    calls ReadDone() for each; block b uses the last halfFft samples of block b-1 (carried by the
//    handle; zeros for the first block after TurnOn -- the reference reads uninitialised ring memory);
//  * 2^decimation input blocks fill one output slot of 32768 complex floats, then WriteDone();
//  * decimation and sideband are latched at TurnOn; tune bin and rand flag are picked up per batch;
//  * TurnOff stops both rings (waking any waiter) and joins the worker.
// Each batch is one fused-kernel launch over 2^decimation * slots_per_batch blocks.
#pragma once
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "r2iq_cuda.h"
#include "r2iq_iface.h"

class cuda_r2iq : public r2iqControlClass {
public:
    // fuse_fine_tune: run RadioHandlerClass's residual mixer (Core/RadioHandler.cpp:33-37) on the GPU too.
    // setFreqOffset then hands the residual to the device NCO and returns 0, so the caller's `fc` stays 0
    // and its CPU mixer is skipped (RadioHandler.cpp:33 `if (fc != 0.0f)`); the stream is bit-identical.
    explicit cuda_r2iq(int device = -1, int slots_per_batch = 1, bool fuse_fine_tune = false);
    ~cuda_r2iq() override;

    void Init(float gain, ringbuffer<int16_t>* input, ringbuffer<float>* obuffers) override;
    void TurnOn() override;
    void TurnOff(void) override;
    bool IsOn(void) override;
    float setFreqOffset(float offset) override;

    // not part of the reference interface
    const char* lastError() const { return error_.c_str(); }
    unsigned long long batches() const { return batches_.load(); }

private:
    void worker();

    int device_;
    int slots_per_batch_;
    bool fuse_fine_tune_;
    r2iq_cuda_t handle_ = nullptr;
    ringbuffer<int16_t>* in_ = nullptr;
    ringbuffer<float>* out_ = nullptr;
    std::thread thread_;
    std::atomic<bool> run_{false};
    std::atomic<unsigned long long> batches_{0};
    int16_t* stage_in_ = nullptr;    // pinned: one batch of input blocks
    float* stage_out_ = nullptr;     // pinned: one batch of output slots
    std::string error_;
};
