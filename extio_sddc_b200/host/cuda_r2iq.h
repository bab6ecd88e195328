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
// Data path: the two rings are page-locked in place (cudaHostRegister over each ring's single allocation), a
// submitter thread hands every input slot that is already there (up to 16) to r2iq_cuda_process_async straight from
// ring memory and returns the slots to the producer as soon as the copy engine has read them; a publisher thread copies
// finished results from the device straight into output-ring slots, in order.  Up to four submissions are in flight, so
// copies in both directions overlap the kernels; nothing is staged or memcpy'd on the host.
#pragma once
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "r2iq_cuda.h"
#include "r2iq_iface.h"

class cuda_r2iq : public r2iqControlClass {
public:
    // max_batch: most input slots per submission (1..16); fewer are taken whenever fewer are waiting, so a paced
    // producer sees one-block submissions and a backlog is worked off sixteen blocks at a time.
    // fuse_fine_tune: run RadioHandlerClass's residual mixer (Core/RadioHandler.cpp:33-37) on the GPU too.
    // setFreqOffset then hands the residual to the device NCO and returns 0, so the caller's `fc` stays 0
    // and its CPU mixer is skipped (RadioHandler.cpp:33 `if (fc != 0.0f)`).  1: the reference's float32 recurrence,
    // bit-identical stream (R2IQ_NCO_REFERENCE); 2: phases computed on the device, free in the kernel's store
    // epilogue, within the tolerance stated in include/r2iq_cuda.h (R2IQ_NCO_DIRECT).
    explicit cuda_r2iq(int device = -1, int max_batch = 16, int fuse_fine_tune = 0);
    ~cuda_r2iq() override;

    void Init(float gain, ringbuffer<int16_t>* input, ringbuffer<float>* obuffers) override;
    void TurnOn() override;
    void TurnOff(void) override;
    bool IsOn(void) override;
    float setFreqOffset(float offset) override;

    // not part of the reference interface
    // Take the randomizer and sideband flags from another control object: the non-virtual setters of r2iqControlClass
    // (updateRand, setSideband: Core/r2iq.h:22-26) land on whatever object the caller holds, so a wrapper that owns a
    // cuda_r2iq (host/fft_mt_r2iq_cuda.cpp) points this at itself.
    void follow(const r2iqControlClass* controls) { controls_ = controls; }
    const char* lastError() const { return error_.c_str(); }
    unsigned long long batches() const { return batches_.load(); }

private:
    struct Pending {
        uint64_t ticket;
        long blocks;
    };
    static constexpr int kQueue = 8;     // the handle keeps at most four submissions in flight

    void submit_loop();
    void publish_loop();
    void register_rings();
    void unregister_rings();
    void fail_stream();

    int device_;
    int max_batch_;
    int fuse_fine_tune_;
    const r2iqControlClass* controls_ = nullptr;
    r2iq_cuda_t handle_ = nullptr;
    ringbuffer<int16_t>* in_ = nullptr;
    ringbuffer<float>* out_ = nullptr;
    std::thread submitter_, publisher_;
    std::atomic<bool> run_{false};
    std::atomic<unsigned long long> batches_{0};
    // page-locked ring memory
    int16_t* in_base_ = nullptr;
    float* out_base_ = nullptr;
    size_t in_pitch_ = 0, out_pitch_ = 0;   // elements between slots
    int in_count_ = 0, out_count_ = 0;
    bool in_registered_ = false, out_registered_ = false;
    int in_writes_at_start_ = 0;
    // submissions handed from the submitter to the publisher, in order
    Pending queue_[kQueue];
    unsigned long head_ = 0, tail_ = 0;
    std::mutex qmu_;
    std::condition_variable qcv_;
    std::string error_;
};
