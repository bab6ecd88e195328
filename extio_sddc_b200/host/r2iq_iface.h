// r2iq_iface.h -- the operator interface the CUDA backend plugs into.
//
// Inside the reference tree build with -DR2IQ_USE_REFERENCE_HEADERS (and Core/ on the include
// path): the adapter then derives from the reference's own r2iqControlClass (Core/r2iq.h:16-48) and
// uses its ringbuffer<T> (Core/dsp/ringbuffer.h:136-205) unchanged.
//
// Stand-alone (tests in this repo, no reference tree on the GPU box) the same two types are
// provided here with the same public members and semantics, written from the interface contract:
// a fixed ring of equal slots, a blocking producer/consumer hand-off per slot, Stop() releasing
// every waiter, peek*(offset) addressing neighbours of the current slot.
#pragma once

#ifdef R2IQ_USE_REFERENCE_HEADERS
#include "r2iq.h"
#else

#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <vector>

#define NDECIDX 7   // Core/r2iq.h:5

template <typename T>
class ringbuffer {
public:
    explicit ringbuffer(int count = 64) : slots_(count) {}   // default_count, Core/dsp/ringbuffer.h:7

    void setBlockSize(int elements) {   // size in ELEMENTS, rounded up to 8 (ringbuffer.h:154-170)
        if (elements == block_) return;
        block_ = elements;
        pitch_ = (elements + 7) & ~7;
        store_.assign((size_t)slots_ * pitch_, T());
    }
    int getBlockSize() const { return block_; }

    T* getWritePtr() {   // blocks while the ring is full (unless stopped)
        std::unique_lock<std::mutex> lk(mu_);
        if (!stopped_ && (w_ + 1) % slots_ == r_) {
            ++full_;
            not_full_.wait(lk, [this] { return stopped_ || (w_ + 1) % slots_ != r_; });
        }
        return slot(w_);
    }
    void WriteDone() {
        std::lock_guard<std::mutex> lk(mu_);
        w_ = (w_ + 1) % slots_;
        ++writes_;
        not_empty_.notify_all();
    }
    const T* getReadPtr() {   // blocks while the ring is empty (unless stopped)
        std::unique_lock<std::mutex> lk(mu_);
        if (!stopped_ && r_ == w_) {
            ++empty_;
            not_empty_.wait(lk, [this] { return stopped_ || r_ != w_; });
        }
        return slot(r_);
    }
    void ReadDone() {
        std::lock_guard<std::mutex> lk(mu_);
        r_ = (r_ + 1) % slots_;
        not_full_.notify_all();
    }
    T* peekWritePtr(int offset) { return slot((w_ + slots_ + offset) % slots_); }
    T* peekReadPtr(int offset) { return slot((r_ + slots_ + offset) % slots_); }

    void Start() {
        std::lock_guard<std::mutex> lk(mu_);
        r_ = w_ = 0;
        stopped_ = false;
    }
    void Stop() {   // wake every waiter; indices are left "half full" like the reference (ringbuffer.h:66-74)
        std::lock_guard<std::mutex> lk(mu_);
        r_ = 0;
        w_ = slots_ / 2;
        stopped_ = true;
        not_full_.notify_all();
        not_empty_.notify_all();
    }
    int getFullCount() const { return full_; }
    int getEmptyCount() const { return empty_; }
    int getWriteCount() const { return writes_; }

private:
    T* slot(int i) { return store_.data() + (size_t)i * pitch_; }
    int slots_;
    int block_ = 0, pitch_ = 0;
    int r_ = 0, w_ = 0;
    int full_ = 0, empty_ = 0, writes_ = 0;
    bool stopped_ = false;
    std::vector<T> store_;
    std::mutex mu_;
    std::condition_variable not_empty_, not_full_;
};

// Same members as Core/r2iq.h:16-48.
class r2iqControlClass {
public:
    r2iqControlClass() {
        for (int i = 0; i < NDECIDX; i++) mratio[i] = 1 << i;
    }
    virtual ~r2iqControlClass() {}
    int getRatio() { return mratio[mdecimation]; }
    void updateRand(bool v) { randADC = v; }
    bool getRand() const { return randADC; }
    void setSideband(bool lsb) { sideband = lsb; }
    bool getSideband() const { return sideband; }
    void setDecimate(int dec) { mdecimation = dec; }
    virtual void Init(float, ringbuffer<int16_t>*, ringbuffer<float>*) {}
    virtual void TurnOn() { r2iqOn = true; }
    virtual void TurnOff(void) { r2iqOn = false; }
    virtual bool IsOn(void) { return r2iqOn; }
    virtual void DataReady(void) {}
    virtual float setFreqOffset(float) { return 0; }

protected:
    int mdecimation = 0;
    bool r2iqOn = false;
    int mratio[NDECIDX];

private:
    bool randADC = false;
    bool sideband = false;
};

#endif  // R2IQ_USE_REFERENCE_HEADERS
