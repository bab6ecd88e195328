// cuda_r2iq.cpp -- ring-buffer adapter over the C ABI (see cuda_r2iq.h).  No sample arithmetic here.
#include <string>

#include "cuda_r2iq.h"
#ifdef R2IQ_USE_REFERENCE_HEADERS
#include "config.h"
#endif

#include <chrono>
#include <cstring>

namespace {
constexpr int kBlock = R2IQ_TRANSFER_SAMPLES;             // int16 samples per input slot
constexpr int kSlotComplex = R2IQ_TRANSFER_SAMPLES / 2;   // complex samples per output slot (EXT_BLOCKLEN)
constexpr int kMaxBatch = 16;                             // input slots per submission (chunk_blocks of the handle)

// One `new T[max_count * aligned_block_size]` per ring (Core/dsp/ringbuffer.h:154-170): find its first slot, its
// slot pitch and its slot count through the public peek interface (slot i of the allocation is reached from the
// current read index by some offset; the lowest address is the base).
template <typename T>
void ring_extent(ringbuffer<T>* rb, T** base, size_t* pitch, int* count) {
    T* first = rb->peekReadPtr(0);
    *base = nullptr;
    *pitch = 0;
    *count = 0;
    if (!first || rb->getBlockSize() <= 0) return;     // no slots allocated yet
    T* lo = first;
    T* second_lowest = nullptr;
    int n = 1;
    for (; n < 4096; n++) {
        T* p = rb->peekReadPtr(n);
        if (p == first) break;
        if (p < lo) {
            second_lowest = lo;
            lo = p;
        } else if (!second_lowest || p < second_lowest) {
            second_lowest = p;
        }
    }
    *base = lo;
    *count = n;
    *pitch = (n > 1 && second_lowest) ? (size_t)(second_lowest - lo) : (size_t)rb->getBlockSize();
}
}  // namespace

cuda_r2iq::cuda_r2iq(int device, int max_batch, int fuse_fine_tune)
    : r2iqControlClass(), device_(device), max_batch_(max_batch < 1 ? 1 : (max_batch > kMaxBatch ? kMaxBatch : max_batch)),
      fuse_fine_tune_(fuse_fine_tune) {}

cuda_r2iq::~cuda_r2iq() {
    if (run_.load()) TurnOff();
    unregister_rings();
    if (handle_) r2iq_cuda_destroy(handle_);
}

// fft_mt_r2iq::Init (Core/fft_mt_r2iq.cpp:147-227): remember the rings, design and upload the filters.
void cuda_r2iq::Init(float gain, ringbuffer<int16_t>* input, ringbuffer<float>* obuffers) {
    in_ = input;
    out_ = obuffers;
    if (handle_) {
        r2iq_cuda_destroy(handle_);
        handle_ = nullptr;
    }
    r2iq_cuda_config cfg = {};
#ifdef R2IQ_USE_REFERENCE_HEADERS
    cfg.fftn = FFTN_R_ADC;       // inside the reference tree: whatever Core/config.h:49 says (8192 as shipped)
#endif
    cfg.device = device_;
    cfg.gain = gain;
    cfg.chunk_blocks = kMaxBatch;
    if (r2iq_cuda_create(&cfg, &handle_) != R2IQ_OK) {
        error_ = r2iq_cuda_last_error();   // no CPU fallback: the operator stays unusable and says why
        handle_ = nullptr;
        return;
    }
    if (fuse_fine_tune_ == 2) r2iq_cuda_set_fine_tune_mode(handle_, R2IQ_NCO_DIRECT);
}

// Page-lock the two rings (once per allocation: setBlockSize reallocates only when the size changes,
// Core/dsp/ringbuffer.h:154-170), so that the copy engines move data straight between ring slots and the device:
// no staging buffer, no extra memcpy per block (what Core/arch/linux/FX3handler.cpp:139-147 writes is what the GPU reads).
void cuda_r2iq::register_rings() {
    // input ring: RadioHandlerClass::Start turns the operator on BEFORE the USB layer sizes the input ring
    // (Core/RadioHandler.cpp:165-167), so on a first start there is nothing to lock yet; submit_loop calls again once
    // the first slot has arrived (no copy has touched the ring by then)
    int16_t* ib = nullptr;
    size_t ip = 0;
    int ic = 0;
    ring_extent(in_, &ib, &ip, &ic);
    if (ib && !(ib == in_base_ && ip == in_pitch_ && ic == in_count_)) {
        if (in_registered_) r2iq_cuda_host_unregister(in_base_);
        in_base_ = ib;
        in_pitch_ = ip;
        in_count_ = ic;
        // a ring that cannot be page-locked still works: the copies are then staged by the driver (slower, not wrong)
        in_registered_ = r2iq_cuda_host_register(ib, sizeof(int16_t) * ip * (size_t)ic) == R2IQ_OK;
    }
    float* ob = nullptr;
    size_t op = 0;
    int oc = 0;
    ring_extent(out_, &ob, &op, &oc);
    if (ob && !(ob == out_base_ && op == out_pitch_ && oc == out_count_)) {
        if (out_registered_) r2iq_cuda_host_unregister(out_base_);
        out_base_ = ob;
        out_pitch_ = op;
        out_count_ = oc;
        out_registered_ = r2iq_cuda_host_register(ob, sizeof(float) * op * (size_t)oc) == R2IQ_OK;
    }
}

void cuda_r2iq::unregister_rings() {
    if (in_registered_) r2iq_cuda_host_unregister(in_base_);
    if (out_registered_) r2iq_cuda_host_unregister(out_base_);
    in_registered_ = out_registered_ = false;
    in_base_ = nullptr;
    out_base_ = nullptr;
    in_pitch_ = out_pitch_ = 0;
    in_count_ = out_count_ = 0;
}

float cuda_r2iq::setFreqOffset(float offset) {
    if (!handle_) return 0.0f;
    r2iq_cuda_set_decimate(handle_, mdecimation);   // the residual scales with getRatio()
    const float residual = r2iq_cuda_set_freq_offset(handle_, offset);
    if (!fuse_fine_tune_) return residual;
    r2iq_cuda_set_fine_tune(handle_, residual);      // same restart rule as RadioHandler.cpp:292-297
    return 0.0f;
}

// fft_mt_r2iq::TurnOn (Core/fft_mt_r2iq.cpp:111-129)
void cuda_r2iq::TurnOn() {
    if (!handle_ || !in_ || !out_) return;
    r2iq_cuda_set_decimate(handle_, mdecimation);
    r2iq_cuda_set_sideband(handle_, (controls_ ? controls_ : this)->getSideband() ? 1 : 0);
    if (r2iq_cuda_turn_on(handle_) != R2IQ_OK) {
        error_ = r2iq_cuda_last_error();
        return;
    }
    this->r2iqOn = true;
    run_.store(true);
    in_->Start();
    out_->Start();
    register_rings();
    in_writes_at_start_ = in_->getWriteCount();
    head_ = tail_ = 0;
    submitter_ = std::thread([this] { submit_loop(); });
    publisher_ = std::thread([this] { publish_loop(); });
}

// fft_mt_r2iq::TurnOff (Core/fft_mt_r2iq.cpp:131-143)
void cuda_r2iq::TurnOff(void) {
    this->r2iqOn = false;
    run_.store(false);
    if (in_) in_->Stop();
    if (out_) out_->Stop();
    if (handle_) r2iq_cuda_turn_off(handle_);      // wakes a submitter waiting for a held slot, drains the device
    {
        std::lock_guard<std::mutex> lk(qmu_);
    }
    qcv_.notify_all();
    if (submitter_.joinable()) submitter_.join();
    if (publisher_.joinable()) publisher_.join();
}

bool cuda_r2iq::IsOn(void) { return this->r2iqOn; }

void cuda_r2iq::fail_stream() {
    error_ = r2iq_cuda_last_error();
    run_.store(false);
    this->r2iqOn = false;
    qcv_.notify_all();
}

// The worker loop of the reference (Core/fft_mt_r2iq_impl.hpp:15-33), input side: block until a slot is there, then
// take EVERY slot the producer has already published (up to kMaxBatch, up to the end of the ring's allocation) as one
// submission, straight from ring memory; hand the slots back as soon as the copy engine has read them.
void cuda_r2iq::submit_loop() {
    long consumed = 0;
    while (run_.load()) {
        const int16_t* src = in_->getReadPtr();          // blocks like the reference worker
        if (!run_.load()) return;
        if (!in_base_) register_rings();                 // first start: the ring was sized after TurnOn
        long avail = (long)(in_->getWriteCount() - in_writes_at_start_) - consumed;
        if (avail < 1) avail = 1;                        // getReadPtr returned: at least this one
        long n = avail < max_batch_ ? avail : max_batch_;
        const long slot_index = in_pitch_ ? (long)((src - in_base_) / (long)in_pitch_) : 0;
        if (slot_index + n > in_count_) n = in_count_ - slot_index;      // contiguous run only; the rest next time round
        if ((size_t)kBlock != in_pitch_) n = 1;          // a ring with padded slots: one slot per submission
        r2iq_cuda_set_rand(handle_, (controls_ ? controls_ : this)->getRand() ? 1 : 0);  // sampled per batch (impl.hpp:40)
        uint64_t ticket = 0;
        if (r2iq_cuda_process_async(handle_, src, n, nullptr, nullptr, &ticket) != R2IQ_OK) {
            if (run_.load()) fail_stream();
            return;
        }
        {
            std::lock_guard<std::mutex> lk(qmu_);
            queue_[head_ % kQueue] = {ticket, n};
            head_++;
        }
        qcv_.notify_all();
        if (r2iq_cuda_wait_input(handle_, ticket) != R2IQ_OK) {
            if (run_.load()) fail_stream();
            return;
        }
        for (long i = 0; i < n; i++) in_->ReadDone();    // the producer may refill them while the GPU computes
        consumed += n;
        batches_.fetch_add(1);
    }
}

// Output side (impl.hpp:100-148): results leave in submission order; 2^d input blocks fill one 32768-complex slot,
// copied from the device straight into the ring slot, which is published when its last piece has arrived.
void cuda_r2iq::publish_loop() {
    const int d = r2iq_cuda_get_decimate(handle_);           // latched for this run, like impl.hpp:3
    const long per_block = kSlotComplex >> d;                // complex samples one input block yields
    float* dst = nullptr;
    long filled = 0;                                         // complex samples already in the open output slot
    while (true) {
        Pending pd;
        {
            std::unique_lock<std::mutex> lk(qmu_);
            qcv_.wait(lk, [this] { return tail_ < head_ || !run_.load(); });
            if (tail_ >= head_) return;
            if (!run_.load()) return;
            pd = queue_[tail_ % kQueue];
        }
        long done = 0;
        const long total = pd.blocks * per_block;
        while (done < total) {
            if (!dst) {
                dst = out_->getWritePtr();                   // blocks while the consumer is behind
                if (!run_.load()) return;
                filled = 0;
            }
            long n = kSlotComplex - filled;
            if (n > total - done) n = total - done;
            if (r2iq_cuda_read(handle_, pd.ticket, done, n, dst + 2 * filled) != R2IQ_OK) {
                if (run_.load()) fail_stream();
                return;
            }
            done += n;
            filled += n;
            if (filled == kSlotComplex) {
                out_->WriteDone();
                dst = nullptr;
            }
        }
        r2iq_cuda_release(handle_, pd.ticket);
        {
            std::lock_guard<std::mutex> lk(qmu_);
            tail_++;
        }
    }
}
