// cuda_r2iq.cpp -- ring-buffer adapter over the C ABI (see cuda_r2iq.h).  No sample arithmetic here.
#include <string>

#include "cuda_r2iq.h"
#ifdef R2IQ_USE_REFERENCE_HEADERS
#include "config.h"
#endif

#include <cstring>

namespace {
constexpr int kBlock = R2IQ_TRANSFER_SAMPLES;        // int16 samples per input slot
constexpr int kSlotComplex = R2IQ_TRANSFER_SAMPLES / 2;   // complex samples per output slot (EXT_BLOCKLEN)
}

cuda_r2iq::cuda_r2iq(int device, int slots_per_batch, bool fuse_fine_tune)
    : r2iqControlClass(), device_(device), slots_per_batch_(slots_per_batch < 1 ? 1 : slots_per_batch),
      fuse_fine_tune_(fuse_fine_tune) {}

cuda_r2iq::~cuda_r2iq() {
    if (run_.load()) TurnOff();
    if (stage_in_) r2iq_cuda_host_free(stage_in_);
    if (stage_out_) r2iq_cuda_host_free(stage_out_);
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
    cfg.chunk_blocks = 0;
    if (r2iq_cuda_create(&cfg, &handle_) != R2IQ_OK) {
        error_ = r2iq_cuda_last_error();   // no CPU fallback: the operator stays unusable and says why
        handle_ = nullptr;
        return;
    }
    const size_t max_blocks = (size_t)slots_per_batch_ << (NDECIDX - 1);
    if (!stage_in_) stage_in_ = (int16_t*)r2iq_cuda_host_alloc_on(sizeof(int16_t) * max_blocks * kBlock, device_);
    if (!stage_out_) stage_out_ = (float*)r2iq_cuda_host_alloc_on(sizeof(float) * 2 * (size_t)slots_per_batch_ * kSlotComplex, device_);
    if (!stage_in_ || !stage_out_) error_ = r2iq_cuda_last_error();
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
    if (!handle_ || !in_ || !out_ || !stage_in_ || !stage_out_) return;
    r2iq_cuda_set_decimate(handle_, mdecimation);
    r2iq_cuda_set_sideband(handle_, getSideband() ? 1 : 0);
    if (r2iq_cuda_turn_on(handle_) != R2IQ_OK) {
        error_ = r2iq_cuda_last_error();
        return;
    }
    this->r2iqOn = true;
    run_.store(true);
    in_->Start();
    out_->Start();
    thread_ = std::thread([this] { worker(); });
}

// fft_mt_r2iq::TurnOff (Core/fft_mt_r2iq.cpp:131-143)
void cuda_r2iq::TurnOff(void) {
    this->r2iqOn = false;
    run_.store(false);
    if (in_) in_->Stop();
    if (out_) out_->Stop();
    if (thread_.joinable()) thread_.join();
    if (handle_) r2iq_cuda_turn_off(handle_);
}

bool cuda_r2iq::IsOn(void) { return this->r2iqOn; }

void cuda_r2iq::worker() {
    const int d = r2iq_cuda_get_decimate(handle_);          // latched for this run, like impl.hpp:3
    const int blocks_per_slot = 1 << d;
    const int n_blocks = blocks_per_slot * slots_per_batch_;
    while (run_.load()) {
        // gather one batch of input slots in order (blocks in getReadPtr like the reference worker)
        for (int b = 0; b < n_blocks; b++) {
            const int16_t* src = in_->getReadPtr();
            if (!run_.load()) return;
            memcpy(stage_in_ + (size_t)b * kBlock, src, sizeof(int16_t) * kBlock);
            in_->ReadDone();
        }
        r2iq_cuda_set_rand(handle_, getRand() ? 1 : 0);     // sampled per batch (impl.hpp:40)
        if (r2iq_cuda_process(handle_, stage_in_, n_blocks, nullptr, stage_out_) != R2IQ_OK) {
            error_ = r2iq_cuda_last_error();
            run_.store(false);
            this->r2iqOn = false;
            return;
        }
        batches_.fetch_add(1);
        for (int s = 0; s < slots_per_batch_; s++) {
            float* dst = out_->getWritePtr();
            if (!run_.load()) return;
            memcpy(dst, stage_out_ + 2 * (size_t)s * kSlotComplex, sizeof(float) * 2 * kSlotComplex);
            out_->WriteDone();
        }
    }
}
