// GPU test of the drop-in adapter `cuda_r2iq : r2iqControlClass` through the ring-buffer contract,
// with mock producers in the style of the reference's unittest (no hardware):
//   R2IQ count   unittest/core_test.cpp:149-174   every callback slot carries 32768 complex samples for all rates
//   Continuity   the stream delivered through the rings equals one direct C-ABI call over the same blocks
//   TuneTest     unittest/core_test.cpp:176-197   retune every few ms while streaming must not break the stream
//   Start/stop   unittest/stability_test.cpp:255-301 repeated TurnOn/TurnOff cycles over all rates
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "cuda_r2iq.h"

static int failures = 0;
#define CHECK(c) do { if (!(c)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); failures++; } } while (0)

static void make_block(std::vector<int16_t>& v, long first_sample) {
    for (int i = 0; i < R2IQ_TRANSFER_SAMPLES; i++) {
        long n = first_sample + i;
        v[i] = (int16_t)(8000.0 * sin(2.0 * M_PI * 0.13 * n) + 3000.0 * sin(2.0 * M_PI * 0.127 * n + 1.0) + (n * 7919 % 257) - 128);
    }
}

// feed n_blocks through the rings, return everything the operator published
// `prebuilt` (optional): 16 pre-generated blocks cycled by memcpy, like fx3handler_benchmark
// (unittest/benchmark_test.cpp:78-107), so that the producer is not the bottleneck of a throughput check.
// `discard`: the consumer only touches every slot (first and last sample summed into the returned vector's single
// element), like the counting callbacks of unittest/benchmark_test.cpp -- a throughput check must not be bounded by
// this test allocating and page-faulting gigabytes.
static std::vector<float> stream(cuda_r2iq& op, ringbuffer<int16_t>& in, ringbuffer<float>& out, int d, int n_blocks,
                                 bool retune, const std::vector<int16_t>* prebuilt = nullptr, bool hammer = false,
                                 bool discard = false) {
    op.setDecimate(d);
    op.TurnOn();
    CHECK(op.IsOn());
    std::thread prod([&] {
        std::vector<int16_t> blk(R2IQ_TRANSFER_SAMPLES);
        for (int b = 0; b < n_blocks; b++) {
            const int16_t* src = blk.data();
            if (prebuilt) src = prebuilt->data() + (size_t)(b % 16) * R2IQ_TRANSFER_SAMPLES;
            else make_block(blk, (long)b * R2IQ_TRANSFER_SAMPLES);
            int16_t* p = in.getWritePtr();
            memcpy(p, src, sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES);
            in.WriteDone();
        }
    });
    std::thread tuner;
    std::atomic<bool> tuning{retune};
    if (retune) tuner = std::thread([&] {
        float f = 1e3f;
        while (tuning.load()) {
            op.setFreqOffset(f / 32e6f);
            f += 377e3f;
            if (f > 15e6f) f = 1e3f;
            if (hammer) std::this_thread::yield();   // retune as fast as the thread can: any launch that reads the tune
            else std::this_thread::sleep_for(std::chrono::milliseconds(2));   // bin twice would pair tables with another bin
        }
    });
    const int n_slots = n_blocks >> d;
    std::vector<float> got(discard ? 1 : (size_t)n_slots * R2IQ_TRANSFER_SAMPLES);
    for (int s = 0; s < n_slots; s++) {
        const float* p = out.getReadPtr();
        if (discard) got[0] += p[0] + p[R2IQ_TRANSFER_SAMPLES - 1];
        else memcpy(got.data() + (size_t)s * R2IQ_TRANSFER_SAMPLES, p, sizeof(float) * R2IQ_TRANSFER_SAMPLES);
        out.ReadDone();
    }
    prod.join();
    tuning.store(false);
    if (tuner.joinable()) tuner.join();
    op.TurnOff();
    CHECK(!op.IsOn());
    return got;
}

int main() {
    ringbuffer<int16_t> in;
    ringbuffer<float> out;
    in.setBlockSize(R2IQ_TRANSFER_SAMPLES);                 // unittest/core_test.cpp:73
    out.setBlockSize(32768 * 2 * sizeof(float));            // Core/RadioHandler.cpp:166 (elements, oversized)
    cuda_r2iq op;
    const float gain = 1.08e-8f;
    op.Init(gain, &in, &out);
    if (op.lastError()[0]) { printf("FAIL Init: %s\n", op.lastError()); return 1; }
    op.updateRand(false);
    op.setFreqOffset(0.25f);

    // direct C-ABI reference for continuity
    r2iq_cuda_config cfg = {-1, gain, 0};
    r2iq_cuda_t h = nullptr;
    CHECK(r2iq_cuda_create(&cfg, &h) == R2IQ_OK);

    for (int d = 0; d <= 5; d++) {
        const int n_blocks = 64;
        std::vector<float> got = stream(op, in, out, d, n_blocks, false);
        CHECK(got.size() == (size_t)(n_blocks >> d) * 65536);       // 32768 complex per slot
        std::vector<int16_t> all((size_t)n_blocks * R2IQ_TRANSFER_SAMPLES), blk(R2IQ_TRANSFER_SAMPLES);
        for (int b = 0; b < n_blocks; b++) {
            make_block(blk, (long)b * R2IQ_TRANSFER_SAMPLES);
            memcpy(all.data() + (size_t)b * R2IQ_TRANSFER_SAMPLES, blk.data(), sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES);
        }
        std::vector<float> ref(got.size());
        r2iq_cuda_set_decimate(h, d);
        r2iq_cuda_set_freq_offset(h, 0.25f);
        r2iq_cuda_turn_on(h);
        CHECK(r2iq_cuda_process(h, all.data(), n_blocks, nullptr, ref.data()) == R2IQ_OK);
        r2iq_cuda_turn_off(h);
        CHECK(memcmp(ref.data(), got.data(), sizeof(float) * got.size()) == 0);
        double e = 0;
        for (float v : got) e += (double)v * v;
        CHECK(e > 0);
        printf("d=%d: %zu complex through the rings, identical to the direct call, %llu batches\n", d, got.size() / 2, op.batches());
    }
    {   // real-time margin of the live path: one launch per output slot, pageable ring slots, two PCIe copies
        std::vector<int16_t> pre((size_t)16 * R2IQ_TRANSFER_SAMPLES), one(R2IQ_TRANSFER_SAMPLES);
        for (int b = 0; b < 16; b++) {
            make_block(one, (long)b * R2IQ_TRANSFER_SAMPLES);
            memcpy(pre.data() + (size_t)b * R2IQ_TRANSFER_SAMPLES, one.data(), sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES);
        }
        for (int d : {0, 5}) {
            const int n_blocks = 16384;
            auto t0 = std::chrono::steady_clock::now();
            std::vector<float> got = stream(op, in, out, d, n_blocks, false, &pre, false, true);
            double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            double msps = (double)n_blocks * R2IQ_TRANSFER_SAMPLES / sec / 1e6;
            printf("live path d=%d: %.0f Msps through the rings (RX-888 full rate is 128 Msps), %llu submissions\n", d, msps, op.batches());
            // round 1 (one blocking launch per output slot, two host memcpys per block): 365 / 4769 Msps.  The producer
            // and the consumer of this test are single memcpy threads; thresholds sit well below what was measured
            // (profiles/r02_live_path.txt) so that a slower host does not fail the suite.
            CHECK(msps > (d == 0 ? 1500.0 : 3000.0));   // measured 3514 / 6508-8388 (d = 5 is bound by this test's producer thread)
        }
    }
    {   // added latency at the RX-888's real rate: the producer publishes one block every 512 us, the consumer stamps
        // the arrival of the output slot that carries it (d = 0: one slot per block)
        const int n_blocks = 400;
        std::vector<double> t_in(n_blocks), t_out(n_blocks);
        op.setDecimate(0);
        op.TurnOn();
        auto t0 = std::chrono::steady_clock::now();
        auto now = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
        std::thread prod([&] {
            for (int b = 0; b < n_blocks; b++) {
                const double due = (b + 1) * (65536.0 / 128e6);        // the block is complete when its last sample arrives
                while (now() < due) std::this_thread::yield();
                int16_t* p = in.getWritePtr();
                memset(p, 0x11, sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES);
                t_in[b] = now();
                in.WriteDone();
            }
        });
        for (int s = 0; s < n_blocks; s++) {
            (void)out.getReadPtr();
            t_out[s] = now();
            out.ReadDone();
        }
        prod.join();
        op.TurnOff();
        double mean = 0, worst = 0;
        for (int b = 50; b < n_blocks; b++) {          // skip the start-up transient (first launches build tables)
            const double l = t_out[b] - t_in[b];
            mean += l;
            if (l > worst) worst = l;
        }
        mean /= (n_blocks - 50);
        const double blk = 65536.0 / 128e6;
        printf("paced 128 Msps producer, d=0: added latency mean %.0f us (%.2f blocks), worst %.0f us (%.2f blocks)\n", mean * 1e6,
               mean / blk, worst * 1e6, worst / blk);
        CHECK(mean < 2.0 * blk);
    }
    {   // retune while streaming
        std::vector<float> got = stream(op, in, out, 2, 128, true);
        bool finite = true;
        for (float v : got) finite = finite && std::isfinite(v);
        CHECK(finite);
        printf("retune-while-streaming: %zu complex, all finite\n", got.size() / 2);
        for (int d : {0, 2, 5}) {   // the same with a retune storm (regression: tables and kernel parameter from two different bins)
            std::vector<float> g2 = stream(op, in, out, d, 512, true, nullptr, true);
            bool fin = true;
            double e = 0;
            for (float v : g2) { fin = fin && std::isfinite(v); e += (double)v * v; }
            CHECK(fin);
            CHECK(e > 0 && e < 1e6);
            printf("retune storm d=%d: %zu complex, finite=%d energy=%.3g\n", d, g2.size() / 2, (int)fin, e);
        }
        op.setFreqOffset(0.25f);
    }
    for (int cycle = 0; cycle < 21; cycle++) {   // unittest/stability_test.cpp:255-301: 20+ start/stop cycles over all the rates, stopping mid-stream
        int d = cycle % 7;
        op.setDecimate(d);
        op.TurnOn();
        std::thread prod([&] {
            for (int b = 0; b < (2 << d); b++) {
                int16_t* p = in.getWritePtr();
                memset(p, 0x5A, sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES);   // MockFx3Handler fill
                in.WriteDone();
            }
        });
        (void)out.getReadPtr();
        out.ReadDone();
        prod.join();
        op.TurnOff();
    }
    printf("start/stop cycles done\n");
    {   // fused fine-tune: setFreqOffset returns 0 to the caller and the device applies the residual NCO
        cuda_r2iq fused(-1, 16, 1);
        fused.Init(gain, &in, &out);
        fused.updateRand(false);
        const float r = fused.setFreqOffset(0.2503f);
        CHECK(r == 0.0f);
        std::vector<float> a = stream(fused, in, out, 1, 32, false);
        op.setFreqOffset(0.2503f);
        std::vector<float> b = stream(op, in, out, 1, 32, false);
        op.setFreqOffset(0.25f);
        CHECK(a.size() == b.size());
        double ea = 0, eb = 0, diff = 0;
        for (size_t i = 0; i < a.size(); i++) { ea += (double)a[i] * a[i]; eb += (double)b[i] * b[i]; diff += (double)(a[i] - b[i]) * (a[i] - b[i]); }
        CHECK(fabs(ea - eb) < 1e-4 * eb);   // a pure rotation keeps the energy ...
        CHECK(diff > 1e-3 * eb);            // ... but changes the samples
        printf("fused fine-tune mixer: energy kept (%.6g vs %.6g), samples rotated\n", ea, eb);
        // the direct-phase form (mode 2) gives the same stream within its stated tolerance (1e-6 rel-L2 to the ideal
        // mixer; the reference-bit form differs from the ideal by its own float32 rounding)
        cuda_r2iq direct(-1, 16, 2);
        direct.Init(gain, &in, &out);
        direct.updateRand(false);
        CHECK(direct.setFreqOffset(0.2503f) == 0.0f);
        std::vector<float> c = stream(direct, in, out, 1, 32, false);
        CHECK(c.size() == a.size());
        double dd = 0;
        for (size_t i = 0; i < a.size(); i++) dd += (double)(a[i] - c[i]) * (a[i] - c[i]);
        CHECK(sqrt(dd / ea) < 1e-3);   // the reference-bit form's own float32 frequency error (tests/test_gpu_parity.py)
        printf("direct-phase mixer vs reference-bit mixer: rel-L2 %.3g\n", sqrt(dd / ea));
    }
    r2iq_cuda_destroy(h);
    printf(failures ? "FAILED %d\n" : "ALL PASS\n", failures);
    return failures ? 1 : 0;
}
