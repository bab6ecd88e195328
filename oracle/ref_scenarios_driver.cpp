// oracle/ref_scenarios_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The reference's application-level robustness scenarios, restated at RadioHandlerClass level with cuda_r2iq as the
// operator.  RadioHandler.cpp, config.cpp, the radio models, the CPU mixer and the reference's own
// fft_mt_r2iq are the UNMODIFIED sources compiled in place (oracle/Makefile, FFTN_R_ADC 8192 as shipped); the USB layer
// is a mock in the style of unittest/stability_test.cpp:70-160 that plays a deterministic capture.
//
//   stress      unittest/stability_test.cpp:255-303 StartStopStressTest: 20 Start/Stop cycles, srate_idx = i % 5,
//               100...500 ms each, ~1 ms per USB transfer.  On top of "callbacks arrived": every slot has EXT_BLOCKLEN
//               samples and the first slots of every cycle equal what the reference operator yields for that rate
//               (rel-L2 <= 1e-5, max abs <= 1e-4 of full scale; the IQ of input block 0 excluded, see ref_radio_driver.cpp).
//   continuity  stability_test.cpp:309-365 DataContinuityTest: srate_idx 3 for N seconds (15 by default, argv), a
//               transfer every 1.024 ms (64 Msps real time).  The capture is one tone, so the IQ is one complex
//               exponential: every output sample must continue the previous one (also across slots and submissions);
//               a dropped, repeated or reordered input block anywhere on the GPU path breaks the phase.
//   switching   unittest/benchmark_test.cpp:601-671 RateSwitchingBenchmark: 10 cycles through the five rates,
//               Start -> first callback -> 50 ms -> Stop; switch-time statistics for cuda_r2iq and, beside them, for
//               the reference operator (2 cycles).
// Prints PASS/FAIL lines; exit code 0 when all pass.   usage: ref_radio_scenarios [continuity_seconds=15]
#include <atomic>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "FX3Class.h"
#include "RadioHandler.h"
#include "config.h"
#include "cuda_r2iq.h"

using namespace std::chrono;

namespace {

constexpr int kPeriodBlocks = 8;          // the capture repeats every 8 transfers (tone frequencies are multiples of fs / 2^19)
constexpr double kFs = 64000000.0;

struct Capture {
    std::vector<int16_t> blocks;          // kPeriodBlocks transfers
    explicit Capture(bool single_tone) : blocks((size_t)kPeriodBlocks * transferSamples) {
        const double unit = kFs / ((double)kPeriodBlocks * transferSamples);
        const double f1 = round(7.3e6 / unit) * unit, f2 = round(0.27 * kFs / unit) * unit, f3 = round(0.04 * kFs / unit) * unit;
        for (size_t n = 0; n < blocks.size(); n++) {
            const double t = (double)n / kFs;
            double v = 12000.0 * sin(2.0 * M_PI * f1 * t);
            if (!single_tone) v += 5000.0 * sin(2.0 * M_PI * f2 * t + 0.4) + 2000.0 * sin(2.0 * M_PI * f3 * t + 1.1) + (double)((n * 7919UL) % 613UL) - 306.0;
            blocks[n] = (int16_t)lrint(v);
        }
    }
};

class MockUsb : public fx3class {
public:
    MockUsb(const Capture& cap, int pace_us) : cap_(cap), pace_us_(pace_us) {}
    ~MockUsb() override { StopStream(); }
    bool Open() override { return true; }
    bool Control(FX3Command, uint8_t) override { return true; }
    bool Control(FX3Command, uint32_t) override { return true; }
    bool Control(FX3Command, uint64_t) override { return true; }
    bool SetArgument(uint16_t, uint16_t) override { return true; }
    bool GetHardwareInfo(uint32_t* data) override {
        const uint8_t d[4] = {0, FIRMWARE_VER_MAJOR, FIRMWARE_VER_MINOR, 0};     // no radio: DummyRadio, as in the reference tests
        memcpy(data, d, 4);
        return true;
    }
    bool ReadDebugTrace(uint8_t*, uint8_t) override { return false; }
    bool Enumerate(unsigned char&, char*) override { return true; }
    void StartStream(ringbuffer<int16_t>& input, int) override {
        input.setBlockSize(transferSamples);
        run_ = true;
        seq_ = 0;
        thread_ = std::thread([&input, this] {
            auto next = steady_clock::now();
            while (run_) {
                int16_t* p = input.getWritePtr();
                if (!run_) break;
                const uint32_t s = seq_.fetch_add(1);
                memcpy(p, cap_.blocks.data() + (size_t)(s % kPeriodBlocks) * transferSamples, transferSamples * sizeof(int16_t));
                input.WriteDone();
                if (pace_us_ > 0) {
                    next += microseconds(pace_us_);
                    std::this_thread::sleep_until(next);
                }
            }
        });
    }
    void StopStream() override {
        run_ = false;
        if (thread_.joinable()) thread_.join();
    }
    uint32_t sequence() const { return seq_.load(); }

private:
    const Capture& cap_;
    int pace_us_;
    std::thread thread_;
    std::atomic<bool> run_{false};
    std::atomic<uint32_t> seq_{0};
};

struct Sink {
    std::mutex mu;
    std::atomic<bool> active{false};
    std::atomic<uint64_t> callbacks{0};
    std::atomic<bool> bad_len{false};
    size_t keep_slots = 0;                  // copy the first keep_slots slots of a run
    std::vector<float> iq;
    // continuity
    bool check_phase = false;
    std::complex<double> rot{0, 0}, last{0, 0};
    double amp = 0;
    uint64_t samples = 0, breaks = 0;
    double worst = 0;
    std::atomic<long long> first_ns{0};     // steady-clock time of the first callback of a run
};

void on_iq(void* ctx, const float* data, uint32_t len) {
    Sink* s = (Sink*)ctx;
    if (!s->active.load()) return;
    if (len != EXT_BLOCKLEN) s->bad_len = true;
    const uint64_t idx = s->callbacks.fetch_add(1);
    if (idx == 0) s->first_ns = duration_cast<nanoseconds>(steady_clock::now().time_since_epoch()).count();
    if (idx < s->keep_slots) {
        std::lock_guard<std::mutex> lk(s->mu);
        s->iq.insert(s->iq.end(), data, data + 2 * (size_t)len);
    }
    if (!s->check_phase || idx < 2) return;               // slots 0-1: start-up transient of the zero history
    const std::complex<float>* z = (const std::complex<float>*)data;
    if (idx == 2) {                                       // measure the per-sample rotation and the amplitude once
        std::complex<double> acc = 0;
        double p = 0;
        for (uint32_t n = 1; n < len; n++) {
            acc += std::complex<double>(z[n]) * std::conj(std::complex<double>(z[n - 1]));
            p += std::norm(std::complex<double>(z[n]));
        }
        s->rot = acc / std::abs(acc);
        s->amp = sqrt(p / (len - 1));
        s->last = z[len - 1];
        return;
    }
    std::complex<double> prev = s->last;
    for (uint32_t n = 0; n < len; n++) {
        const std::complex<double> cur(z[n]);
        const double dev = std::abs(cur - prev * s->rot) / s->amp;
        if (dev > s->worst) s->worst = dev;
        if (dev > 5e-3) s->breaks++;      // a lost, repeated or swapped transfer moves the phase by ~0.2 cycle (error ~1)
        prev = cur;
    }
    s->last = prev;
    s->samples += len;
}

struct Radio {
    MockUsb* usb;
    RadioHandlerClass* radio;
    cuda_r2iq* op;
    Sink sink;
    Radio(const Capture& cap, int pace_us, bool cuda) {
        adcnominalfreq = (uint32_t)kFs;
        usb = new MockUsb(cap, pace_us);
        radio = new RadioHandlerClass();
        op = cuda ? new cuda_r2iq() : nullptr;
        radio->Init(usb, on_iq, op, &sink);
        radio->UptRand(false);
        radio->UpdatemodeRF(HFMODE);
        radio->TuneLO(7000000ull);
    }
    bool ok() const { return !op || !op->lastError()[0]; }
    void start(int srate_idx) {
        sink.callbacks = 0;
        sink.first_ns = 0;
        sink.active = true;
        radio->Start(srate_idx);
        radio->TuneLO(7000000ull);
    }
    void stop() {
        sink.active = false;
        radio->Stop();
    }
};

struct Err {
    double rel_l2, max_abs;
};
Err compare(const float* a, const float* b, size_t n) {
    double num = 0, den = 0, mx = 0;
    for (size_t i = 0; i < n; i++) {
        const double d = (double)a[i] - b[i];
        num += d * d;
        den += (double)b[i] * b[i];
        if (fabs(d) > mx) mx = fabs(d);
    }
    return {den > 0 ? sqrt(num / den) : sqrt(num), mx};
}

int g_failures = 0;
void verdict(bool ok, const char* fmt, ...) {
    char msg[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(msg, sizeof(msg), fmt, ap);
    va_end(ap);
    printf("%s %s\n", ok ? "PASS" : "FAIL", msg);
    fflush(stdout);
    if (!ok) g_failures++;
}

constexpr size_t kKeep = 3;   // slots compared per cycle

// the reference operator's first slots at every rate, same capture, same pacing
std::vector<std::vector<float>> reference_slots(const Capture& cap) {
    std::vector<std::vector<float>> ref(5);
    Radio r(cap, 1000, false);
    for (int idx = 0; idx < 5; idx++) {
        r.sink.iq.clear();
        r.sink.keep_slots = kKeep;
        r.start(idx);
        for (int ms = 0; ms < 5000 && r.sink.callbacks.load() < kKeep; ms++) std::this_thread::sleep_for(milliseconds(1));
        r.stop();
        std::lock_guard<std::mutex> lk(r.sink.mu);
        ref[idx] = r.sink.iq;
    }
    return ref;
}

void stress(const Capture& cap) {
    const std::vector<std::vector<float>> ref = reference_slots(cap);
    Radio r(cap, 1000, true);
    if (!r.ok()) return verdict(false, "stress: cuda_r2iq Init: %s", r.op->lastError());
    const double full_scale = 32768.0 * BBRF103_GAINFACTOR * 1024.0;
    int cycles_ok = 0;
    double worst_l2 = 0, worst_abs = 0;
    uint64_t min_callbacks = ~0ull;
    for (int i = 0; i < 20; i++) {
        const int idx = i % 5;
        const int d = 4 - idx;
        r.sink.iq.clear();
        r.sink.keep_slots = kKeep;
        r.start(idx);
        std::this_thread::sleep_for(milliseconds(100 + (i % 5) * 100));
        r.stop();
        const uint64_t cb = r.sink.callbacks.load();
        if (cb < min_callbacks) min_callbacks = cb;
        std::lock_guard<std::mutex> lk(r.sink.mu);
        const size_t have = std::min(r.sink.iq.size(), ref[idx].size());
        const size_t skip = 2 * (size_t)(EXT_BLOCKLEN >> d);
        bool good = cb > 0 && !r.sink.bad_len.load() && have > skip && !r.op->lastError()[0];
        if (good) {
            const Err e = compare(r.sink.iq.data() + skip, ref[idx].data() + skip, have - skip);
            worst_l2 = std::max(worst_l2, e.rel_l2);
            worst_abs = std::max(worst_abs, e.max_abs / full_scale);
            good = e.rel_l2 <= 1e-5 && e.max_abs / full_scale <= 1e-4;
        }
        if (good) cycles_ok++;
        else printf("  cycle %d (srate_idx %d): callbacks=%llu compared=%zu err=%s\n", i, idx, (unsigned long long)cb, have, r.op->lastError());
    }
    verdict(cycles_ok == 20, "stress: %d/20 start/stop cycles over 5 rates delivered verified IQ (fewest callbacks in a cycle %llu, worst rel_l2 %.3e, max_abs/fs %.3e)",
            cycles_ok, (unsigned long long)min_callbacks, worst_l2, worst_abs);
}

// worst single-step deviation with the reference operator in the same harness: the floor that the reference's own CPU
// fine-tune mixer (float32 phase recurrence, Core/pffft/pf_mixer.cpp) sets, printed beside the cuda_r2iq figure
double continuity_floor(const Capture& tone, int seconds_) {
    Radio r(tone, 1024, false);
    r.sink.check_phase = true;
    r.start(3);
    std::this_thread::sleep_for(seconds(seconds_));
    r.stop();
    return r.sink.worst;
}

void continuity(const Capture& tone, int seconds_) {
    const double floor_ref = continuity_floor(tone, seconds_ < 3 ? seconds_ : 3);
    Radio r(tone, 1024, true);
    if (!r.ok()) return verdict(false, "continuity: cuda_r2iq Init: %s", r.op->lastError());
    r.sink.check_phase = true;
    r.sink.keep_slots = 0;
    r.start(3);
    uint32_t last_seq = 0, backwards = 0;
    const auto t0 = steady_clock::now();
    while (steady_clock::now() - t0 < seconds(seconds_)) {
        std::this_thread::sleep_for(milliseconds(100));
        const uint32_t s = r.usb->sequence();
        if (s < last_seq) backwards++;
        last_seq = s;
    }
    const uint32_t produced = r.usb->sequence();
    const uint64_t cb = r.sink.callbacks.load();
    r.stop();
    const uint64_t expect = produced / 2;               // srate_idx 3 -> decimate 1: two transfers per slot
    const bool ok = produced > 100 && backwards == 0 && !r.sink.bad_len.load() && r.sink.breaks == 0 && r.sink.samples > 0 &&
                    cb + 24 >= expect && !r.op->lastError()[0];
    verdict(ok, "continuity: %d s at 64 Msps real time, %u transfers in, %llu of %llu slots out, %llu samples phase-checked, %llu breaks, "
                "worst step error %.2e of amplitude (reference operator in the same harness: %.2e)",
            seconds_, produced, (unsigned long long)cb, (unsigned long long)expect, (unsigned long long)r.sink.samples,
            (unsigned long long)r.sink.breaks, r.sink.worst, floor_ref);
}

struct SwitchStats {
    double mn = 1e9, mx = 0, sum = 0, sq = 0;
    int n = 0, timeouts = 0;
};
SwitchStats switching_run(const Capture& cap, bool cuda, int cycles) {
    SwitchStats st;
    Radio r(cap, 1000, cuda);
    for (int c = 0; c < cycles; c++) {
        for (int idx = 0; idx < 5; idx++) {
            r.sink.keep_slots = 0;
            const auto t0 = steady_clock::now();
            r.start(idx);
            int waited = 0;
            while (r.sink.first_ns.load() == 0 && waited < 50000) {
                std::this_thread::sleep_for(microseconds(100));
                waited++;
            }
            const long long first = r.sink.first_ns.load();
            if (first == 0) st.timeouts++;
            const double ms = (first - duration_cast<nanoseconds>(t0.time_since_epoch()).count()) * 1e-6;
            std::this_thread::sleep_for(milliseconds(50));
            r.stop();
            if (first == 0) continue;
            st.mn = std::min(st.mn, ms);
            st.mx = std::max(st.mx, ms);
            st.sum += ms;
            st.sq += ms * ms;
            st.n++;
        }
    }
    return st;
}

void switching(const Capture& cap) {
    const SwitchStats g = switching_run(cap, true, 10), c = switching_run(cap, false, 2);
    auto line = [](const char* who, const SwitchStats& s) {
        const double avg = s.n ? s.sum / s.n : 0, sd = s.n ? sqrt(std::max(0.0, s.sq / s.n - avg * avg)) : 0;
        printf("  %-22s switches %3d  min %7.2f ms  max %7.2f ms  avg %7.2f ms  stddev %6.2f ms\n", who, s.n, s.mn, s.mx, avg, sd);
    };
    printf("Rate Switch Time Statistics (Start -> first callback; a slot needs 2^decimate transfers of ~1 ms):\n");
    line("cuda_r2iq", g);
    line("fft_mt_r2iq (reference)", c);
    // the slowest rate needs 16 transfers (~16-20 ms of mock USB time) before its first slot exists
    verdict(g.n == 50 && g.timeouts == 0 && g.sum / std::max(1, g.n) < 60.0,
            "switching: %d/50 rate switches delivered data, avg %.2f ms to first callback (reference operator avg %.2f ms over %d)",
            g.n, g.n ? g.sum / g.n : 0.0, c.n ? c.sum / c.n : 0.0, c.n);
}

}  // namespace

int main(int argc, char** argv) {
    const int secs = argc > 1 ? atoi(argv[1]) : 15;
    Capture mixed(false), tone(true);
    stress(mixed);
    switching(mixed);
    continuity(tone, secs);
    printf(g_failures ? "FAILED %d\n" : "ALL PASS\n", g_failures);
    return g_failures ? 1 : 0;
}
