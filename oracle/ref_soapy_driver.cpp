// oracle/ref_soapy_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The SoapySDR consumer: the reference's UNMODIFIED SoapySDDC/Settings.cpp + Streaming.cpp +
// Registration.cpp, compiled in place against the declaration-only SoapySDR headers of oracle/shim/SoapySDR, on a mock
// USB layer.  IQ travels RadioHandlerClass callback -> SoapySDDC::Callback -> 16 buffers of 32768 CF32 samples
// (SoapySDDC/Settings.cpp:14-35) -> acquireReadBuffer / readStream (Streaming.cpp:110-182) -> this program.
//     setSampleRate -> setFrequency -> setupStream(RX, CF32) -> activateStream -> readStream ... -> deactivateStream
// oracle/Makefile links it twice, like ref_libsddc_driver.cpp: with the reference's fft_mt_r2iq, and with
// extio_sddc_b200/host/fft_mt_r2iq_cuda.cpp in its place (SoapySDDC passes a null operator, Settings.cpp:47).
// The device drops whole slots and raises SOAPY_SDR_OVERFLOW when all 16 buffers are full (Settings.cpp:17-21); part 1
// keeps the producer within the buffers so both runs are comparable, part 2 lets it run free and expects the overflow
// event to be reported and streaming to continue.
//   usage: ref_soapy_{cpu,cuda} <sample_rate_hz> <n_input_blocks> <out.bin>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <unistd.h>

#include <SoapySDR/Formats.h>
#include <SoapySDR/Registry.hpp>

#include "SoapySDDC.hpp"

namespace {
std::atomic<long long> g_read_elems{0};
std::atomic<bool> g_paced{true};
long long g_elems_per_block = 0;     // complex samples one transfer yields at the running decimation
int g_blocks = 0;

void make_block(int16_t* dst, long first) {
    const double fs = 64000000.0;
    for (int i = 0; i < (int)transferSamples; i++) {
        const long n = first + i;
        const double t = (double)n / fs;
        const double v = 10000.0 * sin(2.0 * M_PI * 9.6e6 * t) + 4000.0 * sin(2.0 * M_PI * 0.33 * fs * t + 0.2) +
                         3000.0 * sin(2.0 * M_PI * 0.02 * fs * t + 1.7) + (double)((n * 4099L) % 811L - 405L);
        dst[i] = (int16_t)lrint(v);
    }
}

class MockUsb : public fx3class {
public:
    ~MockUsb() override { StopStream(); }
    bool Open() override { return true; }
    bool Control(FX3Command, uint8_t) override { return true; }
    bool Control(FX3Command, uint32_t) override { return true; }
    bool Control(FX3Command, uint64_t) override { return true; }
    bool SetArgument(uint16_t, uint16_t) override { return true; }
    bool GetHardwareInfo(uint32_t* data) override {
        const uint8_t d[4] = {(uint8_t)RX888r2, FIRMWARE_VER_MAJOR, FIRMWARE_VER_MINOR, 0};
        memcpy(data, d, 4);
        return true;
    }
    bool ReadDebugTrace(uint8_t*, uint8_t) override { return false; }
    bool Enumerate(unsigned char&, char* name) override {
        strcpy(name, "mock");
        return true;
    }
    void StartStream(ringbuffer<int16_t>& input, int) override {
        input.setBlockSize(transferSamples);
        run_ = true;
        thread_ = std::thread([&input, this] {
            const long long room = 12LL * EXT_BLOCKLEN;     // of the device's 16 buffers
            int b = 0;
            for (; run_ && b < g_blocks; b++) {
                while (run_ && (long long)(b + 1) * g_elems_per_block - g_read_elems.load() > room)
                    std::this_thread::sleep_for(std::chrono::microseconds(50));
                int16_t* p = input.getWritePtr();
                if (!run_) break;
                make_block(p, (long)b * transferSamples);
                input.WriteDone();
            }
            while (run_ && g_paced.load()) std::this_thread::sleep_for(std::chrono::milliseconds(1));
            std::vector<int16_t> again(transferSamples);
            make_block(again.data(), 0);
            while (run_) {                                   // part 2: as fast as the ring takes it, content irrelevant
                int16_t* p = input.getWritePtr();
                if (!run_) break;
                memcpy(p, again.data(), transferSamples * sizeof(int16_t));
                input.WriteDone();
            }
        });
    }
    void StopStream() override {
        run_ = false;
        if (thread_.joinable()) thread_.join();
    }

private:
    std::thread thread_;
    std::atomic<bool> run_{false};
};
}  // namespace

extern "C" fx3class* CreateUsbHandler() { return new MockUsb(); }

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s <sample_rate_hz> <n_input_blocks> <out.bin>\n", argv[0]);
        return 2;
    }
    const double rate = atof(argv[1]);
    g_blocks = atoi(argv[2]);
    int d = 0;
    while ((32000000.0 / (1 << d)) > rate + 1 && d < 4) d++;
    g_elems_per_block = EXT_BLOCKLEN >> d;
    const long long want = (long long)(g_blocks >> d) * EXT_BLOCKLEN;

    // the module's own factory, as the SoapySDR loader would call it (Registration.cpp:28-35)
    if (SoapySDR::Registry::last_name() != "SDDC" || !SoapySDR::Registry::last_make()) return 3;
    SoapySDR::Device* made = SoapySDR::Registry::last_make()(SoapySDR::Kwargs());
    SoapySDDC* dev = dynamic_cast<SoapySDDC*>(made);
    if (!dev) return 4;
    dev->setSampleRate(SOAPY_SDR_RX, 0, rate);
    dev->setFrequency(SOAPY_SDR_RX, 0, 9500000.0);
    SoapySDR::Stream* st = dev->setupStream(SOAPY_SDR_RX, SOAPY_SDR_CF32);
    const size_t mtu = dev->getStreamMTU(st);
    std::vector<float> got;
    got.reserve((size_t)want * 2);
    std::vector<float> buf(2 * 8192);
    dev->activateStream(st);
    dev->setFrequency(SOAPY_SDR_RX, 0, 9500000.0);
    int timeouts = 0, overflows = 0;
    while ((long long)got.size() < want * 2 && timeouts < 50) {
        void* buffs[1] = {buf.data()};
        int flags = 0;
        long long t_ns = 0;
        const int n = dev->readStream(st, buffs, 8192, flags, t_ns, 100000);
        if (n == SOAPY_SDR_TIMEOUT) {
            timeouts++;
            continue;
        }
        if (n == SOAPY_SDR_OVERFLOW) {
            overflows++;
            continue;
        }
        if (n < 0) return 5;
        timeouts = 0;
        got.insert(got.end(), buf.begin(), buf.begin() + 2 * (size_t)n);
        g_read_elems += n;
    }
    const bool part1 = (long long)got.size() >= want * 2 && overflows == 0;

    // part 2: free-running producer, reader asleep -> the 16 buffers fill, slots are dropped, the next acquire reports
    // SOAPY_SDR_OVERFLOW once (Streaming.cpp:141-147) and data flows again afterwards
    g_paced = false;
    for (int ms = 0; ms < 20000 && !dev->_overflowEvent.load(); ms++)      // the float64 FFT of the shim makes the CPU link slow
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
    int seen_overflow = 0, after = 0;
    for (int i = 0; i < 400 && after < 64; i++) {
        void* buffs[1] = {buf.data()};
        int flags = 0;
        long long t_ns = 0;
        const int n = dev->readStream(st, buffs, 8192, flags, t_ns, 100000);
        if (n == SOAPY_SDR_OVERFLOW) seen_overflow++;
        else if (n > 0 && seen_overflow) after++;
    }
    dev->deactivateStream(st);
    dev->closeStream(st);
    const bool part2 = seen_overflow >= 1 && after >= 64;

    FILE* f = fopen(argv[3], "wb");
    if (!f || fwrite(got.data(), sizeof(float), got.size(), f) != got.size()) return 6;
    fclose(f);
    printf("DONE rate=%.0f d=%d blocks=%d mtu=%zu elems=%zu want=%lld overflows_paced=%d overflow_events_free_running=%d reads_after_overflow=%d ssi_marks=%d %s\n",
           rate, d, g_blocks, mtu, got.size() / 2, want, overflows, seen_overflow, after, soapy_shim_ssi_count(),
           part1 && part2 ? "OK" : "BAD");
    fflush(stdout);
    _exit(part1 && part2 ? 0 : 1);   // the device object is left to the OS, as a SoapySDR application exiting would
}
