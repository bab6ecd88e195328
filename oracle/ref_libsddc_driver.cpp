// oracle/ref_libsddc_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The consumer one level above RadioHandlerClass: the reference's UNMODIFIED libsddc/libsddc.cpp
// (compiled in place with -DLIBSDDC_TESTING as unittest/CMakeLists.txt:15-33 does, plus RadioHandler.cpp, config.cpp,
// the radio models and the CPU mixer) is driven through its public C API
//     sddc_open -> sddc_set_rf_mode -> sddc_set_adc_random -> sddc_set_sample_rate -> sddc_set_tuner_frequency ->
//     sddc_set_async_params -> sddc_start_streaming -> sddc_read_sync ... -> sddc_stop_streaming -> sddc_close
// on a mock USB layer (CreateUsbHandler / usb_device_* defined here) that plays a deterministic capture.  IQ reaches the
// caller the way libsddc delivers it: application callback -> 1 MiB byte ring -> sddc_read_sync
// (libsddc/libsddc.cpp:79-126,659-695).  The byte stream read back is written to a file.
//
// oracle/Makefile links this file twice from the same sources:
//   _ref/ref_libsddc_cpu    with the reference's own Core/fft_mt_r2iq*.cpp             (the reference, untouched)
//   _ref/ref_libsddc_cuda   with extio_sddc_b200/host/fft_mt_r2iq_cuda.cpp in their place (the B200 operator behind the
//                           very same `new fft_mt_r2iq()` of RadioHandler.cpp:94-95; nothing else differs)
// and tests/test_host_adapter.py compares the two files.
//
// The sync ring drops the oldest bytes when the reader falls behind (libsddc.cpp:111-121); to keep the two runs
// comparable the mock producer only hands out an input block while the IQ it will yield still fits in the ring beside
// what has not been read yet.
//   usage: ref_libsddc_{cpu,cuda} <sample_rate_hz> <n_input_blocks> <out.bin> [rand=0|1]
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <unistd.h>

#include "FX3Class.h"
#include "config.h"
#include "libsddc.h"
#include "arch/linux/usb_device.h"

extern "C" void sddc_test_reset_sync_buffer(void);   // libsddc.cpp:794 (LIBSDDC_TESTING hook)

namespace {
std::atomic<long long> g_read_bytes{0};
std::atomic<long long> g_async_bytes{0};
std::atomic<int> g_async_calls{0};
long long g_out_bytes_per_block = 0;   // IQ bytes one input block yields at the running decimation
int g_blocks = 0;
bool g_rand = false;

void make_block(int16_t* dst, long first) {
    const double fs = 64000000.0;
    for (int i = 0; i < (int)transferSamples; i++) {
        const long n = first + i;
        const double t = (double)n / fs;
        const double v = 9000.0 * sin(2.0 * M_PI * 7.1e6 * t) + 5000.0 * sin(2.0 * M_PI * 0.29 * fs * t + 0.7) +
                         2500.0 * sin(2.0 * M_PI * 0.051 * fs * t + 2.0) + (double)((n * 6151L) % 997L - 498L);
        int16_t s = (int16_t)lrint(v);
        if (g_rand && (s & 1)) s = (int16_t)(s ^ (int16_t)0xFFFE);
        dst[i] = s;
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
    bool Enumerate(unsigned char&, char*) override { return true; }
    void StartStream(ringbuffer<int16_t>& input, int) override {
        input.setBlockSize(transferSamples);
        run_ = true;
        thread_ = std::thread([&input, this] {
            const long long room = 1024 * 1024 - 256 * 1024;      // one callback's worth (32768 complex) always fits on top
            for (int b = 0; b < g_blocks && run_; b++) {
                while (run_ && (long long)(b + 1) * g_out_bytes_per_block - g_read_bytes.load() > room)
                    std::this_thread::sleep_for(std::chrono::microseconds(50));
                int16_t* p = input.getWritePtr();
                if (!run_) break;
                make_block(p, (long)b * transferSamples);
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

void on_async(uint32_t size, uint8_t*, void*) {   // libsddc.cpp:83-90: the raw callback sees every output slot too
    g_async_bytes += size;
    g_async_calls++;
}
}  // namespace

// the USB layer libsddc.cpp links against (Core/arch/linux/FX3handler.cpp:13, usb_device.c), mocked
extern "C" fx3class* CreateUsbHandler() { return new MockUsb(); }
extern "C" int usb_device_count_devices() { return 1; }
extern "C" int usb_device_get_device_list(struct usb_device_info** infos) {
    *infos = nullptr;
    return 0;
}
extern "C" int usb_device_free_device_list(struct usb_device_info*) { return 0; }

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s <sample_rate_hz> <n_input_blocks> <out.bin> [rand]\n", argv[0]);
        return 2;
    }
    const double rate = atof(argv[1]);
    g_blocks = atoi(argv[2]);
    g_rand = argc > 4 && atoi(argv[4]) != 0;
    int d = 0;
    while ((32000000.0 / (1 << d)) > rate + 1 && d < 4) d++;     // libsddc.cpp:20-27: 32 MS/s >> d at a 64 Msps ADC
    g_out_bytes_per_block = (long long)(EXT_BLOCKLEN >> d) * 2 * (long long)sizeof(float);
    const long long want = (long long)(g_blocks >> d) * EXT_BLOCKLEN * 2 * (long long)sizeof(float);

    char fw[] = "/tmp/ref_libsddc_fw_XXXXXX";
    const int fd = mkstemp(fw);
    if (fd < 0 || write(fd, "fw", 2) != 2) return 3;
    close(fd);
    sddc_t* dev = sddc_open(0, fw);
    remove(fw);
    if (!dev || sddc_get_status(dev) != SDDC_STATUS_READY) {
        fprintf(stderr, "sddc_open failed: %s\n", sddc_get_last_error() ? sddc_get_last_error() : "?");
        return 4;
    }
    sddc_set_rf_mode(dev, HF_MODE);
    sddc_set_adc_random(dev, g_rand ? 1 : 0);
    if (sddc_set_sample_rate(dev, rate) != 0) return 5;
    sddc_set_tuner_frequency(dev, 7000000.0);
    sddc_set_async_params(dev, 0, 0, on_async, nullptr);
    sddc_test_reset_sync_buffer();
    std::vector<uint8_t> got;
    got.reserve((size_t)want);
    const auto t0 = std::chrono::steady_clock::now();
    sddc_start_streaming(dev);
    sddc_set_tuner_frequency(dev, 7000000.0);   // residual recomputed with the running decimation, like the applications
    std::vector<uint8_t> buf(65536);
    int idle = 0;
    while ((long long)got.size() < want && idle < 10) {
        int n = 0;
        sddc_read_sync(dev, buf.data(), (int)buf.size(), &n);
        if (n == 0) {
            idle++;      // sddc_read_sync waited a second for nothing
            continue;
        }
        idle = 0;
        got.insert(got.end(), buf.begin(), buf.begin() + n);
        g_read_bytes += n;
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    sddc_stop_streaming(dev);
    sddc_close(dev);
    FILE* f = fopen(argv[3], "wb");
    if (!f || fwrite(got.data(), 1, got.size(), f) != got.size()) return 6;
    fclose(f);
    printf("DONE rate=%.0f d=%d blocks=%d bytes=%zu want=%lld async_calls=%d async_bytes=%lld seconds=%.3f\n", rate, d, g_blocks,
           got.size(), want, g_async_calls.load(), g_async_bytes.load(), sec);
    return (long long)got.size() >= want ? 0 : 1;
}
