// CPU test of the capture-file reader/writer against files written by the REFERENCE's libsddc/wavewrite.c
// (tests/golden/ref_capture*.wav, made by tests/golden/make_golden_wav.py):
//   * the reader recovers every field the reference wrote and finds the payload;
//   * the header re-packed from the parsed fields is byte-identical to the reference's 120 bytes;
//   * write -> finalize -> read round trip, and an un-finalised recording (sizes still 0) reads to end of file.
// argv[1] = directory with the golden files, argv[2] = scratch directory
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "wav_capture.h"

using namespace r2iq_host;
static int failures = 0;
#define CHECK(c) do { if (!(c)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); failures++; } } while (0)

static void golden(const std::string& path, int channels, uint32_t rate, uint32_t freq, uint64_t payload) {
    FILE* f = fopen(path.c_str(), "rb");
    CHECK(f != nullptr);
    if (!f) return;
    WavInfo w;
    CHECK(wav_read_header(f, &w) == 0);
    CHECK(w.format_tag == 1 && w.channels == channels && w.bits_per_sample == 16 && w.sample_rate == rate);
    CHECK(w.has_auxi && w.center_freq == freq && w.ad_sample_rate == rate && w.if_frequency == 0 && w.bandwidth == 0);
    CHECK(w.data_offset == (uint64_t)kWavHeaderBytes && w.data_bytes == payload);
    CHECK(w.start.year >= 2024 && w.start.month >= 1 && w.start.month <= 12 && w.stop.year >= w.start.year);
    CHECK(ftell(f) == kWavHeaderBytes);
    unsigned char ref[kWavHeaderBytes], mine[kWavHeaderBytes];
    fseek(f, 0, SEEK_SET);
    CHECK(fread(ref, kWavHeaderBytes, 1, f) == 1);
    wav_pack_header(w, w.data_bytes, mine);
    CHECK(memcmp(ref, mine, kWavHeaderBytes) == 0);
    fclose(f);
    printf("%s: header byte-identical to libsddc/wavewrite.c, %llu payload bytes\n", path.c_str(), (unsigned long long)w.data_bytes);
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const std::string gold = argv[1], tmp = argv[2];
    golden(gold + "/ref_capture.wav", 1, 128000000u, 0u, 512);
    golden(gold + "/ref_capture_iq.wav", 2, 2000000u, 7100000u, 1024);

    std::vector<int16_t> x(1000);
    for (size_t i = 0; i < x.size(); i++) x[i] = (int16_t)(i * 37 - 9000);
    const std::string p = tmp + "/wav_roundtrip.wav";
    {   // write -> finalize -> read
        FILE* f = fopen(p.c_str(), "wb+");
        WavInfo w;
        w.sample_rate = 64000000u;
        w.center_freq = 5000000u;
        CHECK(wav_write_header(f, &w) == 0);
        CHECK(fwrite(x.data(), 2, x.size(), f) == x.size());
        CHECK(wav_finalize(f, &w, 2 * x.size()) == 0);
        fclose(f);
        f = fopen(p.c_str(), "rb");
        WavInfo r;
        CHECK(wav_read_header(f, &r) == 0);
        CHECK(r.sample_rate == 64000000u && r.center_freq == 5000000u && r.ad_sample_rate == 64000000u && r.channels == 1);
        CHECK(r.data_bytes == 2 * x.size());
        std::vector<int16_t> y(x.size());
        CHECK(fread(y.data(), 2, y.size(), f) == y.size());
        CHECK(x == y);
        fclose(f);
    }
    {   // recording interrupted before waveFinalizeHeader: sizes still zero -> payload runs to end of file
        FILE* f = fopen(p.c_str(), "wb+");
        WavInfo w;
        w.sample_rate = 128000000u;
        CHECK(wav_write_header(f, &w) == 0);
        CHECK(fwrite(x.data(), 2, 600, f) == 600);
        fclose(f);
        f = fopen(p.c_str(), "rb");
        WavInfo r;
        CHECK(wav_read_header(f, &r) == 0);
        CHECK(r.data_bytes == 1200 && r.data_offset == (uint64_t)kWavHeaderBytes);
        fclose(f);
    }
    {   // not a WAVE file / truncated
        FILE* f = fopen(p.c_str(), "wb+");
        fwrite("RIFFxxxxAVI ", 12, 1, f);
        fclose(f);
        f = fopen(p.c_str(), "rb");
        WavInfo r;
        CHECK(wav_read_header(f, &r) < 0);
        fclose(f);
    }
    {   // float IQ output of the replayer: a standard WAV block alignment, and a restored time span survives finalise
        FILE* f = fopen(p.c_str(), "wb+");
        WavInfo w;
        w.format_tag = 3;
        w.channels = 2;
        w.bits_per_sample = 32;
        w.sample_rate = 2000000;
        w.reference_block_align = false;
        w.start.year = 2025; w.start.month = 3; w.start.day = 9; w.start.hour = 11;
        CHECK(wav_write_header(f, &w) == 0);
        CHECK(w.stop.year == 0);                       // open recording
        WavTime stop = w.start;
        stop.hour = 12;
        w.stop = stop;                                 // the capture's own stop time
        std::vector<float> iq(200, 0.25f);
        CHECK(fwrite(iq.data(), 4, iq.size(), f) == iq.size());
        CHECK(wav_finalize(f, &w, 800) == 0);
        unsigned char h[kWavHeaderBytes];
        fseek(f, 0, SEEK_SET);
        CHECK(fread(h, kWavHeaderBytes, 1, f) == 1);
        CHECK(h[32] == 8 && h[33] == 0);               // nBlockAlign = channels * bits / 8
        fseek(f, 0, SEEK_SET);
        WavInfo r;
        CHECK(wav_read_header(f, &r) == 0);
        CHECK(r.stop.year == 2025 && r.stop.hour == 12 && r.start.hour == 11 && r.format_tag == 3 && r.data_bytes == 800);
        fclose(f);
        // and an open recording that nobody stamps gets "now"
        f = fopen(p.c_str(), "wb+");
        WavInfo v;
        CHECK(wav_write_header(f, &v) == 0 && wav_finalize(f, &v, 0) == 0 && v.stop.year >= 2024);
        fclose(f);
    }
    remove(p.c_str());
    printf(failures ? "FAILED %d\n" : "ALL PASS\n", failures);
    return failures ? 1 : 0;
}
