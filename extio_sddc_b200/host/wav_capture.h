// wav_capture.h -- the capture file format either side of the r2iq path.
//
// The reference records raw ADC streams as RIFF/WAVE files with an rfspace/HDSDR "auxi" chunk
// (libsddc/wavehdr.h:26-102, written by libsddc/wavewrite.c:134-196,245-263; used by
// libsddc/sddc_stream_test.c:185-188: 16-bit, one channel, the ADC rate).  This is a reader/writer of that exact
// layout -- 120-byte header: riff(12) fmt(24) auxi(76) data(8), little-endian, packed -- written from the format
// description, so that recorded RX-888 captures can be replayed through the GPU operator (tools: r2iq_replay)
// and IQ can be written back in the same container (two channels, auxi.centerFreq = tuned frequency).
// Quirks kept bit for bit: fmt.nBlockAlign = nChannels (wavewrite.c:147), riff size = 112 + data bytes,
// auxi size 68, stop time = start time until finalised.
#pragma once
#include <cstdint>
#include <cstdio>

namespace r2iq_host {

struct WavTime {   // Wind_SystemTime, wavehdr.h:34-44
    uint16_t year, month, day_of_week, day, hour, minute, second, milliseconds;
};

struct WavInfo {
    int format_tag = 1;          // 1 = PCM (what the reference writes), 3 = IEEE float (accepted on read, used for cf32 IQ)
    int channels = 1;
    int bits_per_sample = 16;
    uint32_t sample_rate = 0;
    bool has_auxi = false;
    WavTime start = {}, stop = {};
    uint32_t center_freq = 0, ad_sample_rate = 0, if_frequency = 0, bandwidth = 0;
    int32_t iq_offset = 0;
    uint64_t data_offset = 0;    // byte offset of the first sample
    uint64_t data_bytes = 0;     // payload size; taken from the file length when the header was never finalised
    // nBlockAlign as the reference's recorder writes it (= nChannels, wavewrite.c:147; byte parity with its files) or
    // as the WAV format defines it (channels * bits / 8: what standard readers expect; used for the float IQ output)
    bool reference_block_align = true;
};

constexpr int kWavHeaderBytes = 120;

// Serialise the reference's waveFileHeader (wavehdr.h:96-102) for `info` with `data_bytes` of payload.
void wav_pack_header(const WavInfo& info, uint64_t data_bytes, unsigned char out[kWavHeaderBytes]);
// waveWriteHeader (wavewrite.c:171-179): start time = stop time = now (UTC) unless info.start.year != 0.
int wav_write_header(FILE* f, WavInfo* info);
// waveFinalizeHeader (wavewrite.c:245-263): stop time = now unless the caller has set one (stop.year != 0) after
// wav_write_header, sizes fixed, header rewritten.
int wav_finalize(FILE* f, WavInfo* info, uint64_t data_bytes);
// Chunk walker: RIFF/WAVE, "fmt ", optional "auxi", "data"; unknown chunks are skipped; a data size of 0 or one
// running past the end of the file (recording interrupted before waveFinalizeHeader) means "until end of file".
// Returns 0 on success, <0 on malformed input; leaves the stream positioned at the first sample.
int wav_read_header(FILE* f, WavInfo* info);

}  // namespace r2iq_host
