// wav_capture.cpp -- see wav_capture.h.  Host-side file format code; no DSP, no CUDA.
#include "wav_capture.h"

#include <cstring>
#include <ctime>
#include <sys/time.h>

namespace r2iq_host {

namespace {

void put16(unsigned char* p, uint32_t v) {
    p[0] = (unsigned char)(v & 0xFF);
    p[1] = (unsigned char)((v >> 8) & 0xFF);
}
void put32(unsigned char* p, uint32_t v) {
    put16(p, v & 0xFFFF);
    put16(p + 2, v >> 16);
}
uint32_t get16(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
uint32_t get32(const unsigned char* p) { return get16(p) | (get16(p + 2) << 16); }

void put_time(unsigned char* p, const WavTime& t) {
    const uint16_t v[8] = {t.year, t.month, t.day_of_week, t.day, t.hour, t.minute, t.second, t.milliseconds};
    for (int i = 0; i < 8; i++) put16(p + 2 * i, v[i]);
}
WavTime get_time(const unsigned char* p) {
    WavTime t;
    uint16_t* v[8] = {&t.year, &t.month, &t.day_of_week, &t.day, &t.hour, &t.minute, &t.second, &t.milliseconds};
    for (int i = 0; i < 8; i++) *v[i] = (uint16_t)get16(p + 2 * i);
    return t;
}

WavTime now_utc() {   // waveSetCurrTime, wavewrite.c:91-112
    struct timeval tv;
    struct tm t;
    gettimeofday(&tv, nullptr);
    gmtime_r(&tv.tv_sec, &t);
    WavTime w;
    w.year = (uint16_t)(t.tm_year + 1900);
    w.month = (uint16_t)(t.tm_mon + 1);
    w.day_of_week = (uint16_t)t.tm_wday;
    w.day = (uint16_t)t.tm_mday;
    w.hour = (uint16_t)t.tm_hour;
    w.minute = (uint16_t)t.tm_min;
    w.second = (uint16_t)t.tm_sec;
    w.milliseconds = (uint16_t)(tv.tv_usec / 1000);
    return w;
}

}  // namespace

void wav_pack_header(const WavInfo& info, uint64_t data_bytes, unsigned char out[kWavHeaderBytes]) {
    memset(out, 0, kWavHeaderBytes);
    const uint32_t bytes_per_frame = (uint32_t)(info.bits_per_sample / 8 * info.channels);
    memcpy(out, "RIFF", 4);
    put32(out + 4, (uint32_t)(kWavHeaderBytes - 8 + data_bytes));   // sizeof(waveFileHeader) - 8, + data at finalise
    memcpy(out + 8, "WAVE", 4);
    memcpy(out + 12, "fmt ", 4);
    put32(out + 16, 16);
    put16(out + 20, (uint32_t)info.format_tag);
    put16(out + 22, (uint32_t)info.channels);
    put32(out + 24, info.sample_rate);
    put32(out + 28, info.sample_rate * bytes_per_frame);
    put16(out + 32, info.reference_block_align ? (uint32_t)info.channels : bytes_per_frame);   // see WavInfo
    put16(out + 34, (uint32_t)info.bits_per_sample);
    memcpy(out + 36, "auxi", 4);
    put32(out + 40, 68);                                            // 2 * 16 + 9 * 4
    put_time(out + 44, info.start);
    put_time(out + 60, info.stop);
    put32(out + 76, info.center_freq);
    put32(out + 80, info.ad_sample_rate);
    put32(out + 84, info.if_frequency);
    put32(out + 88, info.bandwidth);
    put32(out + 92, (uint32_t)info.iq_offset);
    // Unused2..5 stay zero (96..111)
    memcpy(out + 112, "data", 4);
    put32(out + 116, (uint32_t)data_bytes);
}

int wav_write_header(FILE* f, WavInfo* info) {
    if (!f || !info) return -1;
    if (info->start.year == 0) info->start = now_utc();
    info->stop = info->start;                                       // "to fix", wavewrite.c:153
    if (info->ad_sample_rate == 0) info->ad_sample_rate = info->sample_rate;   // wavewrite.c:155
    info->has_auxi = true;
    unsigned char h[kWavHeaderBytes];
    wav_pack_header(*info, 0, h);
    info->data_offset = kWavHeaderBytes;
    info->data_bytes = 0;
    info->stop = WavTime{};           // open recording: wav_finalize stamps the stop time unless the caller sets one
    return fwrite(h, kWavHeaderBytes, 1, f) == 1 ? 0 : -1;
}

int wav_finalize(FILE* f, WavInfo* info, uint64_t data_bytes) {
    if (!f || !info) return -1;
    if (info->stop.year == 0) info->stop = now_utc();
    info->data_bytes = data_bytes;
    unsigned char h[kWavHeaderBytes];
    wav_pack_header(*info, data_bytes, h);
    if (fseek(f, 0, SEEK_SET)) return -1;
    return fwrite(h, kWavHeaderBytes, 1, f) == 1 ? 0 : -1;
}

int wav_read_header(FILE* f, WavInfo* info) {
    if (!f || !info) return -1;
    *info = WavInfo();
    if (fseek(f, 0, SEEK_END)) return -2;
    const long file_len = ftell(f);
    if (file_len < 12 || fseek(f, 0, SEEK_SET)) return -2;
    unsigned char b[80];
    if (fread(b, 12, 1, f) != 1 || memcmp(b, "RIFF", 4) || memcmp(b + 8, "WAVE", 4)) return -3;
    bool have_fmt = false;
    long pos = 12;
    while (pos + 8 <= file_len) {
        if (fseek(f, pos, SEEK_SET) || fread(b, 8, 1, f) != 1) return -4;
        const uint32_t size = get32(b + 4);
        if (!memcmp(b, "fmt ", 4)) {
            if (size < 16 || fread(b + 8, 16, 1, f) != 1) return -5;
            info->format_tag = (int)get16(b + 8);
            info->channels = (int)get16(b + 10);
            info->sample_rate = get32(b + 12);
            info->bits_per_sample = (int)get16(b + 22);
            have_fmt = true;
        } else if (!memcmp(b, "auxi", 4)) {
            if (size >= 52 && fread(b + 8, 52, 1, f) == 1) {
                info->has_auxi = true;
                info->start = get_time(b + 8);
                info->stop = get_time(b + 24);
                info->center_freq = get32(b + 40);
                info->ad_sample_rate = get32(b + 44);
                info->if_frequency = get32(b + 48);
                info->bandwidth = get32(b + 52);
                info->iq_offset = (int32_t)get32(b + 56);
            }
        } else if (!memcmp(b, "data", 4)) {
            if (!have_fmt) return -6;
            info->data_offset = (uint64_t)pos + 8;
            const uint64_t rest = (uint64_t)file_len - info->data_offset;
            info->data_bytes = (size == 0 || size > rest) ? rest : size;   // header never finalised: until end of file
            if (info->channels < 1 || info->bits_per_sample < 8) return -7;
            return fseek(f, (long)info->data_offset, SEEK_SET) ? -4 : 0;
        }
        pos += 8 + (long)size + (long)(size & 1);                    // chunks are word aligned
    }
    return -8;   // no data chunk
}

}  // namespace r2iq_host
