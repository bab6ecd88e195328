// r2iq_replay -- replay a recorded ADC capture through the B200 operator.
//
//   r2iq_replay IN.wav OUT.wav --decimate D [--tune-hz F] [--rand] [--sideband] [--gain G] [--fftn N]
//                              [--batch-blocks B] [--device I]
//
// IN.wav is what libsddc/sddc_stream_test.c:185-188 records: RIFF/WAVE with an "auxi" chunk, 16-bit, ONE channel,
// real ADC samples at the ADC rate (the header's nSamplesPerSec).  The capture is cut into 65536-sample transfers
// (Core/config.h:80-81) exactly as the USB path delivers them; a trailing partial transfer is dropped (the
// reference never sees one).  Each batch goes through r2iq_cuda_process with history carried on the handle, so the
// result equals one pass over the whole capture.  --tune-hz F is what RadioHandlerClass::TuneLO hands to
// setFreqOffset in HF mode (Core/RadioHandler.cpp:280-300): offset = F / (Fs/2); the residual below one tune-bin
// step is applied by the device fine-tune mixer (r2iq_cuda_set_fine_tune), as OnDataPacket does on the CPU.
// OUT.wav: same container, two channels of IEEE float32 (format tag 3) at Fs / 2^(D+1), auxi.centerFreq = F,
// auxi.ADsamplerate = Fs.  Pure host-side plumbing; no DSP here, and no CPU path if the GPU is missing.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "r2iq_cuda.h"
#include "wav_capture.h"

using namespace r2iq_host;

static int usage() {
    fprintf(stderr, "usage: r2iq_replay IN.wav OUT.wav --decimate D [--tune-hz F] [--rand] [--sideband] [--gain G] "
                    "[--fftn 2048|8192] [--batch-blocks B] [--device I]\n");
    return 2;
}

int main(int argc, char** argv) {
    if (argc < 3) return usage();
    const char* in_path = argv[1];
    const char* out_path = argv[2];
    int d = 0, fftn = 2048, device = -1, rnd = 0, lsb = 0;
    long batch = 256;
    double tune_hz = -1.0;
    float gain = 1.08e-8f;   // RX888mk2_GAINFACTOR, Core/config.h:52-55
    for (int i = 3; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&](double& v) { if (i + 1 >= argc) return false; v = atof(argv[++i]); return true; };
        double v = 0;
        if (a == "--decimate" && val(v)) d = (int)v;
        else if (a == "--tune-hz" && val(v)) tune_hz = v;
        else if (a == "--gain" && val(v)) gain = (float)v;
        else if (a == "--fftn" && val(v)) fftn = (int)v;
        else if (a == "--batch-blocks" && val(v)) batch = (long)v;
        else if (a == "--device" && val(v)) device = (int)v;
        else if (a == "--rand") rnd = 1;
        else if (a == "--sideband") lsb = 1;
        else return usage();
    }
    if (d < 0 || d >= R2IQ_NDECIDX || batch < 1) return usage();
    batch = ((batch + (1L << d) - 1) >> d) << d;          // whole output slots per batch

    FILE* fin = fopen(in_path, "rb");
    if (!fin) { fprintf(stderr, "r2iq_replay: cannot open %s\n", in_path); return 1; }
    WavInfo wi;
    int rc = wav_read_header(fin, &wi);
    if (rc != 0 || wi.channels != 1 || wi.bits_per_sample != 16 || wi.format_tag != 1) {
        fprintf(stderr, "r2iq_replay: %s is not a 16-bit one-channel PCM capture (rc=%d ch=%d bits=%d tag=%d)\n", in_path, rc,
                wi.channels, wi.bits_per_sample, wi.format_tag);
        return 1;
    }
    const double fs = wi.has_auxi && wi.ad_sample_rate ? (double)wi.ad_sample_rate : (double)wi.sample_rate;
    const long total_blocks = (long)(wi.data_bytes / 2 / R2IQ_TRANSFER_SAMPLES);

    r2iq_cuda_config cfg = {};
    cfg.device = device;
    cfg.gain = gain;
    cfg.fftn = fftn;
    r2iq_cuda_t h = nullptr;
    if (r2iq_cuda_create(&cfg, &h) != R2IQ_OK) {
        fprintf(stderr, "r2iq_replay: %s\n", r2iq_cuda_last_error());   // no GPU, no result
        return 1;
    }
    r2iq_cuda_set_decimate(h, d);
    r2iq_cuda_set_rand(h, rnd);
    r2iq_cuda_set_sideband(h, lsb);
    if (tune_hz >= 0.0) {
        const float residual = r2iq_cuda_set_freq_offset(h, (float)(tune_hz / (fs / 2.0)));
        r2iq_cuda_set_fine_tune(h, residual);
    }
    r2iq_cuda_turn_on(h);
    const long out_per_block = r2iq_cuda_out_per_block(h);

    int16_t* hin = (int16_t*)r2iq_cuda_host_alloc_on(sizeof(int16_t) * (size_t)batch * R2IQ_TRANSFER_SAMPLES, device);
    float* hout = (float*)r2iq_cuda_host_alloc_on(sizeof(float) * 2 * (size_t)batch * (size_t)out_per_block, device);
    FILE* fout = fopen(out_path, "wb+");
    if (!hin || !hout || !fout) { fprintf(stderr, "r2iq_replay: %s\n", fout ? r2iq_cuda_last_error() : "cannot create output"); return 1; }
    WavInfo wo;
    wo.format_tag = 3;
    wo.channels = 2;
    wo.bits_per_sample = 32;
    wo.reference_block_align = false;        // a standard float WAV: nBlockAlign = 8
    wo.sample_rate = (uint32_t)(fs / 2.0 / (double)(1 << d) + 0.5);
    wo.center_freq = tune_hz >= 0.0 ? (uint32_t)(tune_hz + 0.5) : 0u;
    wo.ad_sample_rate = (uint32_t)(fs + 0.5);
    wo.start = wi.start;
    wo.stop = wi.stop;
    wav_write_header(fout, &wo);
    wo.stop = wi.stop;   // keep the recording's own time span (a stop time of year 0 means: stamp at finalise)

    uint64_t out_bytes = 0;
    long done = 0;
    // blocks 2^d-aligned per batch; a tail shorter than one output slot still yields its share of samples
    while (done < total_blocks) {
        const long nb = total_blocks - done < batch ? total_blocks - done : batch;
        if (fread(hin, sizeof(int16_t) * R2IQ_TRANSFER_SAMPLES, (size_t)nb, fin) != (size_t)nb) { fprintf(stderr, "r2iq_replay: short read\n"); return 1; }
        if (r2iq_cuda_process(h, hin, nb, nullptr, hout) != R2IQ_OK) { fprintf(stderr, "r2iq_replay: %s\n", r2iq_cuda_last_error()); return 1; }
        const size_t nf = (size_t)nb * (size_t)out_per_block;
        if (fwrite(hout, 2 * sizeof(float), nf, fout) != nf) { fprintf(stderr, "r2iq_replay: short write\n"); return 1; }
        out_bytes += 2 * sizeof(float) * nf;
        done += nb;
    }
    wav_finalize(fout, &wo, out_bytes);
    fclose(fout);
    fclose(fin);
    r2iq_cuda_stats st;
    r2iq_cuda_get_stats(h, &st);
    printf("r2iq_replay: %ld blocks (%.3f s at %.0f Msps) -> %llu IQ samples at %.3f MS/s, tune bin %d, %llu launches\n", total_blocks,
           (double)total_blocks * R2IQ_TRANSFER_SAMPLES / fs, fs / 1e6, (unsigned long long)(out_bytes / 8), wo.sample_rate / 1e6,
           r2iq_cuda_get_tunebin(h), (unsigned long long)st.kernel_launches);
    r2iq_cuda_turn_off(h);
    r2iq_cuda_host_free(hin);
    r2iq_cuda_host_free(hout);
    r2iq_cuda_destroy(h);
    return 0;
}
