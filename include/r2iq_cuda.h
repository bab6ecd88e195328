/* r2iq_cuda.h -- C ABI of the B200 (sm_100a) real->IQ down-converter backend.
 *
 * This is the drop-in boundary: plain C, opaque handle, pointers and sizes only.  It carries
 * exactly the operations the reference's `r2iqControlClass` / `fft_mt_r2iq` operator exposes to
 * its single caller RadioHandlerClass, so a `cuda_r2iq : r2iqControlClass` adapter (see
 * extio_sddc_b200/host/cuda_r2iq.h and INTEGRATION.md) can sit beside the FFTW/AVX/NEON workers.
 * Each entry point cites the reference interface it replaces (paths relative to the reference
 * tree, ShadowLive/ExtIO_sddc).
 *
 * Conventions: every function returning int yields 0 on success and a negative R2IQ_E_* code on
 * failure; r2iq_cuda_last_error() then returns a thread-local description (same style as
 * libsddc/libsddc.cpp:31-46).  Nothing throws across this boundary.  There is NO CPU fallback:
 * without a usable CUDA device r2iq_cuda_create fails.
 *
 * Stream model: the ADC stream is a sequence of blocks of R2IQ_TRANSFER_SAMPLES int16 samples
 * (Core/config.h:80-81).  Every block yields exactly R2IQ_TRANSFER_SAMPLES/2 >> d complex float
 * samples (interleaved re,im), d = decimation index 0..6; 2^d input blocks fill one 32768-complex
 * output slot of the reference's output ring (Core/fft_mt_r2iq_impl.hpp:117-148).  Block b needs
 * the last halfFft samples of block b-1 ("history", impl.hpp:32,42).
 */
#ifndef R2IQ_CUDA_H
#define R2IQ_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define R2IQ_API __attribute__((visibility("default")))
#else
#define R2IQ_API
#endif

#define R2IQ_TRANSFER_SAMPLES 65536   /* Core/config.h:80-81  transferSize / sizeof(int16_t) */
#define R2IQ_NDECIDX 7                /* Core/r2iq.h:5 */
#define R2IQ_FFTN 2048                /* FFTN_R_ADC of the north-star configuration (Core/config.h:49 ships 8192) */
#define R2IQ_HALF_FFT (R2IQ_FFTN / 2) /* Core/fft_mt_r2iq.h:18 */

enum {
    R2IQ_OK = 0,
    R2IQ_E_INVALID = -1,    /* bad argument */
    R2IQ_E_CUDA = -2,       /* CUDA runtime error (sticky: the handle refuses further work) */
    R2IQ_E_NOMEM = -3,
    R2IQ_E_STATE = -4,      /* call order violated (e.g. process before filters exist) */
    R2IQ_E_NODEVICE = -5    /* no CUDA device / wrong architecture: there is no CPU path */
};

typedef struct r2iq_cuda_s* r2iq_cuda_t;

typedef struct r2iq_cuda_config {
    int device;            /* CUDA device ordinal; -1 = current device */
    float gain;            /* RadioHardware::getGain(), the `gain` argument of fft_mt_r2iq::Init */
    int chunk_blocks;      /* host-buffer path: blocks per pipelined H2D/compute/D2H chunk; 0 = default (256) */
    int fftn;              /* FFTN_R_ADC (Core/config.h:49): 0 or 2048 = north-star size, 8192 = the value the reference
                              tree ships with.  halfFft = fftn/2 sets history length, filter length (halfFft complex bins,
                              halfFft/4+1 taps) and the tune-bin range; the channelizer exists for 2048 only. */
} r2iq_cuda_config;

typedef struct r2iq_cuda_stats {
    uint64_t kernel_launches;    /* fused kernels launched since create */
    uint64_t blocks_processed;   /* input blocks */
    uint64_t h2d_bytes;          /* host->device bytes moved by r2iq_cuda_process */
    uint64_t d2h_bytes;          /* device->host bytes moved by r2iq_cuda_process */
    int sm_count;
    int cc_major, cc_minor;
} r2iq_cuda_stats;

/* Thread-local description of the last failure on this thread. */
R2IQ_API const char* r2iq_cuda_last_error(void);
/* Library / kernel build identification, e.g. "r2iq-b200 0.1 sm_100a fftn=2048,8192". */
R2IQ_API const char* r2iq_cuda_version(void);

/* fft_mt_r2iq::fft_mt_r2iq() + fft_mt_r2iq::Init(gain, in, out)  (Core/fft_mt_r2iq.cpp:39-49,147-227).
 * Designs the seven decimation filters on the host exactly as Init does (KaiserWindow, Core/fir.cpp:48-105;
 * tap placement and gain, fft_mt_r2iq.cpp:185-205), uploads their spectra, and sets the constructor
 * defaults: decimation 0, tune bin halfFft/4, rand off, sideband off. */
R2IQ_API int r2iq_cuda_create(const r2iq_cuda_config* cfg, r2iq_cuda_t* out);
/* fft_mt_r2iq::~fft_mt_r2iq()  (Core/fft_mt_r2iq.cpp:71-98). */
R2IQ_API int r2iq_cuda_destroy(r2iq_cuda_t h);

/* filterHw[d] (Core/fft_mt_r2iq.h:101): replace / read back one filter spectrum, halfFft complex
 * floats interleaved.  A host that designs its own filter (Core/fir.cpp) uploads it once per
 * Init/setDecimate as the north star describes. */
R2IQ_API int r2iq_cuda_set_filter(r2iq_cuda_t h, int d, const float* hw_interleaved, int n_complex);
R2IQ_API int r2iq_cuda_get_filter(r2iq_cuda_t h, int d, float* hw_interleaved, int n_complex);

/* r2iqControlClass::setDecimate / getRatio  (Core/r2iq.h:22,30).  Like the reference, a new
 * decimation takes effect at the next TurnOn (the worker latches it once, impl.hpp:3-6). */
R2IQ_API int r2iq_cuda_set_decimate(r2iq_cuda_t h, int d);
R2IQ_API int r2iq_cuda_get_decimate(r2iq_cuda_t h);
R2IQ_API int r2iq_cuda_get_ratio(r2iq_cuda_t h);

/* fft_mt_r2iq::setFreqOffset  (Core/fft_mt_r2iq.cpp:101-109): offset = f / (Fs/2) in [0,1);
 * quantises the tune bin to a multiple of 4 (truncation toward zero) and returns the residual
 * (mtunebin/halfFft - offset) * ratio for the fine-tune mixer.  Picked up by the next block. */
R2IQ_API float r2iq_cuda_set_freq_offset(r2iq_cuda_t h, float offset);
/* Direct access to mtunebin (Core/fft_mt_r2iq.h:90); must be a multiple of 4 in [0, halfFft). */
R2IQ_API int r2iq_cuda_set_tunebin(r2iq_cuda_t h, int tunebin);
R2IQ_API int r2iq_cuda_get_tunebin(r2iq_cuda_t h);

/* r2iqControlClass::updateRand/getRand, setSideband/getSideband  (Core/r2iq.h:24-28). */
R2IQ_API int r2iq_cuda_set_rand(r2iq_cuda_t h, int on);
R2IQ_API int r2iq_cuda_get_rand(r2iq_cuda_t h);
R2IQ_API int r2iq_cuda_set_sideband(r2iq_cuda_t h, int lsb);
R2IQ_API int r2iq_cuda_get_sideband(r2iq_cuda_t h);

/* fft_mt_r2iq::TurnOn / TurnOff / IsOn  (Core/fft_mt_r2iq.cpp:111-145).  TurnOn latches decimation
 * and sideband and clears the carried history (a fresh ring starts from zeros here; the reference
 * reads uninitialised ring memory for the first block, impl.hpp:32).  TurnOff drains outstanding work. */
R2IQ_API int r2iq_cuda_turn_on(r2iq_cuda_t h);
R2IQ_API int r2iq_cuda_turn_off(r2iq_cuda_t h);
R2IQ_API int r2iq_cuda_is_on(r2iq_cuda_t h);

/* Complex output samples per input block at the latched decimation: 32768 >> d. */
R2IQ_API long r2iq_cuda_out_per_block(r2iq_cuda_t h);

/* The worker loop body (Core/fft_mt_r2iq_impl.hpp:15-151) for n_blocks consecutive blocks held in
 * HOST memory.  blocks: n_blocks*65536 int16.  history: the halfFft samples preceding blocks[0], or
 * NULL to continue the stream (history carried from the previous call; zeros after TurnOn).
 * iq_out: n_blocks*(32768>>d) complex floats, interleaved.  Copies host->device, runs the fused
 * kernel and copies device->host in pipelined chunks; returns when iq_out is complete.
 * Pinned host memory (r2iq_cuda_host_alloc) is needed for the copies to overlap. */
R2IQ_API int r2iq_cuda_process(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history, float* iq_out);

/* The same, without waiting: the worker loop of the reference overlaps its producer, its FFTs and its consumer through
 * the two rings (Core/fft_mt_r2iq_impl.hpp:15-33,100-148); over PCIe that overlap needs submissions in flight.
 *  process_async  enqueues host->device copy, fused kernel and (when iq_out is not NULL) the copy back for
 *                 n_blocks <= chunk_blocks blocks and returns a ticket; history is carried exactly as by
 *                 r2iq_cuda_process (the tail of `blocks` is saved before the call returns).  Up to 4 submissions
 *                 are in flight; a further one first waits for the oldest.  With iq_out == NULL the result stays on
 *                 the device until r2iq_cuda_release: fetch it in pieces with r2iq_cuda_read (the adapter copies
 *                 straight into output-ring slots as they become free).
 *  wait_input     returns when the submission's blocks have left the caller's buffer (the input-ring slots may be
 *                 handed back to the producer: ReadDone, Core/dsp/ringbuffer.h:29-42).
 *  wait           returns when everything enqueued for the ticket has finished (iq_out complete).
 *  read           copies complex samples [first, first + n) of a held result to dst and waits for the copy.
 *  release        gives a held submission's device buffers back.
 * Tickets complete in submission order.  Host buffers should be page-locked (r2iq_cuda_host_alloc, or memory the
 * caller owns made so with r2iq_cuda_host_register: the reference's rings are one `new T[]` each,
 * Core/dsp/ringbuffer.h:154-170) for the copies to be asynchronous. */
R2IQ_API int r2iq_cuda_process_async(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history,
                                     float* iq_out, uint64_t* ticket);
R2IQ_API int r2iq_cuda_wait_input(r2iq_cuda_t h, uint64_t ticket);
R2IQ_API int r2iq_cuda_wait(r2iq_cuda_t h, uint64_t ticket);
R2IQ_API int r2iq_cuda_read(r2iq_cuda_t h, uint64_t ticket, long first_complex, long n_complex, float* dst);
R2IQ_API int r2iq_cuda_release(r2iq_cuda_t h, uint64_t ticket);
/* Page-lock / unlock memory the caller allocated (cudaHostRegister); registering a range twice is not an error. */
R2IQ_API int r2iq_cuda_host_register(void* p, size_t bytes);
R2IQ_API int r2iq_cuda_host_unregister(void* p);

/* One ADC stream at several output rates.  In the reference every rate is its own fft_mt_r2iq instance reading
 * the SAME input ring slots in place (Core/fft_mt_r2iq_impl.hpp:20-33; RadioHandlerClass owns the ring,
 * Core/RadioHandler.h:111); here the capture crosses PCIe once and is transformed n_rates times on the device:
 * decimations[r] in [0, 7) selects the rate of iq_outs[r] (n_blocks*(32768>>d) complex floats each), tune bin,
 * rand and the latched sideband as for r2iq_cuda_process, history handling identical (carried across calls).
 * Every output equals what r2iq_cuda_process gives at that decimation.  The latched decimation is left unchanged;
 * the fine-tune mixer (one residual per rate) must be off. */
R2IQ_API int r2iq_cuda_process_rates(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history,
                                     const int* decimations, int n_rates, float* const* iq_outs);

/* Same transform on DEVICE memory, asynchronously on the caller's stream (cudaStream_t passed as
 * void*; NULL = the legacy default stream).  d_in points at [halfFft history samples | n_blocks
 * blocks] contiguous in device memory (16-byte aligned: frames are fetched by TMA bulk copies); d_iq_out receives n_blocks*(32768>>d)
 * complex floats.  No host synchronisation; stream state is not carried.
 * ONE HANDLE, ONE STREAM: a handle's device-path launches share per-handle scratch (the work-counter ring, and in
 * R2IQ_NCO_REFERENCE mode the start-phasor buffer, whose upload also synchronises the stream), so all device-path calls
 * on one handle must be issued on the same stream, or be ordered by the caller.  Use one handle per stream for
 * concurrent work; handles share nothing.  (R2IQ_NCO_DIRECT and mixer-off launches touch no host state at all.)
 * One launch takes at most 49,941,000 blocks (the kernels index frames with 32 bits; 6.5 TB of samples): more is
 * R2IQ_E_INVALID. */
R2IQ_API int r2iq_cuda_process_device(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_iq_out, void* stream);

/* r2iq_cuda_process_device for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Arguments, result and errors as r2iq_cuda_process_device (same kernel, same tables, same bits), but everything the
 * launch needs beside the caller's buffers -- the work counter of the persistent kernel and the per-(filter, d, tune
 * bin) tables, which r2iq_cuda_process_device builds into cudaMalloc'ed buffers on first use -- lives in ONE
 * file-scope `static __device__` array of R2IQ_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the
 * library holds it until the process exits).  The call issues no cudaMalloc / cudaFree and no allocating library
 * routine: a memset of the counter, the kernel, and -- only when the (filter, d, tune bin) differ from the last call
 * on that device -- one cudaMemcpyAsync of the freshly built tables on `stream`.
 * Limits, each R2IQ_E_INVALID / R2IQ_E_STATE with a message naming it:
 *   - FFTN 2048 operators only: a table set is at most 41,216 bytes; the FFTN 8192 tables (~150 KB) do not fit.
 *   - at most R2IQ_STATIC_MAX_BLOCKS blocks per call (1 GiB of samples); larger launches go through
 *     r2iq_cuda_process_device.
 *   - fine-tune mixer off or R2IQ_NCO_DIRECT; the R2IQ_NCO_REFERENCE form needs start phasors that grow with the launch.
 * ALL calls in the process (all handles, all streams) on one device share the array, and it keeps the last call's
 * contents: the counter is cleared by every call, the tables are replaced when they differ.  Calls that overlap on the
 * GPU therefore RACE (two kernels on one counter; tables replaced under a running kernel): issue them on one stream,
 * or order them yourself. */
#define R2IQ_STATIC_SCRATCH_BYTES 49152
#define R2IQ_STATIC_MAX_BLOCKS 8192
R2IQ_API int r2iq_cuda_process_device_static(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, float* d_iq_out, void* stream);

/* Wideband channelizer: n_channels tuned channels from ONE input stream.  Channel c is tuned to tunebins[c]
 * (multiples of 4 in [0, halfFft), the value setFreqOffset would store in mtunebin) and receives exactly the stream a
 * separate fft_mt_r2iq tuned there would produce (Core/fft_mt_r2iq_impl.hpp:76-138) at the latched decimation, with
 * the current rand / sideband settings.  iq_out is channel-major: channel c occupies
 * [c*n_blocks*(32768>>d), (c+1)*n_blocks*(32768>>d)) complex floats.
 *   FFTN 2048, d >= 2: a single pass; the forward FFT and the real-FFT split of every frame pair are computed once and
 *                      shared by all channels (d = 5, 6: one channel per lane; d = 2..4: channel after channel, the
 *                      backward transform across the lanes).
 *   FFTN 8192, d = 4, 5: a single pass as well (teams of four warps; r2iq_wide_channels.cu).
 *   d = 0, 1 and the other FFTN 8192 decimations: one launch of the single-channel kernel per channel (the reference's
 *                      one operator per tuned instance; same result, forward FFT not shared).
 * Host-buffer form: history as r2iq_cuda_process (NULL = zeros; no state is carried).
 * Device form: d_in = [halfFft history | blocks], d_in and d_iq_out 16-byte aligned, asynchronous on `stream`. */
R2IQ_API int r2iq_cuda_process_channels(r2iq_cuda_t h, const int16_t* blocks, long n_blocks, const int16_t* history,
                                        const int* tunebins, int n_channels, float* iq_out);
R2IQ_API int r2iq_cuda_process_channels_device(r2iq_cuda_t h, const int16_t* d_in, long n_blocks, const int* tunebins,
                                               int n_channels, float* d_iq_out, void* stream);

/* Residual fine-tune mixer on the device.  RadioHandlerClass::TuneLO keeps the value
 * returned by setFreqOffset as `fc`, restarts the NCO with shift_limited_unroll_C_sse_init(fc, 0) when it
 * changes, and OnDataPacket rotates every output slot with shift_limited_unroll_C_sse_inp_c when fc != 0
 * (Core/RadioHandler.cpp:33-37,287-297; Core/pffft/pf_mixer.cpp:750-856).  This call is that init: from the
 * next processed block on, every output sample of r2iq_cuda_process / r2iq_cuda_process_device is multiplied
 * by the same phasor sequence, bit-identical to the CPU mixer (phase continues across calls; fc = 0 switches
 * it off; the same fc again leaves the running phase alone).  Returns fc, or 0 when switched off / on error.
 * Not applied by the channelizer entry points. */
R2IQ_API float r2iq_cuda_set_fine_tune(r2iq_cuda_t h, float fc);
/* How the mixer's phasors are formed.
 *  R2IQ_NCO_REFERENCE (default): the reference's float32 recurrence (one renormalised step per 128 samples, run on
 *      the host because every step depends on the one before) -- bit-identical to Core/pffft/pf_mixer.cpp:791-856,
 *      and bounded by that sequential recurrence (about 30 ms per 2^21 steps) and a second pass over the IQ.
 *  R2IQ_NCO_DIRECT: sample n since the last r2iq_cuda_set_fine_tune(fc != 0) is multiplied by exp(2 pi j fc n), the
 *      phase reduced in float64 on the device, in the store epilogue of the fused kernel: no host work, no second
 *      pass.  This is the mixer the reference's recurrence approximates; it is NOT bit-identical to it.  Stated
 *      tolerance: relative L2 error <= 1e-6 and max-abs <= 1e-5 of full scale against the float64 ideal
 *      exp(2 pi j fc n) applied to the unmixed float32 stream, at any n; against the reference's own float32
 *      recurrence the difference is the reference's accumulated rounding (a phase random walk of about
 *      6e-8 rad * sqrt(n / 128)), see tests/test_gpu_parity.py::test_fine_tune_mixer_direct_mode.
 * The mode can be changed only while the mixer is off (the two forms keep separate phase state). */
enum { R2IQ_NCO_REFERENCE = 0, R2IQ_NCO_DIRECT = 1 };
R2IQ_API int r2iq_cuda_set_fine_tune_mode(r2iq_cuda_t h, int mode);
R2IQ_API int r2iq_cuda_get_fine_tune_mode(r2iq_cuda_t h);

/* Block until everything issued through this handle has finished. */
R2IQ_API int r2iq_cuda_sync(r2iq_cuda_t h);

/* Page-locked host buffers for ring slots / captures (replaces `new T[]` in Core/dsp/ringbuffer.h:166). */
R2IQ_API void* r2iq_cuda_host_alloc(size_t bytes);
/* The same, placed on the NUMA node of `device`'s PCIe root port (device < 0: the current device), so that the
 * copy engines of each GPU of a multi-socket box read and write socket-local memory -- the ingest side of the path
 * (Core/arch/linux/FX3handler.cpp:93-147 copies every USB frame into such a slot).  Falls back to the default
 * placement when the node is unknown; r2iq_cuda_host_alloc(bytes) == r2iq_cuda_host_alloc_on(bytes, -1). */
R2IQ_API void* r2iq_cuda_host_alloc_on(size_t bytes, int device);
/* NUMA node of a device's PCIe root port, -1 when unknown or switched off (R2IQ_NUMA=0). */
R2IQ_API int r2iq_cuda_device_numa_node(int device);
R2IQ_API int r2iq_cuda_host_free(void* p);

/* Output shared between the ranks of one box (one process per GPU; SURVEY.md 8e: "gather IQ to one rank").  The
 * collecting rank allocates the destination with device_alloc and exports it; every other rank opens the handle
 * (CUDA IPC, peer access over NVLink) and passes `mapped + its offset` as d_iq_out of r2iq_cuda_process_device /
 * r2iq_cuda_process_channels_device: the fused kernel stores its IQ into the collecting rank's memory while it
 * computes, in place of a gather after the kernel.  The caller synchronises its stream and then the ranks (any
 * barrier) before the collecting rank reads.  The handle is an opaque 64-byte token to be sent between the processes. */
#define R2IQ_IPC_HANDLE_BYTES 64
R2IQ_API void* r2iq_cuda_device_alloc(int device, size_t bytes);
R2IQ_API int r2iq_cuda_device_free(void* p);
R2IQ_API int r2iq_cuda_ipc_export(void* device_ptr, unsigned char handle[R2IQ_IPC_HANDLE_BYTES]);
R2IQ_API int r2iq_cuda_ipc_open(int device, const unsigned char handle[R2IQ_IPC_HANDLE_BYTES], void** device_ptr);
R2IQ_API int r2iq_cuda_ipc_close(void* device_ptr);

/* convert_float<rand>  (Core/fft_mt_r2iq.h:36-51) on its own: n int16 host samples (n even) ->
 * n floats, through the same device code the fused kernel uses.  Parity hook for the stage that
 * must be bit-exact. */
R2IQ_API int r2iq_cuda_convert(r2iq_cuda_t h, const int16_t* in, long n, int rand, float* out);

R2IQ_API int r2iq_cuda_get_stats(r2iq_cuda_t h, r2iq_cuda_stats* out);
/* FFTN_R_ADC this operator was created with (2048 or 8192). */
R2IQ_API int r2iq_cuda_get_fftn(r2iq_cuda_t h);
/* Name of the device kernel a launch at decimation index d uses on this handle, as profilers print it (measurement
 * records name the kernel that ran instead of assuming one). */
R2IQ_API int r2iq_cuda_kernel_name(r2iq_cuda_t h, int d, char* buf, int n);

/* Device-free host logic (usable without a GPU; exercised by the CPU test suite).
 *  design_filter:  filterHw[d] exactly as r2iq_cuda_create designs it (KaiserWindow + tap placement +
 *                  gain + spectrum; Core/fir.cpp:48-105, Core/fft_mt_r2iq.cpp:181-205); hw = halfFft complex.
 *  kaiser:         the Kaiser-windowed sinc taps alone (Core/fir.cpp:48 KaiserWindow, num_taps > 0 form).
 *  freq_to_tunebin: the quantisation of fft_mt_r2iq::setFreqOffset (Core/fft_mt_r2iq.cpp:101-109);
 *                  returns the residual and stores the UNCLAMPED tune bin. */
R2IQ_API int r2iq_cuda_design_filter(float gain, int d, float* hw_interleaved, int n_complex);
R2IQ_API int r2iq_cuda_kaiser(int num_taps, float astop_db, float f_pass, float f_stop, float* taps);
R2IQ_API float r2iq_cuda_freq_to_tunebin(float offset, int d, int* tunebin);

/* ---- Stand-alone batched transforms: the CUDA side of the reference's FFT backend abstraction --------------------
 * Core/fft_backend.h:22-50 (FFTBackend: plan_r2c / plan_c2c / execute_r2c / execute_c2c / destroy_plan / alloc / free)
 * at the sizes Core/fft_benchmark.cpp:12-15 sweeps.  FFTW conventions: unnormalised, forward = exp(-2 pi i nk/n),
 * r2c of n reals -> n/2 + 1 bins.  One plan runs `batch` transforms per launch, laid out back to back (input pitch n,
 * output pitch n/2 + 1 for r2c, n for c2c).  Lengths: r2c 128...8192, c2c 64...4096, powers of two.
 *   fft_plan            plan_r2c / plan_c2c: twiddles (float64, rounded once), kernel choice, a private stream.
 *   fft_execute         execute_r2c / execute_c2c on HOST buffers, blocking like the reference's backends: page-locked
 *                       buffers (r2iq_cuda_host_alloc, the backend's alloc()) of up to 256 KiB are read and written in
 *                       place by the kernel, anything else is staged through device memory.
 *   fft_execute_device  the same kernel on DEVICE buffers, asynchronous on `stream`, `batch` <= any size.
 * The register DFTs are the ones of the fused r2iq kernels (csrc/dft_regs.cuh); there is no CPU path. */
enum { R2IQ_FFT_R2C = 0, R2IQ_FFT_C2C_FORWARD = 1, R2IQ_FFT_C2C_BACKWARD = 2 };
typedef struct r2iq_cuda_fft_s* r2iq_cuda_fft_t;
R2IQ_API int r2iq_cuda_fft_plan(int device, int kind, int n, long batch, r2iq_cuda_fft_t* out);
R2IQ_API int r2iq_cuda_fft_execute(r2iq_cuda_fft_t plan, const void* in_host, void* out_host);
R2IQ_API int r2iq_cuda_fft_execute_device(r2iq_cuda_fft_t plan, const void* d_in, void* d_out, long batch, void* stream);
R2IQ_API long r2iq_cuda_fft_batch(r2iq_cuda_fft_t plan);
R2IQ_API int r2iq_cuda_fft_destroy(r2iq_cuda_fft_t plan);

#ifdef __cplusplus
}
#endif
#endif /* R2IQ_CUDA_H */
