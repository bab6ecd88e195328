"""oracle/r2iq_oracle.py -- TEST INFRASTRUCTURE, not product code.

CPU restatement (NumPy, float64 transforms) of the reference's real->IQ down-converter, the
``fft_mt_r2iq`` path of ShadowLive/ExtIO_sddc.  It exists only to CHECK the CUDA path: it may be
imported by ``tests/``, by ``__graft_entry__.smoke()`` and by ``bench.py``'s cpu_baseline leg, never
by the product package (``extio_sddc_b200``), which has no CPU fallback.

Parity status: PINNED.  The reference has no golden vectors for this path,
so the restatement is pinned against the reference's own code run here: ``oracle/_ref`` is the
unmodified ``Core/fft_mt_r2iq*.cpp`` + ``Core/fir.cpp`` compiled in place (oracle/Makefile) and
``tests/test_oracle.py`` checks this file against it and against the fixtures in ``tests/golden/``
generated from it (``tests/golden/make_golden.py``).

Each function cites the reference lines it restates (paths relative to /root/reference).
The transforms are evaluated in float64 (numpy.fft), i.e. this is the exact-arithmetic answer the
reference's float32 FFTW pipeline approximates to about 1e-7 relative; filter taps are designed in
float32 exactly as ``Core/fir.cpp`` does because those roundings are part of the filter's definition.
"""
from __future__ import annotations

import math

import numpy as np

# Core/config.h:80-81, Core/r2iq.h:5, Core/config.h:62
TRANSFER_SAMPLES = 65536
NDECIDX = 7
EXT_BLOCKLEN = 32768

# Core/config.h:52-55
BBRF103_GAINFACTOR = 7.8e-8
HF103_GAINFACTOR = 1.14e-8
RX888_GAINFACTOR = 0.695e-8
RX888MK2_GAINFACTOR = 1.08e-8

_f32 = np.float32


def half_fft(fftn: int) -> int:
    """Core/fft_mt_r2iq.h:18."""
    return fftn // 2


def fft_per_buf(fftn: int) -> int:
    """Core/fft_mt_r2iq.h:19 -- frames per 65536-sample block (43 at FFTN 2048, 11 at 8192)."""
    return TRANSFER_SAMPLES // (3 * half_fft(fftn) // 2) + 1


def mfft_of(fftn: int, d: int) -> int:
    """Core/fft_mt_r2iq.cpp:43-48 -- inverse FFT size per decimation index."""
    return half_fft(fftn) >> d


def out_per_block(d: int) -> int:
    """Complex samples produced per input block: mfft/2 + (fftPerBuf-1)*3*mfft/4 = 32768 >> d
    (Core/fft_mt_r2iq_impl.hpp:121,128-137)."""
    return (TRANSFER_SAMPLES // 2) >> d


# ---------------------------------------------------------------------------------------------
# Core/fir.cpp
# ---------------------------------------------------------------------------------------------
def _izero(x: np.float32) -> np.float32:
    """Core/fir.cpp:7-25 -- modified Bessel I0 by power series, float32 arithmetic."""
    x2 = _f32(x) / _f32(2.0)
    s = _f32(1.0)
    ds = _f32(1.0)
    di = _f32(1.0)
    errorlimit = _f32(1e-9)
    while True:
        tmp = x2 / di
        tmp = tmp * tmp
        ds = ds * tmp
        s = s + ds
        di = _f32(float(di) + 1.0)
        if not (ds >= errorlimit * s):
            break
    return s


def kaiser_window(num_taps: int, astop: float, norm_fpass: float, norm_fstop: float) -> np.ndarray:
    """Core/fir.cpp:48-105 -- Kaiser-windowed sinc low-pass, float32 throughout (num_taps > 0 form)."""
    K_PI = _f32(3.141592653)
    K_2PI = _f32(2) * K_PI
    astop = _f32(astop)
    fpass = _f32(norm_fpass)
    fstop = _f32(norm_fstop)
    fcut = (fstop + fpass) / _f32(2.0)
    if astop < _f32(20.96):
        beta = _f32(0.0)
    elif astop >= _f32(50.0):
        beta = _f32(0.1102) * (astop - _f32(8.71))
    else:
        beta = _f32(0.5842) * _f32(math.pow(float(astop - _f32(20.96)), 0.4)) + _f32(0.07886) * (astop - _f32(20.96))
    n_taps = int(num_taps)
    fcenter = _f32(0.5) * _f32(n_taps - 1)
    izb = _izero(beta)
    coef = np.zeros(n_taps, dtype=np.float32)
    half = (_f32(n_taps) - _f32(1.0)) / _f32(2.0)
    for n in range(n_taps):
        x = _f32(n) - fcenter
        if _f32(n) == fcenter:
            c = _f32(2.0) * fcut
        else:
            arg = K_2PI * x * fcut
            c = _f32(math.sin(float(arg))) / (K_PI * x)
        x = (_f32(n) - half) / half
        w = _izero(beta * _f32(math.sqrt(float(_f32(1) - x * x)))) / izb
        coef[n] = _f32(1.0) * c * w
    return coef


def design_filter(gain: float, d: int, fftn: int = 2048) -> np.ndarray:
    """Core/fft_mt_r2iq.cpp:181-205 -- filterHw[d]: halfFft complex64 bins.

    Taps: KaiserWindow(halfFft/4+1, 120 dB, 0.85*Bw/128, 1.1*Bw/128), Bw = 64 / 2^d (float32);
    gainadj = gain*2048/FFTN; taps are placed REVERSED at the END of a halfFft buffer
    (``pfilterht[halfFft-1-t] = gainadj*pht[t]``), then an unnormalised forward FFT of size halfFft.
    """
    hf = half_fft(fftn)
    bw = _f32(64.0) / _f32(1 << d)
    pht = kaiser_window(hf // 4 + 1, 120.0, _f32(0.85) * bw / _f32(128.0), _f32(1.1) * bw / _f32(128.0))
    gainadj = _f32(gain) * _f32(2048.0) / _f32(fftn)
    ht = np.zeros(hf, dtype=np.float32)
    for t in range(hf // 4 + 1):
        ht[hf - 1 - t] = gainadj * pht[t]
    return np.fft.fft(ht.astype(np.float64)).astype(np.complex64)


# ---------------------------------------------------------------------------------------------
# Core/fft_mt_r2iq.cpp:101-109
# ---------------------------------------------------------------------------------------------
def set_freq_offset(offset: float, d: int, fftn: int = 2048):
    """Returns (mtunebin, residual) like fft_mt_r2iq::setFreqOffset: the tune bin is quantised to a
    multiple of 4 with truncation toward zero (float32 product), the residual is
    (mtunebin/halfFft - offset) * 2^d for the fine-tune mixer."""
    hf = half_fft(fftn)
    off = _f32(offset)
    tunebin = int(off * _f32(hf) / _f32(4)) * 4
    delta = _f32(_f32(tunebin) / _f32(hf)) - off
    return tunebin, float(delta * _f32(1 << d))


# ---------------------------------------------------------------------------------------------
# Core/fft_mt_r2iq.h:36-51
# ---------------------------------------------------------------------------------------------
def derandomize(x: np.ndarray, rand: bool) -> np.ndarray:
    """convert_float<rand> integer stage: if the LSB is set, XOR with -2 (0xFFFE)."""
    x = np.asarray(x, dtype=np.int16)
    if not rand:
        return x
    return np.where((x & 1) != 0, x ^ np.int16(-2), x).astype(np.int16)


def convert_float(x: np.ndarray, rand: bool) -> np.ndarray:
    """convert_float<rand>: de-randomise then int16 -> float32 (exact)."""
    return derandomize(x, rand).astype(np.float32)


def randomize(clean: np.ndarray) -> np.ndarray:
    """Inverse of the de-randomiser (it is an involution on the upper 15 bits, keyed by the LSB):
    what an RX-888 with the LTC2208 randomizer enabled would put on the wire for `clean`."""
    return derandomize(clean, True)


# ---------------------------------------------------------------------------------------------
# Core/fft_mt_r2iq_impl.hpp -- the worker loop
# ---------------------------------------------------------------------------------------------
def r2iq(blocks: np.ndarray, d: int, tunebin: int, *, gain: float = 1.0, rand: bool = False,
         sideband: bool = False, fftn: int = 2048, history: np.ndarray | None = None,
         filter_hw: np.ndarray | None = None, dtype=np.complex64) -> np.ndarray:
    """Restatement of the per-block loop, vectorised over the frames of a block.

    blocks   : int16, n_blocks*65536 samples (flat or [n_blocks, 65536])
    history  : the halfFft samples preceding block 0 (None = zeros).  The reference reads the tail
               of the previous ring slot (impl.hpp:32,42); for the first block that is whatever the
               ring held, here it is explicit.
    returns  : n_blocks * (32768 >> d) complex samples.
    """
    hf = half_fft(fftn)
    n_fft = fftn
    hop = 3 * hf // 2
    nfr = fft_per_buf(fftn)
    mfft = hf >> d
    blocks = np.asarray(blocks, dtype=np.int16).reshape(-1, TRANSFER_SAMPLES)
    n_blocks = blocks.shape[0]
    H = design_filter(gain, d, fftn) if filter_hw is None else np.asarray(filter_hw, dtype=np.complex64)
    H = H.astype(np.complex128)

    # impl.hpp:76-80 -- band-select parameters
    count = min(mfft // 2, hf - tunebin)
    start = max(0, mfft // 2 - tunebin)

    opb = out_per_block(d)
    out = np.zeros(n_blocks * opb, dtype=np.complex128)
    prev_tail = np.zeros(hf, dtype=np.int16) if history is None else np.asarray(history, dtype=np.int16)
    assert prev_tail.shape[0] == hf
    idx = (np.arange(nfr)[:, None] * hop + np.arange(n_fft)[None, :])
    for b in range(n_blocks):
        # impl.hpp:40-54 -- history + block, de-randomised, as float
        t = np.concatenate([convert_float(prev_tail, rand), convert_float(blocks[b], rand)]).astype(np.float64)
        prev_tail = blocks[b, TRANSFER_SAMPLES - hf:]
        frames = t[idx]                                  # impl.hpp:88: frame k starts at k*3*halfFft/2
        X = np.fft.rfft(frames, axis=1)                  # halfFft+1 bins, unnormalised
        Y = np.zeros((nfr, mfft), dtype=np.complex128)
        # impl.hpp:90-92 -- lower half: X[tunebin+m]*H[m], m in [0,count); rest of the half is zero
        Y[:, :count] = X[:, tunebin:tunebin + count] * H[None, :count]
        # impl.hpp:94-96 -- upper half: X[tunebin-mfft/2+m]*H[halfFft-mfft/2+m], m in [start,mfft/2)
        m = np.arange(start, mfft // 2)
        Y[:, mfft // 2 + m] = X[:, tunebin - mfft // 2 + m] * H[None, hf - mfft // 2 + m]
        y = np.fft.ifft(Y, axis=1) * mfft                # impl.hpp:98: unnormalised backward c2c
        if sideband:                                     # impl.hpp:124-130: copy<true> conjugates
            y = np.conj(y)
        o = out[b * opb:(b + 1) * opb]
        o[:mfft // 2] = y[0, mfft // 4: mfft // 4 + mfft // 2]            # impl.hpp:127|134 (k == 0)
        o[mfft // 2:] = y[1:, :3 * mfft // 4].reshape(-1)                  # impl.hpp:129|136 (k >= 1)
    return out.astype(dtype)


# ---------------------------------------------------------------------------------------------
# synthetic inputs of SURVEY.md section 8(d)
# ---------------------------------------------------------------------------------------------
def tone_blocks(n_blocks: int, amplitude: float = 16384.0, period: int = 64) -> np.ndarray:
    """unittest/benchmark_test.cpp:80-84 -- x[i] = (int16)(A*sin(2*pi*i/period)), phase-continuous."""
    i = np.arange(n_blocks * TRANSFER_SAMPLES, dtype=np.float64)
    return (amplitude * np.sin(2.0 * np.pi * i / period)).astype(np.int16)


def sine_blocks(n_blocks: int, freq_hz: float, fs_hz: float, amplitude: float = 10000.0) -> np.ndarray:
    """unittest/signal_integrity_test.cpp:131-141 -- phase-accumulated sine."""
    n = np.arange(n_blocks * TRANSFER_SAMPLES, dtype=np.float64)
    return (amplitude * np.sin(2.0 * np.pi * freq_hz / fs_hz * n)).astype(np.int16)


def rx888_blocks(n_blocks: int, d: int, tunebin: int, fftn: int = 2048, seed: int = 888,
                 randomized: bool = True) -> np.ndarray:
    """three tones (one inside the pass-band of decimation d at
    tune + 0.1*Fs/2^(d+2), two out of band) of amplitude 8000/4000/4000 plus uniform noise +-512,
    rounded to int16, then passed through the RX-888 randomizer when `randomized`."""
    rng = np.random.default_rng(seed)
    n = np.arange(n_blocks * TRANSFER_SAMPLES, dtype=np.float64)
    f_tune = tunebin / fftn                       # cycles/sample (bin k of an FFTN-point real FFT)
    f_in = f_tune + 0.1 / (1 << (d + 2))
    f_out1 = (f_tune + 0.23) % 0.5
    f_out2 = (f_tune + 0.37) % 0.5
    x = (8000.0 * np.sin(2 * np.pi * f_in * n) + 4000.0 * np.sin(2 * np.pi * f_out1 * n + 0.7)
         + 4000.0 * np.sin(2 * np.pi * f_out2 * n + 1.9) + rng.uniform(-512.0, 512.0, n.shape[0]))
    clean = np.rint(x).astype(np.int16)
    return randomize(clean) if randomized else clean


def rel_l2(a: np.ndarray, b: np.ndarray) -> float:
    a = np.asarray(a).astype(np.complex128)
    b = np.asarray(b).astype(np.complex128)
    den = float(np.sqrt(np.sum(np.abs(b) ** 2)))
    num = float(np.sqrt(np.sum(np.abs(a - b) ** 2)))
    return num / den if den > 0 else num


def full_scale(gain: float) -> float:
    """a full-scale int16 tone comes out with amplitude 32768*gain*1024."""
    return 32768.0 * gain * 1024.0
